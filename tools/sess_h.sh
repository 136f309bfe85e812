for v in 2 8; do LZ77SSS_FUSED_CANDIDATES=$v timeout 120 python tools/fault_repro.py > gpurun_out/r2h_fault_v$v.log 2>&1; echo "variant $v:"; grep -E "no fault|illegal|check" gpurun_out/r2h_fault_v$v.log | head -3 | cut -c1-260; done
for bl in 1024 512; do echo "== BLOCK_LEN $bl"; LZ77SSS_BLOCK_LEN=$bl timeout 300 python tools/run_stats.py 1024 dna 16 profile 2>/dev/null | grep -E "wall_ms|parse_kernel|dirty" | cut -c1-100 | tail -3; done > gpurun_out/r2h_block_len.log 2>&1; cat gpurun_out/r2h_block_len.log
bash tools/gpu_session.sh r2h tests bench ncu_list
