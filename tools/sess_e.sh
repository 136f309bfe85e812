bash tools/gpu_session.sh r2e tests bench
for div in 4 16 64; do echo "== TRACK_DIV $div"; LZ77SSS_TRACK_DIV=$div LZ77SSS_DEBUG_ROUNDS=1 timeout 300 python tools/run_stats.py 1024 dna 16 profile 2>/dev/null | grep -E "wall_ms|parse|dirty" | cut -c1-160 | tail -14; done > gpurun_out/r2e_trackdiv.log 2>&1; cat gpurun_out/r2e_trackdiv.log
for v in 1 3; do LZ77SSS_FUSED_CANDIDATES=$v timeout 300 python tools/fault_repro.py > gpurun_out/r2e_fault_v$v.log 2>&1; echo "variant $v:"; tail -2 gpurun_out/r2e_fault_v$v.log | cut -c1-300; done
bash tools/gpu_session.sh r2e big4
bash tools/gpu_session.sh r2e big16
