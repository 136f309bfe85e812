bash tools/gpu_session.sh r2g tests bench
for g in 32 16 8; do echo "== PARSE_GROUP $g"; LZ77SSS_PARSE_GROUP=$g timeout 300 python tools/run_stats.py 1024 dna 16 profile 2>/dev/null | grep -E "wall_ms|parse_kernel|dirty" | cut -c1-100 | tail -3; done > gpurun_out/r2g_parse_group.log 2>&1; cat gpurun_out/r2g_parse_group.log
for v in 4 5 6 7; do LZ77SSS_FUSED_CANDIDATES=$v timeout 120 python tools/fault_repro.py > gpurun_out/r2g_fault_v$v.log 2>&1; echo "variant $v:"; grep -E "no fault|illegal|failed" gpurun_out/r2g_fault_v$v.log | head -2 | cut -c1-200; done
for tau in 4 8; do
  timeout 600 python bench.py --steps 2 --warmup 1 --tau $tau --size-mib 256 --no-cpu-baseline > gpurun_out/r2g_tau_${tau}_256mib.json 2> gpurun_out/r2g_tau_$tau.err || tail -2 gpurun_out/r2g_tau_$tau.err
  python tools/show_bench.py gpurun_out/r2g_tau_${tau}_256mib.json short
done
timeout 900 python tools/big_run.py --size-mib 1024 --workload dna --exact --steps 1 > gpurun_out/r2g_exact_1g_dna.json 2> gpurun_out/r2g_exact_1g.err || tail -3 gpurun_out/r2g_exact_1g.err
cut -c1-900 gpurun_out/r2g_exact_1g_dna.json
