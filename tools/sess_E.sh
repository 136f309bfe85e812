# session E: launch list of the final state, sharded check with 64-bit positions, exact factorization just below 4 GiB
bash tools/gpu_session.sh r2E ncu_list
timeout 900 python tools/sharded_check.py 4160 2 16 genomes > gpurun_out/r2E_sharded_check_4gib.log 2>&1; tail -6 gpurun_out/r2E_sharded_check_4gib.log | cut -c1-300
timeout 1200 python tools/big_run.py --size-mib 4000 --workload genomes --exact --steps 1 > gpurun_out/r2E_exact_4000mib.json 2> gpurun_out/r2E_exact_4000mib.err || tail -5 gpurun_out/r2E_exact_4000mib.err
cut -c1-1200 gpurun_out/r2E_exact_4000mib.json
