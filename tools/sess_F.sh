# session F: tests, bench, exact factorization of 3 GiB, tau sweep with the final code
bash tools/gpu_session.sh r2F tests bench
timeout 900 python tools/big_run.py --size-mib 3072 --workload genomes --exact --steps 1 > gpurun_out/r2F_exact_3072mib.json 2> gpurun_out/r2F_exact_3072mib.err || tail -5 gpurun_out/r2F_exact_3072mib.err
cut -c1-1000 gpurun_out/r2F_exact_3072mib.json
bash tools/gpu_session.sh r2F sweep
