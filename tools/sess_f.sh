bash tools/gpu_session.sh r2f tests bench rounds
LZ77SSS_FUSED_CANDIDATES=1 LZ77SSS_SYNC_KERNELS=1 timeout 300 python tools/fault_repro.py > gpurun_out/r2f_fault_sync.log 2>&1; echo "fused + sync kernels:"; grep -E "candidates:|failed|fault" gpurun_out/r2f_fault_sync.log | tail -4 | cut -c1-300
for tau in 8 16 32 64 128 256 1024 2048 4096; do
  timeout 300 python bench.py --steps 2 --warmup 1 --tau $tau --no-cpu-baseline > gpurun_out/r2f_tau_$tau.json 2> gpurun_out/r2f_tau_$tau.err || tail -2 gpurun_out/r2f_tau_$tau.err
  python tools/show_bench.py gpurun_out/r2f_tau_$tau.json short
done
timeout 300 python bench.py --steps 2 --warmup 1 --tau 4 --size-mib 256 --no-cpu-baseline > gpurun_out/r2f_tau_4_256mib.json 2> gpurun_out/r2f_tau_4.err || tail -2 gpurun_out/r2f_tau_4.err
python tools/show_bench.py gpurun_out/r2f_tau_4_256mib.json short
bash tools/gpu_session.sh r2f big16
