bash tools/gpu_session.sh r2d tests bench rounds
LZ77SSS_FUSED_CANDIDATES=2 timeout 300 python tools/fault_repro.py > gpurun_out/r2d_fault_checked.log 2>&1; grep -E "check|fault|mode" gpurun_out/r2d_fault_checked.log | head -20
