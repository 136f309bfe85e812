# session x: the no-allocation entry point alone in a fresh process, the sharded check at 1 GiB, then the whole GPU suite
export LZ77SSS_REQUIRE_REF=1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2x_noalloc_alone.log 2>&1; tail -5 gpurun_out/r2x_noalloc_alone.log
for a in "1024 2 16" "1024 3 24" "1024 4 16"; do
  echo "== $a"; timeout 600 python tools/sharded_check.py $a 2>&1 | tail -9 | cut -c1-300
done > gpurun_out/r2x_sharded_check.log 2>&1
grep -E "^==|equal|first diff|sharded" gpurun_out/r2x_sharded_check.log
bash tools/gpu_session.sh r2x tests
