# session w: sharded stream against the single-GPU path at bench sizes (virtual ranks on one GPU)
for a in "256 2 16" "1024 2 16 equal" "1024 2 16" "1024 2 24"; do
  echo "== $a"; timeout 600 python tools/sharded_check.py $a 2>&1 | tail -14 | cut -c1-600
done > gpurun_out/r2w_sharded_check.log 2>&1
cat gpurun_out/r2w_sharded_check.log
