# session y: timeline of a 2-rank sharded run (virtual ranks), ncu --set full of the gap index scatter
LZ77SSS_DEBUG_ROUNDS=1 timeout 600 python tools/sharded_check.py 1024 2 16 > gpurun_out/r2y_sharded_timeline.log 2>&1
tail -4 gpurun_out/r2y_sharded_timeline.log | cut -c1-200
O=gpurun_out; TAG=r2y
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:rs_scatter_kernel<unsigned int" -c 3 \
  -o $O/${TAG}_full_scatter32 -f python tools/run_stats.py 256 dna 16 > $O/${TAG}_ncu_full_scatter32.log 2>&1
ncu -i $O/${TAG}_full_scatter32.ncu-rep --page raw --csv > $O/${TAG}_full_scatter32.csv 2>/dev/null
rm -f $O/${TAG}_full_scatter32.ncu-rep; ls -la $O/${TAG}_full_scatter32.csv
