# session t: decode with one or two hops per round
for h in 1 2; do echo "== DECODE_HOPS $h"; LZ77SSS_DECODE_HOPS=$h timeout 300 python tools/run_stats.py 1024 dna 16 decode 2>&1 | tail -1; done > gpurun_out/r2t_decode_hops.log 2>&1
cat gpurun_out/r2t_decode_hops.log
