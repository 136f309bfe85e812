# session z: radix scatter variants (items per thread u32 / u64, minimum CTAs per SM)
cp lz77_sss_b200/liblz77sss_b200.so /tmp/main.so
for v in main rs_a rs_b rs_c rs_d rs_e; do
  [ $v = main ] && cp /tmp/main.so lz77_sss_b200/liblz77sss_b200.so || cp lz77_sss_b200/_variants/lib_$v.so lz77_sss_b200/liblz77sss_b200.so
  echo "== $v"
  timeout 300 python tools/run_stats.py 1024 dna 16 profile 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('wall', d['wall_ms'], 'sort', d['t_sort_ns'], 'gap', d['t_gap_hash_ns'], 'total', d['t_total_ns'])
    elif 'rs_scatter' in l: print(l.rstrip())
"
done > gpurun_out/r2z_scatter_variants.log 2>&1
cp /tmp/main.so lz77_sss_b200/liblz77sss_b200.so
cat gpurun_out/r2z_scatter_variants.log
