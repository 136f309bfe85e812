# session n: block length / tracking threshold with memoized chain links
for bl in 512 1024 2048; do
  for td in 16 64; do
    echo "== BLOCK_LEN $bl TRACK_DIV $td"
    LZ77SSS_BLOCK_LEN=$bl LZ77SSS_TRACK_DIV=$td timeout 600 python tools/run_stats.py 1024 dna 16 profile 2>&1 | grep -E "wall_ms|parse_kernel|dirty_from" | cut -c1-90 | tail -3
  done
done > gpurun_out/r2n_block_len.log 2>&1
cat gpurun_out/r2n_block_len.log
