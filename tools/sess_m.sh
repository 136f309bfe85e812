# session m: chain links looked up on demand and kept
bash tools/gpu_session.sh r2m tests bench rounds
for pf in 1 0; do
  echo "== PREFETCH_INDEX $pf"
  LZ77SSS_PREFETCH_INDEX=$pf timeout 600 python tools/run_stats.py 1024 dna 16 2>&1 | cut -c1-1200 | tail -1
done > gpurun_out/r2m_prefetch_ab.log 2>&1
cat gpurun_out/r2m_prefetch_ab.log | cut -c1-200
