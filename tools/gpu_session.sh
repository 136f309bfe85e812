#!/bin/bash
# One GPU session: tests, bench (with per-round profile), fault reproducer under compute-sanitizer, large runs, ncu captures.
# usage (on the GPU box, from the repo root):  bash tools/gpu_session.sh <tag> [steps...]
#   steps: tests bench rounds fault big4 big16 ncu_list ncu_full sweep
TAG=$1; shift
STEPS="$*"; [ -z "$STEPS" ] && STEPS="tests bench"
O=gpurun_out
mkdir -p $O
export LZ77SSS_REQUIRE_REF=1
has() { [[ " $STEPS " == *" $1 "* ]]; }
if has tests; then
  (time timeout 1500 python -m pytest tests -m gpu -q --durations=10) > $O/${TAG}_gputests.log 2>&1
  tail -15 $O/${TAG}_gputests.log
fi
if has bench; then
  timeout 900 python bench.py --steps 5 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err || tail -5 $O/${TAG}_bench.err
  python tools/show_bench.py $O/${TAG}_bench.json
fi
if has rounds; then
  LZ77SSS_DEBUG_ROUNDS=1 timeout 600 python tools/run_stats.py 1024 dna 16 profile > $O/${TAG}_rounds.log 2>&1
  tail -60 $O/${TAG}_rounds.log
fi
if has fault; then
  LZ77SSS_FUSED_CANDIDATES=1 LZ77SSS_SYNC_KERNELS=1 timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python tools/fault_repro.py > $O/${TAG}_fault_memcheck.log 2>&1
  tail -40 $O/${TAG}_fault_memcheck.log
  LZ77SSS_FUSED_CANDIDATES=1 timeout 300 python tools/fault_repro.py > $O/${TAG}_fault_plain.log 2>&1; tail -5 $O/${TAG}_fault_plain.log
fi
if has big4; then
  timeout 1500 python tools/big_run.py --size-mib 4160 --workload genomes --phr lpf_opt --ref > $O/${TAG}_big4_genomes.json 2> $O/${TAG}_big4_genomes.err || tail -8 $O/${TAG}_big4_genomes.err
  cat $O/${TAG}_big4_genomes.json | cut -c1-1500
fi
if has big16; then
  timeout 2400 python tools/big_run.py --size-mib 16384 --workload markov --phr lpf_lnf_opt --steps 1 > $O/${TAG}_big16_markov.json 2> $O/${TAG}_big16_markov.err || tail -8 $O/${TAG}_big16_markov.err
  cat $O/${TAG}_big16_markov.json | cut -c1-1500
fi
if has sweep; then
  # the tau sweep of bench/lz77_sss_bench_tau.cpp (no CPU arm: the reference needs minutes per point at small tau);
  # tau = 4 and 8 at 256 MiB (the structures over |S| = n/2 .. n/4 positions do not fit next to a 1 GiB text's slices)
  for tau in 4 8 16 32 64 128 256 1024 2048 4096; do
    size=1024; [ $tau -le 8 ] && size=256
    timeout 300 python bench.py --steps 2 --warmup 1 --tau $tau --size-mib $size --no-cpu-baseline > $O/${TAG}_tau_$tau.json 2> $O/${TAG}_tau_$tau.err || tail -3 $O/${TAG}_tau_$tau.err
    python tools/show_bench.py $O/${TAG}_tau_$tau.json short
  done
fi
if has ncu_list; then
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/${TAG}_launches.csv python tools/run_stats.py 1024 dna 16 > $O/${TAG}_ncu_list.log 2>&1
  python tools/launch_summary.py $O/${TAG}_launches.csv > $O/${TAG}_launches_summary.csv 2>&1; head -30 $O/${TAG}_launches_summary.csv
  gzip -f $O/${TAG}_launches.csv
fi
if has ncu_full; then
  # one ncu --set full capture per hot kernel (3 launches each) of a 256 MiB run; the raw pages as csv, the reports dropped
  for spec in "parse:parse_kernel" "entries:gap_entries_sparse_kernel" "scatter32:rs_scatter_kernel<unsigned int, 512" "sss:sss_cell_kernel" \
              "dirty:dirty_from_bitmap_kernel" "decode:decode_round_kernel" "splice:splice_repair_kernel" "repair:repair_entries_kernel"; do
    tag=${spec%%:*}; k=${spec#*:}
    timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$k" -c 3 \
      -o $O/${TAG}_full_$tag -f python tools/run_stats.py 256 dna 16 decode > $O/${TAG}_ncu_full_$tag.log 2>&1
    ncu -i $O/${TAG}_full_$tag.ncu-rep --page raw --csv > $O/${TAG}_full_$tag.csv 2>/dev/null
    rm -f $O/${TAG}_full_$tag.ncu-rep
  done
  ls -la $O | tail -12
fi
