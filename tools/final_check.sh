export LZ77SSS_REQUIRE_REF=1
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/r2Z_gputests.log 2>&1; tail -4 gpurun_out/r2Z_gputests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2Z_smoke.log 2>&1; tail -2 gpurun_out/r2Z_smoke.log
timeout 600 python bench.py > gpurun_out/r2Z_bench.json 2> gpurun_out/r2Z_bench.err || tail -5 gpurun_out/r2Z_bench.err
python tools/show_bench.py gpurun_out/r2Z_bench.json | cut -c1-300 | head -4
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r2Z_bench_reference.json 2> gpurun_out/r2Z_bench_reference.err || tail -5 gpurun_out/r2Z_bench_reference.err
cut -c1-400 gpurun_out/r2Z_bench_reference.json
