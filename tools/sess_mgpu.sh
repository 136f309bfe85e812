# usage: bash tools/sess_mgpu.sh <N> <tag>
N=$1; TAG=$2
export LZ77SSS_BENCH_VERBOSE=1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 2 > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err
tail -c 1500 gpurun_out/${TAG}_bench_${N}gpu.json; grep -E "rank 0|rank 1" gpurun_out/${TAG}_bench_${N}gpu.err | tail -4 | cut -c1-400
