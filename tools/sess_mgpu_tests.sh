# usage: bash tools/sess_mgpu_tests.sh <N> <tag>: the sharded GPU tests (virtual ranks on one GPU), then the N-GPU bench
export LZ77SSS_REQUIRE_REF=1
timeout 900 python -m pytest tests -m gpu -q -k "sharded or shard" > gpurun_out/$2_gputests_sharded.log 2>&1; tail -3 gpurun_out/$2_gputests_sharded.log
bash tools/sess_mgpu.sh $1 $2
