#!/bin/bash
# 8-GPU (or N-GPU) weak-scaling bench under torchrun, as the driver launches it.  usage: bash tools/gpu_n8.sh [N] [tag]
N=${1:-8}; TAG=${2:-n$N}
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 5 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"
wc -l $OUT/bench.json; cut -c1-300 $OUT/bench.json
