#!/bin/bash
# usage: tools/run_multi_gpu.sh <ngpus> [bench args...]   (under gpurun --gpus N)
N=$1; shift
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"
