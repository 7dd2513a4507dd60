#!/bin/bash
# usage: scripts/bench_n.sh N [extra bench args]   -- the driver's launch line for N GPUs of one node
N=$1; shift
if [ "$N" = "1" ]; then python bench.py --gpus 1 "$@"; else
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N "$@"; fi
