#!/bin/bash
# usage: scripts/run_bench_multi.sh N [extra bench args]   (torchrun launch as the driver does it)
N=$1; shift
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@"
