#!/bin/bash
# round 2, GPU call 16 (8 GPUs): the driver's SCALE launch of bench.py at N = 8, both arms
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout -s KILL 600 $T --master-port 29701 bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > gpurun_out/r2o_bench_ref_8gpu.json 2> gpurun_out/r2o_bench_ref_8gpu.err; cut -c1-300 gpurun_out/r2o_bench_ref_8gpu.json
timeout -s KILL 900 $T --master-port 29702 bench.py --gpus 8 --steps 8 --warmup 3 > gpurun_out/r2o_bench_8gpu.json 2> gpurun_out/r2o_bench_8gpu.err; cut -c1-300 gpurun_out/r2o_bench_8gpu.json; tail -3 gpurun_out/r2o_bench_8gpu.err
