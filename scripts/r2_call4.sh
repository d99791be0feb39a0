#!/bin/bash
# round 2, GPU call 4 (2 GPUs): overlapped sharded mode fixed; DP5 error fold A/B on one GPU
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout -s KILL 300 $T --master-port 29612 scripts/shard_parity.py --qubits 24 --method dp5 > gpurun_out/r2d_parity24.log 2>&1; grep '"ok"' gpurun_out/r2d_parity24.log
timeout -s KILL 300 $T --master-port 29613 scripts/shard_parity.py --qubits 26 > gpurun_out/r2d_parity26.log 2>&1; grep '"ok"' gpurun_out/r2d_parity26.log
timeout -s KILL 600 $T --master-port 29614 scripts/time_shard.py 27 30 > gpurun_out/r2d_time_overlap.jsonl 2> gpurun_out/r2d_time_overlap.err; cat gpurun_out/r2d_time_overlap.jsonl; tail -3 gpurun_out/r2d_time_overlap.err
timeout -s KILL 600 python -m pytest tests/test_gpu_evolve.py tests/test_gpu_large.py -x -q > gpurun_out/r2d_pytest.log 2>&1; tail -3 gpurun_out/r2d_pytest.log
timeout -s KILL 300 python bench.py --steps 8 --warmup 3 --no-cpu --no-sharded --no-grid > gpurun_out/r2d_bench_fold.json 2> gpurun_out/r2d_bench_fold.err; cut -c1-200 gpurun_out/r2d_bench_fold.json
QSB200_DP5_ERR_SWEEP=1 timeout -s KILL 300 python bench.py --steps 8 --warmup 3 --no-cpu --no-sharded --no-grid > gpurun_out/r2d_bench_sweep.json 2> gpurun_out/r2d_bench_sweep.err; cut -c1-200 gpurun_out/r2d_bench_sweep.json
