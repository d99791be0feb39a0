#!/bin/bash
# round 2, GPU call 3 (2 GPUs): overlapped sharded mode with one communicator per stream - parity, then timing
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout -s KILL 300 $T --master-port 29612 scripts/shard_parity.py --qubits 24 --method dp5 > gpurun_out/r2c_parity24.log 2>&1; grep '"ok"' gpurun_out/r2c_parity24.log
timeout -s KILL 300 $T --master-port 29613 scripts/shard_parity.py --qubits 26 > gpurun_out/r2c_parity26.log 2>&1; grep '"ok"' gpurun_out/r2c_parity26.log
timeout -s KILL 600 $T --master-port 29614 scripts/time_shard.py 27 30 > gpurun_out/r2c_time_overlap.jsonl 2> gpurun_out/r2c_time_overlap.err; cat gpurun_out/r2c_time_overlap.jsonl; tail -3 gpurun_out/r2c_time_overlap.err
