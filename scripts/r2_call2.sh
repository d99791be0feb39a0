#!/bin/bash
# round 2, GPU call 2 (2 GPUs): multi-process sharded parity, overlapped vs serialised timing at N = 31 and smaller
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout -s KILL 300 $T --master-port 29611 scripts/shard_parity.py --qubits 20 > gpurun_out/r2b_parity20.log 2>&1; tail -2 gpurun_out/r2b_parity20.log
timeout -s KILL 300 $T --master-port 29612 scripts/shard_parity.py --qubits 24 --method dp5 > gpurun_out/r2b_parity24.log 2>&1; tail -2 gpurun_out/r2b_parity24.log
QSB200_SHARD_OVERLAP=0 timeout -s KILL 300 $T --master-port 29613 scripts/shard_parity.py --qubits 22 > gpurun_out/r2b_parity22_serial.log 2>&1; tail -2 gpurun_out/r2b_parity22_serial.log
timeout -s KILL 600 $T --master-port 29614 scripts/time_shard.py 27 30 > gpurun_out/r2b_time_overlap.jsonl 2> gpurun_out/r2b_time_overlap.err; cat gpurun_out/r2b_time_overlap.jsonl; tail -3 gpurun_out/r2b_time_overlap.err
QSB200_SHARD_OVERLAP=0 timeout -s KILL 600 $T --master-port 29615 scripts/time_shard.py 30 > gpurun_out/r2b_time_serial.jsonl 2> gpurun_out/r2b_time_serial.err; cat gpurun_out/r2b_time_serial.jsonl; tail -3 gpurun_out/r2b_time_serial.err
timeout -s KILL 600 python -m pytest tests/test_gpu_shard.py tests/test_gpu_evolve.py -x -q > gpurun_out/r2b_pytest.log 2>&1; tail -3 gpurun_out/r2b_pytest.log
