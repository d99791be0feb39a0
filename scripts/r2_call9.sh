#!/bin/bash
# round 2, GPU call 9 (8 GPUs): chunk-pipelined sharded mode at world 8 - parity, N = 33 timing, centre CTAs per SM A/B
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout -s KILL 300 $T --master-port 29612 scripts/shard_parity.py --qubits 25 --method dp5 > gpurun_out/r2i_parity25_w8.log 2>&1; grep '"ok"' gpurun_out/r2i_parity25_w8.log || tail -5 gpurun_out/r2i_parity25_w8.log
timeout -s KILL 600 $T --master-port 29613 scripts/time_shard.py 30 > gpurun_out/r2i_time_w8.jsonl 2> gpurun_out/r2i_time_w8.err; cut -c1-500 gpurun_out/r2i_time_w8.jsonl; grep -o '"parity_max_rel_err": [0-9.e-]*\|"max_abs_diff": [0-9.e-]*\|"strong": {[^}]*}' gpurun_out/r2i_time_w8.jsonl | tr '\n' ' '; echo; tail -2 gpurun_out/r2i_time_w8.err
QSB200_CENTRE_CTAS=296 timeout -s KILL 600 $T --master-port 29614 scripts/time_shard.py 30 > gpurun_out/r2i_time_w8_c296.jsonl 2> gpurun_out/r2i_time_w8_c296.err; cut -c1-500 gpurun_out/r2i_time_w8_c296.jsonl
QSB200_SHARD_OVERLAP=0 timeout -s KILL 600 $T --master-port 29615 scripts/time_shard.py 30 > gpurun_out/r2i_time_w8_serial.jsonl 2> gpurun_out/r2i_time_w8_serial.err; cut -c1-500 gpurun_out/r2i_time_w8_serial.jsonl
