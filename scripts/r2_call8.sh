#!/bin/bash
# round 2, GPU call 8 (2 GPUs): chunk-pipelined overlapped sharded mode - parity, then timing for K = 1, 2, 4, 8 groups
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout -s KILL 300 $T --master-port 29612 scripts/shard_parity.py --qubits 24 --method dp5 > gpurun_out/r2h_parity24.log 2>&1; grep '"ok"' gpurun_out/r2h_parity24.log || tail -5 gpurun_out/r2h_parity24.log
QSB200_SHARD_PIPE=3 timeout -s KILL 300 $T --master-port 29613 scripts/shard_parity.py --qubits 21 > gpurun_out/r2h_parity21_p3.log 2>&1; grep '"ok"' gpurun_out/r2h_parity21_p3.log || tail -5 gpurun_out/r2h_parity21_p3.log
for P in 2 0 1 3; do
  QSB200_SHARD_PIPE=$P timeout -s KILL 600 $T --master-port 2962$P scripts/time_shard.py 30 > gpurun_out/r2h_time_pipe$P.jsonl 2> gpurun_out/r2h_time_pipe$P.err
  echo "pipe_lb=$P"; cut -c1-420 gpurun_out/r2h_time_pipe$P.jsonl; grep -o '"parity_max_rel_err": [0-9.e-]*\|"max_abs_diff": [0-9.e-]*\|"strong": {[^}]*}' gpurun_out/r2h_time_pipe$P.jsonl | tr '\n' ' '; echo
done
