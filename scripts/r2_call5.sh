#!/bin/bash
# round 2, GPU call 5 (2 GPUs): persistent centre pass in the overlapped sharded mode; centre OCC=3 A/B on one GPU
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout -s KILL 300 $T --master-port 29612 scripts/shard_parity.py --qubits 24 --method dp5 > gpurun_out/r2e_parity24.log 2>&1; grep '"ok"' gpurun_out/r2e_parity24.log
timeout -s KILL 600 $T --master-port 29614 scripts/time_shard.py 30 > gpurun_out/r2e_time_overlap.jsonl 2> gpurun_out/r2e_time_overlap.err; cat gpurun_out/r2e_time_overlap.jsonl; tail -3 gpurun_out/r2e_time_overlap.err
for c in 74 296; do
  QSB200_CENTRE_CTAS=$c timeout -s KILL 600 $T --master-port 29615 scripts/time_shard.py 30 > gpurun_out/r2e_time_overlap_c$c.jsonl 2> gpurun_out/r2e_time_overlap_c$c.err; cut -c1-120 gpurun_out/r2e_time_overlap_c$c.jsonl
done
for occ in 2 3; do
  QSB200_CENTRE_OCC=$occ timeout -s KILL 300 python scripts/time_happly.py --n 28 --attempts 2 --profile > gpurun_out/r2e_centre_occ$occ.json 2>&1; cat gpurun_out/r2e_centre_occ$occ.json
done
