#!/bin/bash
# 8 GPUs: sharded N = 33 with 16 chunk groups, and with the double-buffered centre kernel (HBM passes have slack here)
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
for V in "QSB200_SHARD_PIPE=4" "QSB200_CENTRE_PIPE=1" "QSB200_CENTRE_PIPE=1 QSB200_SHARD_PIPE=4"; do
  tag=$(echo $V | tr ' =' '__')
  env $V timeout -s KILL 400 $T --master-port 29811 scripts/time_shard.py 30 > gpurun_out/r2t_w8_$tag.jsonl 2> gpurun_out/r2t_w8_$tag.err
  echo "[$V]"; grep '^{' gpurun_out/r2t_w8_$tag.jsonl | cut -c1-330; grep -o '"parity_max_rel_err": [0-9.e-]*\|"strong": {[^}]*}' gpurun_out/r2t_w8_$tag.jsonl | tr '\n' ' '; echo
done
