#!/bin/bash
# 8 GPUs, N = 33: size of the centre group (3 = rank bits only: 8 KB rows per peer, one more local layout whose passes
# can hide under the centre pass now that the strided passes are pipelined)
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
for G in 3 4; do
  QSB200_CENTRE_GROUP=$G timeout -s KILL 400 $T --master-port 2990$G scripts/time_shard.py 30 > gpurun_out/r2x_w8_cg$G.jsonl 2> gpurun_out/r2x_w8_cg$G.err
  echo "[centre group $G]"; grep '^{' gpurun_out/r2x_w8_cg$G.jsonl | cut -c1-420; grep -o '"parity_max_rel_err": [0-9.e-]*\|"hbm_passes_ms_per_happly": [0-9.]*\|"pass_centre_per_happly": [0-9.]*' gpurun_out/r2x_w8_cg$G.jsonl | tr '\n' ' '; echo
done
