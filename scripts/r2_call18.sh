#!/bin/bash
# usage: scripts/r2_call18.sh N   - the driver's SCALE launch of bench.py at N GPUs
N=$1
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout -s KILL 900 $T --master-port 2970$N bench.py --gpus $N --steps 8 --warmup 3 > gpurun_out/r2q_bench_${N}gpu.json 2> gpurun_out/r2q_bench_${N}gpu.err; grep '^{' gpurun_out/r2q_bench_${N}gpu.json | cut -c1-200; tail -3 gpurun_out/r2q_bench_${N}gpu.err
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2q_bench_${N}gpu.json") if l.startswith("{")][-1])
s=d["sharded"]
print(d["value"], d["ms_per_step"], {k:s[k] for k in ("n_qubits","ms_per_happly","parity_max_rel_err","layouts")}, s["vs_unsharded"], s["strong"], s["nvlink"]["achieved_GBps_per_direction"])
PY
