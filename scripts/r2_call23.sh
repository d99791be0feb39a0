#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2u_pytest_gpu.log 2>&1; tail -4 gpurun_out/r2u_pytest_gpu.log
for F in 1 0; do
  QSB200_DP5_FUSE=$F timeout -s KILL 300 python bench.py --steps 8 --warmup 3 --no-cpu --no-sharded --no-grid > gpurun_out/r2u_bench_fuse$F.json 2> gpurun_out/r2u_bench_fuse$F.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r2u_bench_fuse$F.json"))
k=d["roofline"]["kernels"]
print("fuse=$F", round(d["value"],3), round(d["ms_per_step"],2), d["gpu_launches"], round(d["raw_norm"],9), {n: (round(v["avg_ms"],3), v["launches_per_step"]) for n,v in k.items()})
PY
done
