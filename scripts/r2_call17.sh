#!/bin/bash
mkdir -p gpurun_out
for PF in 0 1 2 3; do
  QSB200_LAST_PF=$PF timeout -s KILL 300 python bench.py --steps 8 --warmup 3 --no-cpu --no-sharded --no-grid > gpurun_out/r2p_bench_pf$PF.json 2> gpurun_out/r2p_bench_pf$PF.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r2p_bench_pf$PF.json"))
k=d["roofline"]["kernels"]
print("pf=$PF", round(d["ms_per_step"],2), {n: round(v["avg_ms"],3) for n,v in k.items()})
PY
done
for PF in 0 2; do QSB200_LAST_PF=$PF timeout -s KILL 300 python scripts/time_happly.py --n 28 --attempts 2 --profile | cut -c1-60,230-600; done
