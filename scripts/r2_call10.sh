#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_gpu_evolve.py tests/test_gpu_large.py tests/test_gpu_happly.py -x -q > gpurun_out/r2j_pytest.log 2>&1; tail -4 gpurun_out/r2j_pytest.log
for n in 29 30; do timeout -s KILL 300 python scripts/time_happly.py --n $n --attempts 2 --profile > gpurun_out/r2j_happly_n$n.json 2>&1; cat gpurun_out/r2j_happly_n$n.json; done
python - <<'PY' > gpurun_out/r2j_c1_n13.json 2>&1
import json, os, sys, time
sys.path.insert(0, os.getcwd())
from quantumsimulations_b200 import DipolarRareParams, simulate_rare
p = DipolarRareParams(is_spin_three_half=False)
out = {}
for tag, env in (("dense", None), ("stepped", "0")):
    if env is None: os.environ.pop("QSB200_DENSE_OUTPUT", None)
    else: os.environ["QSB200_DENSE_OUTPUT"] = env
    simulate_rare(p)
    t0 = time.perf_counter(); _, _, st = simulate_rare(p, return_stats=True); w = time.perf_counter() - t0
    out[tag] = {"wall_s": w, "attempts": st.n_accepted + st.n_rejected, "h_applies": st.n_rhs}
print(json.dumps(out))
PY
cat gpurun_out/r2j_c1_n13.json
