"""Wall / device time of spin-3/2 trajectories at the reference's full trajectory size (30 s, 20000 outputs, the
sweep's tolerances; SW:1201-1252 with is_spin_three_half=True).  Writes gpurun_out/<tag>_time_spin32.json.

    python scripts/time_spin32.py [tag]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import dataclasses  # noqa: E402

import numpy as np  # noqa: E402

import quantumsimulations_b200 as q  # noqa: E402
from bench import sweep_params  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
out = {"what": "is_spin_three_half=True, full-size trajectories (t_final 30 s, 20000 outputs, atol 1e-10, rtol 1e-9)"}
q.simulate_rare(dataclasses.replace(sweep_params(3, 50e3, t_final=1e-3, steps=300), is_spin_three_half=True))   # warm-up
for n_sea in (6, 8):
    p = dataclasses.replace(sweep_params(n_sea, 50e3, t_final=30.0, steps=20000), is_spin_three_half=True)
    t0 = time.perf_counter()
    _, obs, st = q.simulate_rare(p, return_stats=True, raw_norm=True)
    wall = time.perf_counter() - t0
    out[f"n_sea_{n_sea}"] = {"index_bits": n_sea + 2, "wall_s": wall, "device_ms": st.device_ms, "column_h_applies": st.n_rhs,
                             "raw_norm_drift": float(np.max(np.abs(obs["raw_norm"] - 1.0))), "path": st.path}
    print(n_sea, out[f"n_sea_{n_sea}"], flush=True)
plist = []
for d in np.linspace(0.0, 150e3, 13):
    for variant in ("center_off", "center_on", "shell_off"):
        plist.append(dataclasses.replace(sweep_params(6, float(d), variant=variant, t_final=30.0, steps=20000),
                                         is_spin_three_half=True))
t0 = time.perf_counter()
res, stats = q.simulate_rare_batch(plist)
out["sweep_39_full_size"] = {"wall_s": time.perf_counter() - t0, "paths": sorted(set(s.path for s in stats)),
                             "device_ms_sum": float(sum(s.device_ms for s in stats if s.path == 4))}
print(out["sweep_39_full_size"], flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", f"{tag}_time_spin32.json"), "w") as f:
    json.dump(out, f, indent=1)
