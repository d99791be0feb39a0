#!/usr/bin/env python
"""Wall time of the C2 sweep (39 trajectories) and the C3 grid (741) at the CI-sized window, on-chip path.

    python scripts/time_sweep.py
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from bench import F1A, sweep_params  # noqa: E402
from quantumsimulations_b200 import simulate_rare_batch  # noqa: E402

V = ("center_off", "center_on", "shell_off")
c2 = [sweep_params(6, d, v) for d in np.linspace(0.0, 3.0 * F1A, 13) for v in V]
c3 = [sweep_params(6, d, v, f1a=f) for f in np.arange(5e3, 50e3 + 1, 2.5e3) for d in np.linspace(0.0, 3.0 * f, 13) for v in V]
simulate_rare_batch(c2[:3])  # context + module load
out = {}
for name, plist, kw in (("c2_taylor", c2, {"method": "taylor"}), ("c2_dp5", c2, {"method": "dp5"}), ("c3_taylor", c3, {"method": "taylor"})):
    t0 = time.perf_counter()
    res, st = simulate_rare_batch(plist, **kw)
    out[name + "_s"] = round(time.perf_counter() - t0, 4)
    out[name + "_kernel_ms"] = round(max(s.device_ms for s in st), 1)
    out[name + "_iz_last"] = float(res[-1][1]["Iz_sea"][-1])
print(json.dumps(out))
