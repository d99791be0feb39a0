#!/usr/bin/env python
"""Wall / device time of the HBM-resident propagator path (11 <= N <= 14) on full-size sweep trajectories
(t_final = 30 s, 20000 outputs, the sweep's tolerances), next to the stepped Taylor path on a short window of the same
trajectory (extrapolated to the full length).

    python scripts/time_propagator_hbm.py [n_sea ...]        # default 10 11 12
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import sweep_params  # noqa: E402
from quantumsimulations_b200 import api, compile_table, simulate_rare  # noqa: E402

for n_sea in [int(a) for a in sys.argv[1:]] or [10, 11, 12]:
    p = sweep_params(n_sea, 50e3, t_final=30.0, steps=20000)
    assert api.resolve_method(p, None) == ("taylor", True)
    t0 = time.perf_counter()
    _, out, st = simulate_rare(p, return_stats=True, raw_norm=True)
    wall = time.perf_counter() - t0
    w = sweep_params(n_sea, 50e3, t_final=3.0e-4, steps=3)
    t0 = time.perf_counter()
    _, _, ss = simulate_rare(w, method="taylor", return_stats=True)
    wall_s = time.perf_counter() - t0
    print(json.dumps({"n_qubits": n_sea + 1, "outputs": 20000, "wall_s": round(wall, 3), "device_ms": round(st.device_ms, 1),
                      "path": st.path, "column_h_applies": st.n_rhs, "integrator_steps": st.n_accepted + st.n_rejected,
                      "max_raw_norm_dev": float(abs(out["raw_norm"] - 1.0).max()),
                      "norm_bound_t": compile_table(p).spectral_bound() * 30.0,
                      "stepped_3e-4s_window": {"wall_s": round(wall_s, 3), "h_applies": ss.n_rhs,
                                               "extrapolated_full_trajectory_s": round(wall_s * 30.0 / 3.0e-4, 0)}}), flush=True)
