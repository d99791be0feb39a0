#!/usr/bin/env python
"""Time the tiled H-apply at N spins through a Taylor session (order p => p H-applies per attempt).

    QSB200_FUSE_LB=0 python scripts/time_happly.py --n 28 --attempts 3 [--profile]

Prints ms per H-apply (CUDA events on the plan's stream) and, with --profile, the per-kernel-class times.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from bench import sweep_params  # noqa: E402
from quantumsimulations_b200 import compile_table, engine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=28)
ap.add_argument("--attempts", type=int, default=3)
ap.add_argument("--order", type=int, default=8)
ap.add_argument("--method", default="taylor")
ap.add_argument("--profile", action="store_true")
a = ap.parse_args()

table = compile_table(sweep_params(a.n - 1, 50e3))
opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, max_step=1e-5, nsteps=0, method=a.method, order=a.order)
with engine.Plan(table) as plan:
    plan.session_begin(opts)
    plan.session_step(1)
    t, st0, _ = plan.session_step(0)
    t, st, ms = plan.session_step(a.attempts)
    n_rhs = st.n_rhs - st0.n_rhs
    out = {"n": a.n, "fuse_lb": os.environ.get("QSB200_FUSE_LB"), "fuse_lag": os.environ.get("QSB200_FUSE_LAG"), "method": a.method, "ms_per_attempt": ms / a.attempts,
           "ms_per_happly": ms / n_rhs, "n_rhs": n_rhs, "obs": [float(x) for x in plan.session_observe()[:4]]}
    if a.profile:
        engine.profile(True)
        plan.session_step(1)
        pms, cnt, byt = engine.profile_read(with_bytes=True)
        engine.profile(False)
        out["profile_ms_per_launch"] = {k: round(pms[k] / cnt[k], 3) for k in pms if cnt[k]}
        out["profile_GBps"] = {k: round(byt[k] / pms[k] / 1e6) for k in pms if cnt[k]}
        out["profile_launches"] = {k: cnt[k] for k in pms if cnt[k]}
    plan.session_end()
print(json.dumps(out))
