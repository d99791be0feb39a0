"""GPU experiment: accuracy (vs reference-generated exact goldens) and wall time of the on-chip path per method.

    python scripts/sweep_methods.py [--short]
"""
import dataclasses
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from quantumsimulations_b200 import DipolarRareParams, simulate_rare_batch  # noqa: E402

KEYS = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")
meta = json.load(open(os.path.join(ROOT, "tests/golden/traj_cases.json")))
arrs = np.load(os.path.join(ROOT, "tests/golden/traj_cases.npz"))
names = [k for k in meta if k.startswith("n7_d")]
short = "--short" in sys.argv


def run(method, order, atol, rtol):
    os.environ["QSB200_TAYLOR_ORDER"] = str(order)
    plist = []
    for k in names:
        p = DipolarRareParams(**meta[k]["params"])
        p = dataclasses.replace(p, solver_atol=atol, solver_rtol=rtol)
        if short:
            p = dataclasses.replace(p, t_final=p.t_final / 10, steps=(p.steps - 1) // 10 + 1)
        plist.append(p)
    t0 = time.perf_counter()
    res, stats = simulate_rare_batch(plist, method=method)
    wall = time.perf_counter() - t0
    dev = 0.0
    if not short:
        for k, (t, out) in zip(names, res):
            dev = max(dev, max(float(np.max(np.abs(out[q] - arrs[f"{k}/{q}"]))) for q in KEYS))
    steps = sum(s.n_accepted + s.n_rejected for s in stats)
    rhs = sum(s.n_rhs for s in stats)
    mx = max(s.n_accepted + s.n_rejected for s in stats)
    print(f"{method:7s} p={order:2d} atol={atol:.0e} rtol={rtol:.0e}  wall={wall:7.3f}s kernel={stats[0].device_ms/1e3:7.3f}s "
          f"steps={steps:9d} max_steps={mx:8d} rhs={rhs:10d} rej={sum(s.n_rejected for s in stats):7d} max|E-X|={dev:.2e}", flush=True)


if __name__ == "__main__":
    run("dp5", 0, 1e-10, 1e-9)
    if "--only-dp5" in sys.argv:
        sys.exit(0)
    run("dp5", 0, 1e-13, 1e-12)
    for order in (8, 12, 16, 20):
        run("taylor", order, 1e-10, 1e-9)
    for order in (12, 16, 20):
        run("taylor", order, 1e-13, 1e-12)
