#!/usr/bin/env python
"""Wall / device time of the propagator path on the C2 sweep and the C3 grid, CI-sized and at the reference's full
trajectory size (t_final = 30 s, 20000 outputs), split into the column-integration and the matvec-chain launches.

    python scripts/time_propagator.py [--no-grid]
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from bench import F1A, sweep_params  # noqa: E402
from quantumsimulations_b200 import api, compile_table, engine, simulate_rare_batch  # noqa: E402

V = ("center_off", "center_on", "shell_off")


def c2(**kw):
    return [sweep_params(6, d, v, **kw) for d in np.linspace(0.0, 3.0 * F1A, 13) for v in V]


def c3(**kw):
    return [sweep_params(6, d, v, f1a=f, **kw) for f in np.arange(5e3, 50e3 + 1, 2.5e3)
            for d in np.linspace(0.0, 3.0 * f, 13) for v in V]


def timed(plist, **kw):
    t0 = time.perf_counter()
    res, st = simulate_rare_batch(plist, **kw)
    return round(time.perf_counter() - t0, 4), round(max(s.device_ms for s in st), 2), res, st


FULL = dict(t_final=30.0, steps=20000)
simulate_rare_batch(c2()[:3], method="propagator")  # context + module load
out = {}
out["c2_ci_taylor_s"], out["c2_ci_taylor_kernel_ms"], ref, _ = timed(c2(), method="taylor")
out["c2_ci_propagator_s"], out["c2_ci_propagator_kernel_ms"], res, _ = timed(c2(), method="propagator")
out["c2_ci_max_dev_vs_taylor"] = max(float(np.max(np.abs(a[1][k] - b[1][k]))) for a, b in zip(ref, res) for k in api.TRACE_KEYS)
out["c2_full_propagator_s"], out["c2_full_propagator_kernel_ms"], res, st = timed(c2(**FULL), method="propagator", raw_norm=True)
out["c2_full_column_h_applies"] = int(sum(s.n_rhs for s in st))
out["c2_full_max_raw_norm_dev"] = max(float(np.max(np.abs(r[1]["raw_norm"] - 1.0))) for r in res)
# columns alone: the same call with 1 output
plist = c2(**FULL)
tables = [compile_table(p) for p in plist]
opts = [api._options_from(p, "propagator")[0] for p in plist]
dt = 30.0 / 19999
out["c2_full_squarings"] = [engine.propagator_squarings(t, dt) for t in tables[::3]]
for name, n_out, cap in (("build_only", 1, None), ("build_plus_chain", 20000, None), ("build_only_no_squaring", 1, 0)):
    t0 = time.perf_counter()
    _, st = engine.propagate_batch(tables, n_out, dt, opts, max_squarings=cap)
    out[f"c2_full_{name}_s"] = round(time.perf_counter() - t0, 4)
    out[f"c2_full_{name}_kernel_ms"] = round(st[0].device_ms, 2)
    out[f"c2_full_{name}_column_h_applies"] = int(sum(s.n_rhs for s in st))
if "--no-grid" not in sys.argv:
    out["c3_ci_taylor_s"], out["c3_ci_taylor_kernel_ms"], _, _ = timed(c3(), method="taylor")
    out["c3_ci_propagator_s"], out["c3_ci_propagator_kernel_ms"], _, _ = timed(c3(), method="propagator")
    out["c3_full_propagator_s"], out["c3_full_propagator_kernel_ms"], _, _ = timed(c3(**FULL), method="propagator")
print(json.dumps(out))
