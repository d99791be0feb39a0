#!/usr/bin/env python
"""Sharded H-applies with ranks as threads of ONE process (all on device 0): ring (TMA) vs register-staged centre pass.

    python scripts/shard_ring_check.py --n 28 --world 2
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from bench import sweep_params  # noqa: E402
from quantumsimulations_b200 import compile_table, engine, shard  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=28)
ap.add_argument("--world", type=int, default=2)
a = ap.parse_args()
table = compile_table(sweep_params(a.n - 1, 50e3))
opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, max_step=1e-5, nsteps=0, method="taylor", order=4)
t_out = np.array([0.0, 2.0e-8])
with shard.LocalShardGroup(table, a.world, devices=[0] * a.world) as grp:
    obs, st = grp.evolve(t_out, opts)
print(json.dumps({"n": a.n, "world": a.world, "ring": os.environ.get("QSB200_SHARD_RING", "1"), "obs": obs[-1].tolist(),
                  "layouts": engine.tile_layouts(a.n, a.world), "rhs": st.n_rhs}))
