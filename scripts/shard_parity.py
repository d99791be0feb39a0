#!/usr/bin/env python
"""Multi-process parity of the sharded-state path (run under torchrun, one rank per GPU, NCCL):

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/shard_parity.py [--qubits 20]

Every rank also integrates the SAME trajectory unsharded on its own GPU and compares: the observables must agree to
rounding (same kernels, different partition) and the gathered sharded state must equal the unsharded one.
Prints one JSON line on rank 0 and exits non-zero on any mismatch.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from bench import sweep_params  # noqa: E402
from quantumsimulations_b200 import compile_table, engine, shard  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--qubits", dest="n", type=int, default=20)
ap.add_argument("--method", default="taylor")
a = ap.parse_args()
if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
    os.environ["NCCL_DEBUG"] = "WARN"
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()

table = compile_table(sweep_params(a.n - 1, 50e3))
t = np.linspace(0.0, 4e-7, 5)
opts = engine.SolverOptions(atol=1e-12, rtol=1e-11, max_step=1e-5, nsteps=100000, method=a.method, order=12)
with engine.Plan(table, local) as plan:
    ref, st_ref = plan.evolve(t, opts)
    plan.session_begin(opts)
    plan.session_step(2)
    ref_state = plan.session_state()
    plan.session_end()
sp = shard.sharded_plan_from_env(table, local)
got, st = sp.evolve(t, opts)
sp.session_begin(opts)
sp.session_step(2)
mine = sp.session_state()
sp.session_end()
sl = sp.local_slice()
sp.close()
d_obs = float(np.max(np.abs(got - ref)))
d_state = float(np.max(np.abs(mine - ref_state[sl])))
ok = d_obs < 1e-11 and d_state < 1e-13 and st.n_accepted == st_ref.n_accepted and st.n_rejected == st_ref.n_rejected
flags = [None] * world
dist.all_gather_object(flags, (ok, d_obs, d_state))
if rank == 0:
    print(json.dumps({"n": a.n, "world": world, "method": a.method, "steps": st.n_accepted,
                      "max_obs_diff": max(f[1] for f in flags), "max_state_diff": max(f[2] for f in flags),
                      "ok": all(f[0] for f in flags)}), flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if all(f[0] for f in flags) else 1)
