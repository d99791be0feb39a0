"""Trajectory checkpoint / resume and the per-trajectory result cache.

The reference keeps no solver state: every run of ``run_sweep_sea_detuning`` names a fresh directory
(sweep_sea_detuning.py:483-485) and every ``simulate_rare`` call integrates from t = 0
(dipolar_ensemble_with_rare.py:653).  Two granularities are offered here:

* ``TrajectoryCache`` - finished trajectories keyed by their parameters.  An interrupted sweep (39 or 741
  one-CTA trajectories) re-integrates only what is missing; this is the unit that matters for the on-chip path,
  whose whole trajectory is one kernel launch.
* ``evolve_checkpointed`` - the tiled path (N >= 11, state resident in HBM): every ``every`` output times the
  session's (t, controller step size, counters, state vector) is written to one ``.npz``; a later call with the
  same arguments picks the trajectory up there.  A checkpoint is always taken AT an output time, where the
  controller has just accepted a step, so the resumed run makes the same step-size decisions: its traces are
  bitwise those of the uninterrupted run (tests/test_gpu_evolve.py).

Both write atomically (temporary file + rename).
"""
from __future__ import annotations

import hashlib
import json
import os
import tempfile
from dataclasses import asdict, replace
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

from .model import DipolarRareParams

FORMAT = 2


def _atomic_savez(path: str, _before_rename: Optional[Callable[[], None]] = None, **arrays) -> None:
    """Write to a temporary file, then rename.  ``_before_rename`` runs between the two phases (a sharded run puts
    its rank barrier there, so either every rank's file is replaced or - if a rank died while writing - none is)."""
    d = os.path.dirname(os.path.abspath(path)) or "."
    os.makedirs(d, exist_ok=True)
    fd, tmp = tempfile.mkstemp(prefix=".tmp_", suffix=".npz", dir=d)
    os.close(fd)
    try:
        np.savez(tmp, **arrays)
        if _before_rename is not None:
            _before_rename()
        os.replace(tmp, path)
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)


def _plain(v):
    """JSON-stable form of one field: floats by their exact hex representation, numpy scalars coerced first."""
    if isinstance(v, (bool, np.bool_)):
        return bool(v)
    if isinstance(v, (int, np.integer)):
        return int(v)
    if isinstance(v, (float, np.floating)):
        return float(v).hex()
    return v


def effective_solver(params: DipolarRareParams, method: Optional[str] = None, tol_scale: Optional[float] = None) -> Dict:
    """Everything besides ``params`` that selects the numerics of ``simulate_rare(params, method=, tol_scale=)``:
    the RESOLVED integrator (``auto`` and the QSB200_METHOD / QSB200_TOL_SCALE / QSB200_TAYLOR_ORDER overrides are
    resolved here, not stored as given), whether the propagator path runs, order, effective tolerances and the
    library version."""
    from . import api, engine

    integrator, propagate = api.resolve_method(params, method)
    o, _ = api._options_from(params, method, tol_scale)
    return {"integrator": integrator, "propagator": bool(propagate), "order": int(o.order), "atol": float(o.atol),
            "rtol": float(o.rtol), "qsb_version": int(engine.load_library().qsb_version())}


def trajectory_key(params: DipolarRareParams, **solver) -> str:
    """Stable hash of everything that determines a trajectory's traces: the parameter record (exact float
    representation, as ``params_{tag}.json`` stores it) and the solver choice.  ``method`` / ``tol_scale`` keywords
    are resolved through ``effective_solver`` first, so changing the numerics - by argument, by environment override
    or by upgrading the library - never serves stale traces."""
    solver = dict(solver)
    if solver.pop("_resolve", True):
        solver.update(effective_solver(params, solver.pop("method", None), solver.pop("tol_scale", None)))
    payload = {"format": FORMAT, "params": {k: _plain(v) for k, v in asdict(params).items()},
               "solver": {k: _plain(v) for k, v in sorted(solver.items())}}
    return hashlib.sha256(json.dumps(payload, sort_keys=True).encode()).hexdigest()[:32]


class TrajectoryCache:
    """Directory of finished trajectories, one ``{key}.npz`` each (t + the 7 traces of simulate_rare)."""

    def __init__(self, root: str):
        self.root = root
        os.makedirs(root, exist_ok=True)

    def path(self, key: str) -> str:
        return os.path.join(self.root, key + ".npz")

    def get(self, key: str) -> Optional[Tuple[np.ndarray, Dict[str, np.ndarray]]]:
        p = self.path(key)
        if not os.path.exists(p):
            return None
        with np.load(p) as z:
            return z["t"], {k: z[k] for k in z.files if k != "t"}

    def put(self, key: str, t: np.ndarray, obs: Dict[str, np.ndarray]) -> None:
        _atomic_savez(self.path(key), t=t, **obs)


def cached_batch(simulate_batch: Callable[[Sequence[DipolarRareParams]], List], cache: TrajectoryCache,
                 **solver) -> Callable[[Sequence[DipolarRareParams]], List]:
    """Wrap a batch solver (list[params] -> list[(t, obs)]) so that trajectories already in ``cache`` are served
    from disk and only the missing ones are integrated (in ONE batch) and then stored.  ``solver``: whatever else
    selects the numerics (method=..., tol_scale=...) - it becomes part of the key."""

    def run(plist: Sequence[DipolarRareParams]) -> List:
        keys = [trajectory_key(p, **solver) for p in plist]
        out: List = [cache.get(k) for k in keys]
        todo = [i for i, r in enumerate(out) if r is None]
        # identical parameter records inside one batch are integrated once
        first: Dict[str, int] = {}
        uniq = [i for i in todo if first.setdefault(keys[i], i) == i]
        if uniq:
            for i, r in zip(uniq, simulate_batch([plist[i] for i in uniq])):
                cache.put(keys[i], r[0], r[1])
                out[i] = r
            for i in todo:
                if out[i] is None:
                    out[i] = out[first[keys[i]]]
        return out

    return run


def _identity(plan, opts, psi0: Optional[np.ndarray]) -> str:
    """Digest of what a checkpoint must agree with besides the time grid and tolerances: the coupling table as the C
    ABI sees it (h, J, masks, n), the initial state, max_step / nsteps, and the rank layout."""
    bm = plan.table.bit_major()
    hsh = hashlib.sha256()
    hsh.update(np.ascontiguousarray(bm["h"], dtype=np.float64).tobytes())
    hsh.update(np.ascontiguousarray(bm["J"], dtype=np.float64).tobytes())
    hsh.update(json.dumps([int(plan.table.n), int(plan.table.sea_mask), int(plan.table.rare_mask),
                           float(opts.max_step or 0.0).hex(), int(opts.nsteps or 0),
                           int(getattr(plan, "rank", 0)), int(getattr(plan, "world", 1))]).encode())
    if psi0 is None:
        hsh.update(b"basis:%d" % int(plan.table.psi0_index))
    else:
        hsh.update(np.ascontiguousarray(psi0, dtype=np.complex128).tobytes())
    return hsh.hexdigest()[:32]


def evolve_checkpointed(plan, t_out: np.ndarray, opts, path: str, every: int = 1, psi0: Optional[np.ndarray] = None,
                        stop_after: Optional[int] = None):
    """``Plan.evolve`` for the tiled path with a checkpoint file.

    Returns ``(obs[n_out, NOBS], stats, done)``; ``done`` is False when ``stop_after`` output times were processed
    in this call and the trajectory is not finished yet (call again to continue - that is also how tests
    simulate an interruption).  For a sharded plan every rank passes its own ``path``.
    """
    from . import engine

    t_out = np.ascontiguousarray(t_out, dtype=np.float64)
    n_out = len(t_out)
    obs = np.full((n_out, engine.NOBS), np.nan)
    k0 = 0
    counters = np.zeros(3, dtype=np.int64)
    ident = _identity(plan, opts, psi0)
    world = int(getattr(plan, "world", 1))
    have = os.path.exists(path)
    if world > 1:
        # every rank must resume from the SAME output time (a crash between two ranks' writes leaves different files)
        flags = plan.allgather([1.0 if have else 0.0])
        if len(set(flags.ravel().tolist())) != 1:
            raise ValueError(f"{path}: some ranks have a checkpoint and some do not - remove the files and restart")
    if have:
        with np.load(path) as z:
            if int(z["format"]) != FORMAT or not np.array_equal(z["t_out"], t_out) or str(z["method"]) != opts.method \
                    or int(z["order"]) != int(opts.order) or float(z["atol"]) != float(opts.atol) or float(z["rtol"]) != float(opts.rtol):
                raise ValueError(f"{path}: checkpoint belongs to a different trajectory (time grid or solver options differ)")
            if str(z["identity"]) != ident:
                raise ValueError(f"{path}: checkpoint belongs to a different trajectory (Hamiltonian table, initial "
                                 "state, max_step, nsteps or rank/world differ)")
            k0 = int(z["k_done"])
            h_saved = float(z["h"])
            if world > 1:
                got = plan.allgather([float(k0), h_saved])
                if np.any(got != got[0]):
                    raise ValueError(f"{path}: the ranks' checkpoints were taken at different output times "
                                     f"(k_done = {got[:, 0].astype(int).tolist()}) - remove the files and restart")
            obs[:k0 + 1] = z["obs"]
            counters = z["counters"].astype(np.int64)
            plan.session_begin(replace(opts, first_step=h_saved), t0=float(t_out[k0]), psi0=z["psi"])
    else:
        plan.session_begin(opts, t0=float(t_out[0]), psi0=psi0)
        obs[0] = plan.session_observe()
    barrier = (lambda: plan.allgather([0.0])) if world > 1 else None
    st = None
    processed = 0
    try:
        for k in range(k0 + 1, n_out):
            if stop_after is not None and processed >= stop_after:
                break
            _, st = plan.session_advance(float(t_out[k]))
            obs[k] = plan.session_observe()
            processed += 1
            if (k - k0) % max(1, int(every)) == 0 or k == n_out - 1 or (stop_after is not None and processed >= stop_after):
                tot = counters + np.array([st.n_accepted, st.n_rejected, st.n_rhs], dtype=np.int64)
                _atomic_savez(path, barrier, format=FORMAT, t_out=t_out, k_done=k, obs=obs[:k + 1], h=st.last_h,
                              psi=plan.session_state(), counters=tot, method=opts.method, order=int(opts.order),
                              atol=float(opts.atol), rtol=float(opts.rtol), identity=ident)
        done = not np.isnan(obs[-1, 0])
    finally:
        plan.session_end()
    total = counters if st is None else counters + np.array([st.n_accepted, st.n_rejected, st.n_rhs], dtype=np.int64)
    stats = engine.Stats(int(total[0]), int(total[1]), int(total[2]), st.last_h if st else 0.0, 0.0, 1, 0)
    return obs, stats, done
