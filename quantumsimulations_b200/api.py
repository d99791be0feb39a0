"""Reference-facing call surface: ``build_hamiltonian_rare``, ``initial_state_rare``, ``simulate_rare``
(same names, argument meaning, return shapes and error behaviour as dipolar_ensemble_with_rare.py:453-680)
plus the additive batch entry point ``simulate_rare_batch``.

QuTiP is not a dependency: instead of ``qt.Qobj`` the builder returns ``SpinOperator`` handles around
the compiled coupling table (dense / CSR views are available for small N).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import engine
from .model import DipolarRareParams
from .table import SpinTable, compile_table, local_dim4

TRACE_KEYS = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")


class SpinOperator:
    """Matrix-free handle for H or one of the six observables (stands in for qt.Qobj, DE:570-587)."""

    def __init__(self, table: SpinTable, kind: str = "H"):
        self.table = table
        self.kind = kind
        # a spin-3/2 rare spin is the last tensor factor, local dimension 4 (two index bits; DE:28-34)
        local = [2] * table.n if table.pauli_terms is None else [2] * (table.n - 2) + [4]
        self.dims = [local, list(local)]
        self.shape = (table.dim, table.dim)

    def _as_table(self) -> SpinTable:
        if self.kind == "H":
            return self.table
        t = self.table
        if t.pauli_terms is not None and self.kind in ("Ix_R", "Iy_R", "Iz_R"):
            raise NotImplementedError(f"{self.kind} of a spin-3/2 rare spin is not a one-bit operator: use to_csr()")
        n = t.n
        z = np.zeros(n)
        zz = np.zeros((n, n))
        fields = {"x": z.copy(), "y": z.copy(), "z": z.copy()}
        axis = self.kind[1]
        sites = t.sea_sites if self.kind.endswith("_sea") else [t.rare_site]
        fields[axis][sites] = 1.0
        return SpinTable(n, fields["x"], fields["y"], fields["z"], zz, zz.copy(), zz.copy(), t.sea_sites, t.rare_site,
                         t.psi0_index)

    def apply(self, psi: np.ndarray, device: int = 0) -> np.ndarray:
        """H @ psi on the GPU (one H-apply through the engine's kernels).  A local-dimension-4 table lives in the frame
        rotating about J_z (table.py): H_lab psi = R H_table R^dagger psi with the diagonal R applied here."""
        r = self.table.frame_phases() if self.kind == "H" else None
        with engine.Plan(self._as_table(), device) as plan:
            if r is None:
                return plan.apply_h(psi)
            return r * plan.apply_h(np.conj(r) * np.asarray(psi, dtype=np.complex128))

    def _pauli_csr(self):
        """Lab-frame observable of a local-dimension-4 table from its weighted Pauli strings (host, inspection only)."""
        import scipy.sparse as sp

        t = self.table
        col = TRACE_KEYS.index(self.kind)
        dim = t.dim
        x = np.arange(dim, dtype=np.int64)
        m = sp.csr_matrix((dim, dim), dtype=complex)
        for xm, zm, w, c in t.pauli_terms:
            if c != col:
                continue
            par = np.zeros(dim, dtype=np.int64)
            for b in range(t.n):
                if (zm >> b) & 1:
                    par ^= (x >> b) & 1
            val = w * (1j ** bin(xm & zm).count("1")) * (1.0 - 2.0 * par)       # P |x> = i^ny (-1)^popc(x & zm) |x ^ xm>
            m = m + sp.coo_matrix((val, (x ^ xm, x)), shape=(dim, dim)).tocsr()
        r = t.frame_phases()
        return (sp.diags(r) @ m @ sp.diags(np.conj(r))).tocsr()

    def to_csr(self):
        """scipy CSR matrix (host, small N only; for inspection and tests - not used by the solver)."""
        import scipy.sparse as sp

        if self.table.n > 16:
            raise MemoryError("refusing to materialise a 2^N x 2^N operator for N > 16")
        if self.kind != "H" and self.table.pauli_terms is not None:
            return self._pauli_csr()
        t = self._as_table()
        bm = t.bit_major()
        h, J = bm["h"], bm["J"]
        n, dim = t.n, t.dim
        x = np.arange(dim, dtype=np.int64)
        s = np.stack([1.0 - 2.0 * ((x >> b) & 1) for b in range(n)])
        diag = 0.5 * (h[2] @ s)
        rows, cols, vals = [], [], []
        for b in range(n):
            for c in range(b + 1, n):
                if J[2, b, c] != 0.0:
                    diag = diag + 0.25 * J[2, b, c] * s[b] * s[c]
                if J[0, b, c] != 0.0 or J[1, b, c] != 0.0:
                    rows.append(x)
                    cols.append(x ^ (1 << b) ^ (1 << c))
                    vals.append((0.25 * (J[0, b, c] - s[b] * s[c] * J[1, b, c])).astype(complex))
            if h[0, b] != 0.0 or h[1, b] != 0.0:
                rows.append(x)
                cols.append(x ^ (1 << b))
                vals.append(0.5 * (h[0, b] - 1j * s[b] * h[1, b]))
        rows.append(x)
        cols.append(x)
        vals.append(diag.astype(complex))
        m = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(dim, dim)).tocsr()
        if self.kind == "H" and t.frame_phi is not None:       # back to the lab frame (table.py)
            r = t.frame_phases()
            m = (sp.diags(r) @ m @ sp.diags(np.conj(r))).tocsr()
        m.eliminate_zeros()
        return m

    def full(self) -> np.ndarray:
        return self.to_csr().toarray()


def build_hamiltonian_rare(params: DipolarRareParams) -> Tuple[SpinOperator, Dict[str, SpinOperator]]:
    """(H, {Ix_sea, Iy_sea, Iz_sea, Iz_R, Ix_R, Iy_R}) - compiled, never materialised (DE:453-588)."""
    table = compile_table(params)
    obs = {k: SpinOperator(table, k) for k in TRACE_KEYS}
    return SpinOperator(table, "H"), obs


def initial_state_rare(params: DipolarRareParams) -> np.ndarray:
    """The product state of DE:591-606 as a dense ket (2^N complex128); the engine itself only needs its index."""
    table = compile_table(params)
    psi = np.zeros(table.dim, dtype=np.complex128)
    psi[table.psi0_index] = 1.0
    return psi


# Default integrator of the reference-facing calls.  H is time-independent in the reference (DE:514, DE:570), so
# the order-p Taylor pair applies; measured on the sweep's own 12 golden points (profiles/r1_methods.log) it needs
# ~10x fewer H-applies than DP5 at equal tolerances.  Its cost grows only logarithmically with the tolerance
# (1.4x for 1000x tighter), while the reference's own trajectories are integrator-error-limited at its tolerances
# (SURVEY.md TL;DR 6), so the requested atol/rtol are tightened by DEFAULT_TOL_SCALE before they reach the C ABI.
# "auto" = the Taylor integrator, used through the propagator path (below) wherever that is the cheaper way to the
# same traces: N <= 10 and more output intervals than basis states (steps - 1 > 2^N).
DEFAULT_METHOD = "auto"
DEFAULT_TAYLOR_ORDER = 20
DEFAULT_TOL_SCALE = {"taylor": 1e-3, "dp5": 1.0}
# method = "propagator" (or "propagator-dp5"): N <= 10 only.  The reference's output grid is uniform (DE:626) and its
# H constant, so U = exp(-i H dt) is integrated ONCE (2^N column trajectories over one output interval, by the
# integrator named after the dash, default taylor) and every output is one dense matvec (engine.propagate_batch).
# Cheaper than stepping whenever steps - 1 > 2^N; the sweep's own 20000-output, 30 s trajectories (SW:1201-1252) drop
# from ~10^9 H-applies to 128 integrations over 1.5 ms + 2 x 10^4 matvecs.
PROPAGATOR = "propagator"
# The propagator's error is applied steps - 1 times, so the columns are integrated tighter than a stepped trajectory
# would be: measured at the reference's full size (20000 applications) tol x 1e-3 leaves 2.3e-6 on the traces, and the
# Taylor pair's cost grows only logarithmically with the tolerance.
PROPAGATOR_TOL_SCALE = {"taylor": 1e-5, "dp5": 1.0}


# "auto" for 11 <= N <= 14: the HBM-resident propagator (engine.propagate_hbm) costs a few seconds per trajectory
# whatever the trajectory; stepping costs ~3 H-applies per unit of ||H|| t at ~25 us each at these sizes.  The
# propagator is picked when the cheap bound ||H|| t_final (SpinTable.spectral_bound) says stepping would take longer.
AUTO_HBM_NORM_T = 2.0e5


def split_method(method: Optional[str], n_qubits: Optional[int] = None, steps: Optional[int] = None,
                 norm_t: Optional[float] = None) -> Tuple[str, bool]:
    """(integrator, use_propagator) from a method string: "auto", "taylor", "dp5", "propagator", "propagator-dp5".
    "auto" needs the trajectory's size (``n_qubits``, ``steps``; for 11 <= N <= 14 also ``norm_t`` = bound on
    ||H|| t_final) to choose; without it, it means stepping."""
    m = method or os.environ.get("QSB200_METHOD", DEFAULT_METHOD)
    if m == "auto":
        if n_qubits is None or steps is None:
            return "taylor", False
        if n_qubits <= engine.SMALL_MAX_QUBITS:
            return "taylor", bool(steps - 1 > (1 << n_qubits))
        if n_qubits <= engine.HBM_PROPAGATOR_MAX_QUBITS:
            return "taylor", bool(norm_t is not None and norm_t > AUTO_HBM_NORM_T and steps >= 3)
        return "taylor", False
    if m.startswith(PROPAGATOR):
        rest = m[len(PROPAGATOR):].lstrip("-")
        m, prop = (rest or "taylor"), True
    else:
        prop = False
    if m not in DEFAULT_TOL_SCALE:
        raise ValueError(f"unknown method {method!r}: auto, taylor, dp5, propagator or propagator-dp5")
    return m, prop


def resolve_method(params: DipolarRareParams, method: Optional[str]) -> Tuple[str, bool]:
    """``split_method`` for one parameter record (computes the ||H|| t_final bound only where "auto" needs it).
    A spin-3/2 rare spin (local dimension 4): N <= 10 bits always goes through the propagator (the stepped on-chip
    kernel cannot report its observables); 11 <= N <= 14 like the spin-1/2 tables (HBM-resident propagator when asked
    for, or when "auto" finds stepping the expensive way); beyond that a stepping session (``_evolve_pauli``)."""
    n, steps = _size_of(params)
    if local_dim4(params):
        integrator, asked = split_method(method)
        if n <= engine.SMALL_MAX_QUBITS:
            return integrator, True
        if n > engine.HBM_PROPAGATOR_MAX_QUBITS:
            if asked:
                raise ValueError(f"method 'propagator' with a spin-3/2 rare spin needs n_sea + 2 <= {engine.HBM_PROPAGATOR_MAX_QUBITS}")
            return integrator, False
        if asked:
            return integrator, True
        if (method or os.environ.get("QSB200_METHOD", DEFAULT_METHOD)) != "auto":
            return integrator, False
        norm_t = compile_table(params).spectral_bound() * float(params.t_final)     # the rule of the spin-1/2 tables
        return integrator, bool(norm_t > AUTO_HBM_NORM_T and steps >= 3)
    norm_t = None
    m = method or os.environ.get("QSB200_METHOD", DEFAULT_METHOD)
    if m == "auto" and engine.SMALL_MAX_QUBITS < n <= engine.HBM_PROPAGATOR_MAX_QUBITS:
        norm_t = compile_table(params).spectral_bound() * float(params.t_final)
    return split_method(method, n, steps, norm_t)


def _size_of(params: DipolarRareParams) -> Tuple[int, int]:
    """(index bits, outputs): the spin-3/2 rare spin takes two bits (table.py)."""
    return params.n_sea + (2 if local_dim4(params) else 1), params.steps


def _options_from(params: DipolarRareParams, method: Optional[str], tol_scale: Optional[float] = None
                  ) -> Tuple[engine.SolverOptions, bool]:
    """Map the reference's solver_* fields (DE:629-651); returns (options, any_override)."""
    integrator, propagate = resolve_method(params, method)
    knobs = (params.solver_atol, params.solver_rtol, params.solver_nsteps, params.solver_max_step)
    override = any(v is not None for v in knobs)
    o = engine.SolverOptions()
    if params.solver_atol is not None:
        o.atol = params.solver_atol
    if params.solver_rtol is not None:
        o.rtol = params.solver_rtol
    if params.solver_nsteps is not None:
        o.nsteps = int(params.solver_nsteps)
    if params.solver_max_step is not None:
        o.max_step = params.solver_max_step
    o.method = integrator
    o.order = int(os.environ.get("QSB200_TAYLOR_ORDER", str(DEFAULT_TAYLOR_ORDER)))
    if tol_scale is None:
        table = PROPAGATOR_TOL_SCALE if propagate else DEFAULT_TOL_SCALE
        tol_scale = float(os.environ.get("QSB200_TOL_SCALE", table[o.method]))
    o.atol = max(o.atol * tol_scale, 1e-300)
    o.rtol = max(o.rtol * tol_scale, 1e-14) if o.rtol > 0 else 0.0
    return o, override


def _check_grid(params: DipolarRareParams) -> np.ndarray:
    if params.steps < 2 or params.t_final <= 0.0:
        raise ValueError("Bad time grid: steps >= 2 and t_final > 0.")
    return np.linspace(0.0, params.t_final, params.steps)


def _pack(obs: np.ndarray, override: bool, extra: bool, by_column: bool = False) -> Dict[str, np.ndarray]:
    """``obs``: (n_out, NOBS), or (NOBS, n_out) with ``by_column`` (the propagator path's device layout: no copy)."""
    cols = obs if by_column else np.ascontiguousarray(obs.T)   # the traces are the (contiguous) rows of cols
    out = {k: cols[i] for i, k in enumerate(TRACE_KEYS)}
    # DE:669: norms of the stored (output-normalised) states; no states are stored when options is None
    out["state_norm"] = cols[6] if override else np.array([])
    if extra:
        out["raw_norm"] = cols[7]
    return out


def simulate_rare(params: DipolarRareParams, *, device: int = 0, method: Optional[str] = None,
                  tol_scale: Optional[float] = None, return_stats: bool = False, raw_norm: bool = False,
                  checkpoint: Optional[str] = None, checkpoint_every: int = 16):
    """Time evolution on the B200 engine; returns ``(t, {6 traces + state_norm})`` like DE:611-680.

    ``method``: "auto" (default: Taylor order 20, through the propagator path when N <= 10 and steps - 1 > 2^N),
    "taylor", "dp5", or "propagator" / "propagator-dp5" (N <= 10: one propagator per trajectory, then one matvec per
    output time; see PROPAGATOR).  ``tol_scale`` multiplies solver_atol/solver_rtol before
    they reach the integrator (default 1e-3 for taylor, 1 for dp5; see DEFAULT_TOL_SCALE).
    Deviation from the reference's error behaviour: on the propagator path (what "auto" picks for N <= 10 and
    steps - 1 > 2^N) ``solver_nsteps`` / ``solver_max_step`` apply to the column integrations over dt / 2^s only, so a
    long output interval never exhausts ``nsteps`` there, where QuTiP raises IntegratorException ("Excess work") and the
    stepped integrators (``method="taylor"`` / ``"dp5"``) raise ``IntegratorError``.  The traces are the same (<= 1e-6
    from exact); ask for a stepped method if the exception itself matters.
    ``raw_norm=True`` adds the integrator's own norm under an extra key (QuTiP discards it).
    ``checkpoint``: file the tiled path (N >= 11) writes its state to every ``checkpoint_every`` output times and
    resumes from if it exists (``checkpoint.evolve_checkpointed``); the on-chip path is one launch per trajectory
    and is cached whole instead (``checkpoint.TrajectoryCache``, ``run_sweep_sea_detuning(cache_dir=...)``).
    """
    t = _check_grid(params)
    table = compile_table(params)
    opts, override = _options_from(params, method, tol_scale)
    if table.pauli_terms is not None:
        if checkpoint is not None:
            raise ValueError("checkpoint= is not available with is_spin_three_half=True")
        obs, stats = _evolve_pauli(table, t, opts, device, resolve_method(params, method)[1])
        out = _pack(obs, override, raw_norm, by_column=True)
        return (t, out, stats) if return_stats else (t, out)
    if resolve_method(params, method)[1]:
        obs, stats = _propagate([table], [t], [opts], device)
        out = _pack(obs[0], override, raw_norm, by_column=True)
        return (t, out, stats[0]) if return_stats else (t, out)
    with engine.Plan(table, device) as plan:
        if checkpoint is not None and table.n > engine.SMALL_MAX_QUBITS:
            from .checkpoint import evolve_checkpointed

            obs, stats, _ = evolve_checkpointed(plan, t, opts, checkpoint, every=checkpoint_every)
        else:
            obs, stats = plan.evolve(t, opts)
    out = _pack(obs, override, raw_norm)
    if return_stats:
        return t, out, stats
    return t, out


def _evolve_pauli(table: SpinTable, t: np.ndarray, opts: engine.SolverOptions, device: int, propagate: bool = False):
    """One trajectory of a local-dimension-4 table (spin-3/2 rare spin on two bits, DE:22-25 / DE:499-501): the time
    evolution runs through the spin-1/2 kernels in the rotated frame, the observables are the table's weighted Pauli
    strings.  N <= 10 bits: the propagator path with every state kept on the device (engine.propagate_expect; the
    reference's grid is uniform, DE:626); ``propagate`` for 11 <= N <= 14: the same through the HBM-resident propagator.
    Otherwise: a stepping session on the tiled path, advanced to every output time
    (``nsteps`` per output interval, like the stepped integrators) and read with ``Plan.session_expect``.
    Returns (obs [NOBS, n_out], Stats)."""
    if table.n <= engine.SMALL_MAX_QUBITS or propagate:
        dt = engine.uniform_step(t)
        if dt is None:
            raise ValueError("the spin-3/2 propagator paths need a uniform output grid (np.linspace, DE:626)")
        return engine.propagate_expect(table, len(t), dt, opts, device, by_column=True)
    obs = np.empty((engine.NOBS, len(t)), dtype=np.float64)
    with engine.Plan(table, device) as plan:
        plan.session_begin(opts, t0=float(t[0]))
        obs[:, 0] = plan.session_expect(table.pauli_terms)
        stats = None
        for k in range(1, len(t)):
            _, stats = plan.session_advance(float(t[k]))
            obs[:, k] = plan.session_expect(table.pauli_terms)
        plan.session_end()
    return obs, stats


def _propagate(tables, grids, opts, device):
    """Propagator path for trajectories of equal N and equal number of outputs (their dt may differ): the on-chip one
    (N <= 10, one batched call) or the HBM-resident one (11 <= N <= 14, one call per trajectory)."""
    if tables[0].n > engine.HBM_PROPAGATOR_MAX_QUBITS:
        raise ValueError(f"method 'propagator' holds a dense 2^N x 2^N matrix: N <= {engine.HBM_PROPAGATOR_MAX_QUBITS}")
    seen: Dict[Tuple[float, float, int], Optional[float]] = {}   # sweeps share one grid: check each distinct one once
    dts = []
    for g in grids:
        key = (float(g[0]), float(g[-1]), len(g))
        if key not in seen:
            seen[key] = engine.uniform_step(g)
        if seen[key] is None:
            raise ValueError("method 'propagator' needs a uniform output grid (np.linspace, DE:626)")
        dts.append(seen[key])
    if tables[0].n > engine.SMALL_MAX_QUBITS:
        res = [engine.propagate_hbm(tb, len(g), dt, o, device, by_column=True) for tb, g, dt, o in zip(tables, grids, dts, opts)]
        return np.stack([r[0] for r in res]), [r[1] for r in res]
    return engine.propagate_batch(tables, len(grids[0]), dts, opts, device, by_column=True)


def simulate_rare_batch(params_list: Sequence[DipolarRareParams], *, device: int = 0, method: Optional[str] = None,
                        tol_scale: Optional[float] = None, raw_norm: bool = False):
    """Independent trajectories; those with N <= 10 share ONE kernel launch (one CTA each), 11 <= N <= 14 on a uniform
    grid may go through the HBM-resident propagator (``resolve_method``), the rest is stepped on the tiled path."""
    n = len(params_list)
    special = [i for i, p in enumerate(params_list) if local_dim4(p)]
    if special:   # local dimension 4 (_evolve_pauli); the rest batches as usual
        rest = [i for i in range(n) if not local_dim4(params_list[i])]
        results, stats_out = [None] * n, [None] * n
        # on-chip sizes: the trajectories of a sweep share bits, grid length and Pauli strings -> ONE call per group
        groups: Dict[Tuple, List[int]] = {}
        single = []
        for i in special:
            p = params_list[i]
            if _size_of(p)[0] <= engine.SMALL_MAX_QUBITS:
                groups.setdefault((p.n_sea, int(p.steps), float(p.phi_rare)), []).append(i)
            else:
                single.append(i)
        for idxs in groups.values():
            tabs = [compile_table(params_list[i]) for i in idxs]
            om = [_options_from(params_list[i], method, tol_scale) for i in idxs]
            grids = [_check_grid(params_list[i]) for i in idxs]
            dts = [engine.uniform_step(g) for g in grids]
            obs, sts = engine.propagate_expect_batch(tabs, len(grids[0]), dts, [o for o, _ in om], device, by_column=True)
            for j, i in enumerate(idxs):
                results[i], stats_out[i] = (grids[j], _pack(obs[j], om[j][1], raw_norm, by_column=True)), sts[j]
        for i in single:
            t, out, st = simulate_rare(params_list[i], device=device, method=method, tol_scale=tol_scale,
                                       return_stats=True, raw_norm=raw_norm)
            results[i], stats_out[i] = (t, out), st
        if rest:
            r, s = simulate_rare_batch([params_list[i] for i in rest], device=device, method=method,
                                       tol_scale=tol_scale, raw_norm=raw_norm)
            for j, i in enumerate(rest):
                results[i], stats_out[i] = r[j], s[j]
        return results, stats_out
    results: List = [None] * n
    base_grids: Dict[Tuple[float, int], np.ndarray] = {}   # a sweep's trajectories share one grid: build it once
    grids = []
    for p in params_list:
        key = (float(p.t_final), int(p.steps))
        if key not in base_grids:
            base_grids[key] = _check_grid(p)
        grids.append(base_grids[key])
    tables = [compile_table(p) for p in params_list]
    om = [_options_from(p, method, tol_scale) for p in params_list]
    props = [resolve_method(p, method)[1] for p in params_list]
    if any(pr and t.n > engine.HBM_PROPAGATOR_MAX_QUBITS for pr, t in zip(props, tables)):
        raise ValueError(f"method 'propagator' holds a dense 2^N x 2^N matrix: N <= {engine.HBM_PROPAGATOR_MAX_QUBITS}")
    small = [i for i in range(n) if tables[i].n <= engine.SMALL_MAX_QUBITS]
    groups: Dict[int, List[int]] = {}
    for i in small:
        groups.setdefault((tables[i].n, len(grids[i])), []).append(i)  # uniform n -> register-resident rows
    stats_out: List = [None] * n
    for _, idxs in groups.items():
        prop = split_method(method, tables[idxs[0]].n, len(grids[idxs[0]]))[1]   # one (N, steps) per group
        if prop:
            obs, stats = _propagate([tables[i] for i in idxs], [grids[i] for i in idxs], [om[i][0] for i in idxs], device)
        else:
            tgrid = np.stack([grids[i] for i in idxs])
            obs, stats = engine.evolve_batch([tables[i] for i in idxs], tgrid, [om[i][0] for i in idxs], device)
        for j, i in enumerate(idxs):
            results[i] = (grids[i].copy(), _pack(obs[j], om[i][1], raw_norm, by_column=prop))
            stats_out[i] = stats[j]
    for i in range(n):
        if results[i] is None and props[i]:     # 11 <= N <= 14: the HBM-resident propagator, one trajectory per call
            obs, st = _propagate([tables[i]], [grids[i]], [om[i][0]], device)
            results[i] = (grids[i].copy(), _pack(obs[0], om[i][1], raw_norm, by_column=True))
            stats_out[i] = st[0]
    for i in range(n):
        if results[i] is None:
            with engine.Plan(tables[i], device) as plan:
                obs, st = plan.evolve(grids[i], om[i][0])
            results[i] = (grids[i].copy(), _pack(obs, om[i][1], raw_norm))
            stats_out[i] = st
    return results, stats_out
