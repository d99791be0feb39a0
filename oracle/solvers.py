"""Oracle solvers for d(psi)/dt = -i H psi on a fixed output grid.  TEST INFRASTRUCTURE ONLY.

X ("exact")  : eigendecomposition propagation psi(t_k) = V exp(-i E t_k) V^H psi0 for dim <= 4096,
               scipy ``expm_multiply`` on the uniform grid above that.  Ground truth.
Q ("zvode")  : the reference-equivalent.  ``qt.sesolve`` (DE:653) with QuTiP 5 defaults resolves to
               SciPy ``ode('zvode', method='adams', order=12)`` integrating to each t_k in turn with
               rhs = CSR(-iH) @ psi; options map 1:1 (atol, rtol, nsteps, max_step; first_step =
               min_step = 0) and QuTiP defaults are atol 1e-8, rtol 1e-6, nsteps 2500.
Both report expectation values on the OUTPUT-NORMALISED ket (QuTiP ``normalize_output=True``) and
also hand back the raw integrator norm, which QuTiP discards.
"""
from __future__ import annotations

import time
from typing import Dict, List, Sequence, Tuple

import numpy as np
import scipy.sparse as sp
from scipy.integrate import ode
from scipy.sparse.linalg import expm_multiply

EXACT_DENSE_LIMIT = 4096


class IntegratorError(RuntimeError):
    """zvode gave up (e.g. ``nsteps`` exhausted inside one output interval) - QuTiP raises IntegratorException."""


def _expectations(ops: Sequence[sp.spmatrix], states: np.ndarray) -> List[np.ndarray]:
    """states: (dim, nt) columns already normalised -> list of complex (nt,) traces."""
    return [np.einsum("ij,ij->j", states.conj(), op @ states) for op in ops]


def exact_states(H: sp.spmatrix, psi0: np.ndarray, tlist: np.ndarray) -> np.ndarray:
    """(dim, nt) array of psi(t_k), machine precision."""
    dim = H.shape[0]
    tlist = np.asarray(tlist, dtype=float)
    if dim <= EXACT_DENSE_LIMIT:
        Hd = H.toarray()
        if np.max(np.abs(Hd.imag)) == 0.0:
            Hd = Hd.real
        E, V = np.linalg.eigh(Hd)
        c0 = V.conj().T @ psi0
        out = np.empty((dim, len(tlist)), dtype=complex)
        step = max(1, (1 << 24) // dim)
        for a in range(0, len(tlist), step):
            ph = np.exp(-1j * np.outer(E, tlist[a : a + step]))
            out[:, a : a + step] = V @ (ph * c0[:, None])
        return out
    dt = np.diff(tlist)
    if not np.allclose(dt, dt[0], rtol=1e-9, atol=0):
        raise ValueError("expm_multiply path needs a uniform grid")
    A = (-1j) * H.tocsc()
    res = expm_multiply(A, psi0.astype(complex), start=tlist[0], stop=tlist[-1], num=len(tlist), endpoint=True)
    return np.ascontiguousarray(res.T)


def evolve_csr(
    H: sp.spmatrix,
    psi0: np.ndarray,
    tlist: np.ndarray,
    e_ops: Sequence[sp.spmatrix],
    backend: str = "exact",
    atol: float = 1e-8,
    rtol: float = 1e-6,
    nsteps: int = 2500,
    max_step: float = 0.0,
    keep_states: bool = False,
) -> Tuple[List[np.ndarray], np.ndarray, List[np.ndarray], Dict[str, float]]:
    """Returns (traces, raw_norms, normalised states (maybe empty), stats)."""
    H = sp.csr_matrix(H)
    psi0 = np.asarray(psi0, dtype=complex).ravel()
    tlist = np.asarray(tlist, dtype=float)
    nt = len(tlist)
    t_start = time.perf_counter()
    stats: Dict[str, float] = {"backend": backend}

    if backend == "exact":
        raw = exact_states(H, psi0, tlist)
    elif backend == "zvode":
        A = (-1j) * H
        n_rhs = [0]

        def rhs(t, y):
            n_rhs[0] += 1
            return A @ y

        r = ode(rhs)
        r.set_integrator(
            "zvode",
            method="adams",
            order=12,
            atol=atol,
            rtol=rtol,
            nsteps=int(nsteps),
            first_step=0,
            max_step=max_step or 0.0,
            min_step=0,
        )
        r.set_initial_value(psi0.copy(), tlist[0])
        raw = np.empty((len(psi0), nt), dtype=complex)
        raw[:, 0] = psi0
        for k in range(1, nt):
            y = r.integrate(tlist[k])
            if not r.successful():
                raise IntegratorError(
                    f"zvode failed before t={tlist[k]!r} (return code {r.get_return_code()}); "
                    "QuTiP reports this as 'Excess work done on this call'"
                )
            raw[:, k] = y
        iwork = r._integrator.iwork
        stats["n_steps"] = int(iwork[10])
        stats["n_rhs"] = int(n_rhs[0])
    else:
        raise ValueError(f"unknown backend {backend!r}")

    norms = np.sqrt(np.sum(np.abs(raw) ** 2, axis=0))
    states = raw / norms[None, :]
    traces = _expectations(e_ops, states)
    stats["wall_s"] = time.perf_counter() - t_start
    kept = [states[:, k].copy() for k in range(nt)] if keep_states else []
    return traces, norms, kept, stats
