"""numpy/scipy stand-in for the handful of QuTiP calls ``dipolar_ensemble_with_rare.py`` makes.

TEST INFRASTRUCTURE (see oracle/__init__.py).  QuTiP is not installable in this image; this module
lets the reference file be imported and executed *unmodified* so its Hamiltonian builder, initial
state and ``simulate_rare`` control flow produce the golden fixtures.  Call sites covered:
``sigmax/sigmay/sigmaz/qeye/jmat`` DE:16-25, ``tensor`` DE:45/603/606, ``basis`` DE:67/72,
``Qobj.eigenstates`` DE:93-95, Qobj ``+ - * scalar`` DE:504-568, ``Qobj(...)`` DE:570,
``sesolve`` DE:653, ``Qobj.norm`` DE:669, ``Qobj.unit`` DE:68.

``sesolve`` follows QuTiP 5 semantics as far as they are observable through DE:611-680: expectation
values on the output-normalised ket, states kept only when ``store_states`` is truthy (or no e_ops).
Which integrator runs underneath is chosen with ``SESOLVE_BACKEND``:
``"exact"`` (eigendecomposition / expm_multiply, machine precision) or ``"zvode"`` (SciPy's
zvode-Adams order 12, the integrator QuTiP's default ``method='adams'`` wraps).
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp

SESOLVE_BACKEND = "exact"

__version__ = "0.0-fake"


class Qobj:
    __array_ufunc__ = None  # make numpy scalars defer to our __rmul__/__radd__

    def __init__(self, data=None, dims=None):
        if isinstance(data, Qobj):
            self.data = data.data.copy()
            self.dims = data.dims if dims is None else dims
            return
        if sp.issparse(data):
            self.data = data.tocsr().astype(complex)
        else:
            arr = np.atleast_2d(np.asarray(data, dtype=complex))
            self.data = sp.csr_matrix(arr)
        self.dims = dims

    # ---- structure --------------------------------------------------------------------------
    @property
    def shape(self):
        return self.data.shape

    @property
    def isket(self):
        return self.data.shape[1] == 1 and self.data.shape[0] > 1

    def full(self):
        return self.data.toarray()

    def dag(self):
        return Qobj(self.data.conj().T.tocsr(), self.dims)

    # ---- arithmetic -------------------------------------------------------------------------
    def _coerce(self, other):
        if isinstance(other, Qobj):
            return other.data
        if np.isscalar(other):
            if other == 0:
                return None
            if self.data.shape[0] == self.data.shape[1]:
                return sp.identity(self.data.shape[0], dtype=complex, format="csr") * other
        raise TypeError(f"cannot combine Qobj with {type(other)}")

    def __add__(self, other):
        o = self._coerce(other)
        return Qobj(self.data.copy() if o is None else self.data + o, self.dims)

    __radd__ = __add__

    def __sub__(self, other):
        o = self._coerce(other)
        return Qobj(self.data.copy() if o is None else self.data - o, self.dims)

    def __rsub__(self, other):
        return (-self).__add__(other)

    def __neg__(self):
        return Qobj(-self.data, self.dims)

    def __mul__(self, other):
        if isinstance(other, Qobj):
            return Qobj(self.data @ other.data, self.dims)
        if np.isscalar(other):
            return Qobj(self.data * complex(other), self.dims)
        return NotImplemented

    def __rmul__(self, other):
        if np.isscalar(other):
            return Qobj(self.data * complex(other), self.dims)
        return NotImplemented

    __matmul__ = __mul__

    def __truediv__(self, other):
        return Qobj(self.data / complex(other), self.dims)

    # ---- kets / spectra ---------------------------------------------------------------------
    def norm(self):
        if self.data.shape[1] == 1 or self.data.shape[0] == 1:
            return float(np.sqrt(np.sum(np.abs(self.data.toarray()) ** 2)))
        return float(np.sum(np.linalg.svd(self.full(), compute_uv=False)))  # trace norm

    def unit(self):
        return self / self.norm()

    def eigenstates(self):
        w, v = np.linalg.eigh(self.full())
        kets = np.empty(len(w), dtype=object)
        for k in range(len(w)):
            kets[k] = Qobj(v[:, k].reshape(-1, 1))
        return w, kets


def _pauli(which):
    m = {
        "x": [[0, 1], [1, 0]],
        "y": [[0, -1j], [1j, 0]],
        "z": [[1, 0], [0, -1]],
    }[which]
    return Qobj(np.array(m, dtype=complex))


def sigmax():
    return _pauli("x")


def sigmay():
    return _pauli("y")


def sigmaz():
    return _pauli("z")


def qeye(n):
    return Qobj(sp.identity(int(n), dtype=complex, format="csr"))


def jmat(j, which):
    d = int(round(2 * j + 1))
    m = j - np.arange(d)                       # basis ordered m = j, j-1, ..., -j
    jp = np.zeros((d, d), dtype=complex)       # J+ |m> = sqrt(j(j+1)-m(m+1)) |m+1>
    for k in range(1, d):
        jp[k - 1, k] = np.sqrt(j * (j + 1) - m[k] * (m[k] + 1))
    if which == "x":
        return Qobj(0.5 * (jp + jp.conj().T))
    if which == "y":
        return Qobj(-0.5j * (jp - jp.conj().T))
    if which == "z":
        return Qobj(np.diag(m).astype(complex))
    raise ValueError(which)


def basis(n, k=0):
    v = np.zeros((int(n), 1), dtype=complex)
    v[int(k), 0] = 1.0
    return Qobj(v)


def tensor(*args):
    ops = list(args[0]) if len(args) == 1 and isinstance(args[0], (list, tuple)) else list(args)
    out = ops[0].data
    for o in ops[1:]:
        out = sp.kron(out, o.data, format="csr")
    return Qobj(out)


class Result:
    def __init__(self, times, expect, states, stats):
        self.times = times
        self.expect = expect
        self.states = states
        self.stats = stats


def sesolve(H, psi0, tlist, e_ops=None, options=None, **_):
    from oracle import solvers  # local import: the fake must stay importable on its own

    e_ops = list(e_ops or [])
    opts = dict(options or {})
    store = opts.get("store_states", None)
    keep_states = bool(store) or (store is None and not e_ops)
    traces, norms_raw, states, stats = solvers.evolve_csr(
        H.data,
        psi0.full().ravel(),
        np.asarray(tlist, dtype=float),
        [op.data for op in e_ops],
        backend=SESOLVE_BACKEND,
        atol=opts.get("atol", 1e-8),
        rtol=opts.get("rtol", 1e-6),
        nsteps=opts.get("nsteps", 2500),
        max_step=opts.get("max_step", 0.0),
        keep_states=keep_states,
    )
    res_states = [Qobj(s.reshape(-1, 1)) for s in states] if keep_states else []
    stats = dict(stats)
    stats["raw_norm"] = norms_raw
    return Result(np.asarray(tlist), [tr for tr in traces], res_states, stats)
