"""K: literal Kronecker-product restatement of the reference's Hamiltonian/observable/initial-state build.

TEST INFRASTRUCTURE ONLY.  Follows DE:453-606 term by term on scipy.sparse matrices (all local
dimensions 2 - the only case the reference can run).  It takes a *duck-typed* params
object (anything with the DipolarRareParams attribute names) plus the geometry/frequency callables,
so it can be driven either by the reference's own helpers (in the build container) or by the
product's mirror of them (anywhere) without importing either.
"""
from __future__ import annotations

from itertools import combinations
from typing import Callable, Dict, Tuple

import numpy as np
import scipy.sparse as sp

_SX = 0.5 * sp.csr_matrix(np.array([[0, 1], [1, 0]], dtype=complex))
_SY = 0.5 * sp.csr_matrix(np.array([[0, -1j], [1j, 0]], dtype=complex))
_SZ = 0.5 * sp.csr_matrix(np.array([[1, 0], [0, -1]], dtype=complex))
_ID = sp.identity(2, dtype=complex, format="csr")


def site_op(local: sp.spmatrix, site: int, n: int) -> sp.csr_matrix:
    """Embed a one-site operator; site 0 is the MOST significant tensor factor (DE:37-45)."""
    out = None
    for k in range(n):
        f = local if k == site else _ID
        out = f if out is None else sp.kron(out, f, format="csr")
    return out.tocsr()


def build(params, freqs_fn: Callable, positions_fn: Callable, couplings_fn: Callable
          ) -> Tuple[sp.csr_matrix, Dict[str, sp.csr_matrix], np.ndarray]:
    """(H, observables, psi0) exactly as DE:483-606 would assemble them."""
    n_sea = params.n_sea
    n_total = n_sea + 1
    rare = n_sea
    if not params.is_center_rare:           # DE:488-489: the centre becomes one more sea spin
        n_sea = n_total
    fr = freqs_fn(params)
    dim = 1 << n_total
    H = sp.csr_matrix((dim, dim), dtype=complex)

    def sea_sum(local):
        acc = sp.csr_matrix((dim, dim), dtype=complex)
        for k in range(n_sea):
            acc = acc + site_op(local, k, n_total)
        return acc

    if params.drive_sea and fr["delta_sea"] != 0.0:                                   # DE:505-508
        H = H + fr["delta_sea"] * sea_sum(_SZ)
    if params.is_center_rare and params.drive_rare and fr["delta_rare"] != 0.0:       # DE:510-512
        H = H + fr["delta_rare"] * site_op(_SZ, rare, n_total)
    if params.drive_sea and fr["omega1_sea"] != 0.0:                                  # DE:515-519
        H = H + fr["omega1_sea"] * (np.cos(params.phi_sea) * sea_sum(_SX) + np.sin(params.phi_sea) * sea_sum(_SY))
    if params.is_center_rare and params.drive_rare and fr["omega1_rare"] != 0.0:      # DE:523-528
        H = H + fr["omega1_rare"] * (
            np.cos(params.phi_rare) * site_op(_SX, rare, n_total)
            + np.sin(params.phi_rare) * site_op(_SY, rare, n_total)
        )

    pos = positions_fn(params.n_sea, params.shell_scale)                              # DE:533-536
    b = couplings_fn(pos, params.dipolar_scale, params.gamma_sea,
                     params.gamma_rare if params.is_center_rare else params.gamma_sea)  # DE:540-545
    for i, j in combinations(range(n_total), 2):                                      # DE:549-568
        zz = site_op(_SZ, i, n_total) @ site_op(_SZ, j, n_total)
        if i < n_sea and j < n_sea:
            xx = site_op(_SX, i, n_total) @ site_op(_SX, j, n_total)
            yy = site_op(_SY, i, n_total) @ site_op(_SY, j, n_total)
            H = H + b[i, j] * (zz - 0.25 * (xx - yy))
        elif i == rare or j == rare:
            H = H + b[i, j] * zz

    obs = {                                                                           # DE:573-587
        "Ix_sea": sea_sum(_SX),
        "Iy_sea": sea_sum(_SY),
        "Iz_sea": sea_sum(_SZ),
        "Iz_R": site_op(_SZ, rare, n_total),
        "Ix_R": site_op(_SX, rare, n_total),
        "Iy_R": site_op(_SY, rare, n_total),
    }

    sea_bit = 0 if params.init_x_sign >= 0 else 1                                     # DE:71, DE:599
    bits = [sea_bit] * n_total
    if params.is_center_rare:                                                         # DE:602, DE:99
        bits[-1] = 0 if (-params.init_x_sign) >= 0 else 1
    idx = 0
    for q in bits:
        idx = 2 * idx + q
    psi0 = np.zeros(dim, dtype=complex)
    psi0[idx] = 1.0
    return H.tocsr(), obs, psi0
