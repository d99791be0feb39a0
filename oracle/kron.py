"""K: literal Kronecker-product restatement of the reference's Hamiltonian/observable/initial-state build.

TEST INFRASTRUCTURE ONLY.  Follows DE:453-606 term by term on scipy.sparse matrices: ``build`` with all local
dimensions 2 (the only case the reference can run as written), ``build_spin32`` with the J = 3/2 rare
spin on a four-level last site (the reference with DE:486 repaired; pinned by tests/golden/spin32_cases.*).  It takes a *duck-typed* params
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


# ---- local dimension 4: the J = 3/2 rare spin (DE:22-25, DE:499-501) ------------------------------------------------
def _jmat32():
    """Spin-3/2 matrices in the basis m = +3/2, +1/2, -1/2, -3/2 (what ``qt.jmat(1.5, ...)`` returns, DE:22-24)."""
    m = 1.5 - np.arange(4)
    jp = np.zeros((4, 4), dtype=complex)                 # J+ |m> = sqrt(j(j+1) - m(m+1)) |m+1>
    for k in range(1, 4):
        jp[k - 1, k] = np.sqrt(1.5 * 2.5 - m[k] * (m[k] + 1.0))
    jx = 0.5 * (jp + jp.conj().T)
    jy = -0.5j * (jp - jp.conj().T)
    return sp.csr_matrix(jx), sp.csr_matrix(jy), sp.csr_matrix(np.diag(m).astype(complex))


def build_spin32(params, freqs_fn: Callable, positions_fn: Callable, couplings_fn: Callable
                 ) -> Tuple[sp.csr_matrix, Dict[str, sp.csr_matrix], np.ndarray]:
    """(H, observables, psi0) of DE:483-606 for ``is_spin_three_half=True`` with the reference's dims repaired:
    local dimensions [2] * n_sea + [4] (DE:28-34) instead of DE:486's all-2, rare operators the 4 x 4 J matrices
    (DE:499-501).  LAB frame, literal Kronecker products - independent of the product's two-bit encoding, rotated frame
    and Pauli strings.  Needs ``is_center_rare`` (the sea-centre control has no spin-3/2 site)."""
    if not params.is_center_rare:
        raise ValueError("spin-3/2 needs is_center_rare=True")
    n_sea = params.n_sea
    dim = (1 << n_sea) * 4
    jx, jy, jz = _jmat32()

    def sea_op(local, k):
        left = sp.identity(1 << k, dtype=complex, format="csr")
        right = sp.identity((1 << (n_sea - 1 - k)) * 4, dtype=complex, format="csr")
        return sp.kron(sp.kron(left, local), right, format="csr")

    def rare_op(local):
        return sp.kron(sp.identity(1 << n_sea, dtype=complex, format="csr"), local, format="csr")

    def sea_sum(local):
        acc = sp.csr_matrix((dim, dim), dtype=complex)
        for k in range(n_sea):
            acc = acc + sea_op(local, k)
        return acc

    fr = freqs_fn(params)
    H = sp.csr_matrix((dim, dim), dtype=complex)
    if params.drive_sea and fr["delta_sea"] != 0.0:                                   # DE:505-508
        H = H + fr["delta_sea"] * sea_sum(_SZ)
    if params.drive_rare and fr["delta_rare"] != 0.0:                                 # DE:510-512
        H = H + fr["delta_rare"] * rare_op(jz)
    if params.drive_sea and fr["omega1_sea"] != 0.0:                                  # DE:515-519
        H = H + fr["omega1_sea"] * (np.cos(params.phi_sea) * sea_sum(_SX) + np.sin(params.phi_sea) * sea_sum(_SY))
    if params.drive_rare and fr["omega1_rare"] != 0.0:                                # DE:523-528
        H = H + fr["omega1_rare"] * (np.cos(params.phi_rare) * rare_op(jx) + np.sin(params.phi_rare) * rare_op(jy))
    pos = positions_fn(n_sea, params.shell_scale)                                     # DE:533-536
    b = couplings_fn(pos, params.dipolar_scale, params.gamma_sea, params.gamma_rare)  # DE:540-545
    for i, j in combinations(range(n_sea + 1), 2):                                    # DE:549-568
        if j < n_sea:
            zz = sea_op(_SZ, i) @ sea_op(_SZ, j)
            xx = sea_op(_SX, i) @ sea_op(_SX, j)
            yy = sea_op(_SY, i) @ sea_op(_SY, j)
            H = H + b[i, j] * (zz - 0.25 * (xx - yy))
        else:
            H = H + b[i, j] * (sea_op(_SZ, i) @ rare_op(jz))
    obs = {                                                                           # DE:573-587
        "Ix_sea": sea_sum(_SX), "Iy_sea": sea_sum(_SY), "Iz_sea": sea_sum(_SZ),
        "Iz_R": rare_op(jz), "Ix_R": rare_op(jx), "Iy_R": rare_op(jy),
    }
    sea_bit = 0 if params.init_x_sign >= 0 else 1                                     # DE:71, DE:599
    level = 0 if (-params.init_x_sign) >= 0 else 3                                    # DE:99, DE:602: argmax / argmin of Jz
    idx = 0
    for _ in range(n_sea):
        idx = 2 * idx + sea_bit
    idx = 4 * idx + level
    psi0 = np.zeros(dim, dtype=complex)
    psi0[idx] = 1.0
    return H.tocsr(), obs, psi0
