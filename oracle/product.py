"""Closed-form rows of H psi for PRODUCT states - the oracle for sizes no host vector reaches (N = 28 .. 33).

TEST INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py's parity fields).

For psi[x] = prod_b phi[b][x_b] every neighbour amplitude of the reference's CSR row product
(dipolar_ensemble_with_rare.py:504-570 assembled; matrix elements as in oracle/dense.py) is
psi[x] times a ratio of single-site factors:

    psi[x ^ m_b] = psi[x] * r_b(x),   r_b(x) = phi[b][1 - x_b] / phi[b][x_b]

so   (H psi)[x] / psi[x] = d(x) + sum_b (hx_b - i s_b hy_b)/2 r_b + sum_{b<c} (Jxx_bc - s_b s_c Jyy_bc)/4 r_b r_c
with d(x) = sum_b hz_b s_b/2 + sum_{b<c} Jzz_bc s_b s_c/4 and s_b = 1 - 2 x_b: O(n^2) per row, no vector at all.
"""
from __future__ import annotations

import numpy as np


def random_phi(n: int, seed: int) -> np.ndarray:
    """[n][2] complex single-site kets, normalised, both components well away from zero (the ratios stay O(1))."""
    rng = np.random.default_rng(seed)
    theta = rng.uniform(0.25 * np.pi, 0.75 * np.pi, size=n)     # polar angle: |phi_0|, |phi_1| in [0.38, 0.92]
    a, b = rng.uniform(0.0, 2.0 * np.pi, size=(2, n))
    return np.stack([np.cos(theta / 2) * np.exp(1j * a), np.sin(theta / 2) * np.exp(1j * b)], axis=1)


def amplitudes(phi: np.ndarray, rows: np.ndarray) -> np.ndarray:
    """psi[x] for the global indices in ``rows``."""
    rows = np.asarray(rows, dtype=np.uint64)
    out = np.ones(rows.shape, dtype=np.complex128)
    for b in range(phi.shape[0]):
        out *= phi[b][((rows >> np.uint64(b)) & np.uint64(1)).astype(np.int64)]
    return out


def full_state(phi: np.ndarray) -> np.ndarray:
    """The whole 2^n vector (small n only): bit b is factor b counted from the least significant end."""
    out = np.ones(1, dtype=np.complex128)
    for b in range(phi.shape[0]):
        out = np.kron(phi[b], out)
    return out


def apply_h_rows(h: np.ndarray, J: np.ndarray, phi: np.ndarray, rows: np.ndarray) -> np.ndarray:
    """(H psi)[rows] for the bit-major table (h[3][n], J[3][n][n]) and the product state phi; rows: global indices."""
    n = h.shape[1]
    rows = np.asarray(rows, dtype=np.uint64)
    bits = np.stack([((rows >> np.uint64(b)) & np.uint64(1)).astype(np.int64) for b in range(n)])     # (n, R)
    s = 1.0 - 2.0 * bits
    r = np.stack([phi[b][1 - bits[b]] / phi[b][bits[b]] for b in range(n)])                            # (n, R)
    Jx, Jy, Jz = (np.triu(J[a], 1) for a in range(3))
    acc = 0.5 * (h[2] @ s) + 0.25 * np.einsum("br,bc,cr->r", s, Jz, s)
    acc = acc + 0.5 * (h[0] @ r) - 0.5j * ((h[1][:, None] * s * r).sum(axis=0))
    sr = s * r
    acc = acc + 0.25 * np.einsum("br,bc,cr->r", r, Jx, r) - 0.25 * np.einsum("br,bc,cr->r", sr, Jy, sr)
    return acc * amplitudes(phi, rows)
