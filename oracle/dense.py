"""Semantics of the bit-major XYZ coupling table, restated with plain numpy index arithmetic.

TEST INFRASTRUCTURE ONLY.  Input is exactly what the C ABI receives (include/qsb200.h):
``h[3][n]`` = (hx, hy, hz) per BIT and ``J[3][n][n]`` = symmetric (Jxx, Jyy, Jzz) per bit pair, for

    H = sum_b h[a][b] I^a_b + sum_{b<c} J[a][b][c] I^a_b I^a_c,     I = sigma/2,  bit value 0 <-> Iz=+1/2.

Matrix elements (derived from DE:16-18 Pauli conventions):
  diagonal      sum_b hz_b s_b/2 + sum_{b<c} Jzz_bc s_b s_c/4,                    s_b = 1-2*bit_b(x)
  single flip   out[x] += (hx_b - i s_b(x) hy_b)/2 * psi[x ^ m_b]
  double flip   out[x] += (Jxx_bc - s_b(x) s_c(x) Jyy_bc)/4 * psi[x ^ m_b ^ m_c]
"""
from __future__ import annotations

from typing import Dict, Tuple

import numpy as np
import scipy.sparse as sp


def _signs(n: int) -> np.ndarray:
    """(n, 2^n) array of s_b(x) = +1 / -1."""
    x = np.arange(1 << n, dtype=np.int64)
    return np.stack([1.0 - 2.0 * ((x >> b) & 1) for b in range(n)])


def diagonal(h: np.ndarray, J: np.ndarray) -> np.ndarray:
    n = h.shape[1]
    s = _signs(n)
    d = 0.5 * (h[2] @ s)
    for b in range(n):
        for c in range(b + 1, n):
            if J[2, b, c] != 0.0:
                d += 0.25 * J[2, b, c] * s[b] * s[c]
    return d


def apply_h(h: np.ndarray, J: np.ndarray, psi: np.ndarray) -> np.ndarray:
    """out = H psi, matrix-free, O(n^2 2^n); fine up to n ~ 20."""
    n = h.shape[1]
    x = np.arange(1 << n, dtype=np.int64)
    s = _signs(n)
    out = diagonal(h, J) * psi
    for b in range(n):
        if h[0, b] != 0.0 or h[1, b] != 0.0:
            out += 0.5 * (h[0, b] - 1j * s[b] * h[1, b]) * psi[x ^ (1 << b)]
    for b in range(n):
        for c in range(b + 1, n):
            if J[0, b, c] != 0.0 or J[1, b, c] != 0.0:
                out += 0.25 * (J[0, b, c] - s[b] * s[c] * J[1, b, c]) * psi[x ^ (1 << b) ^ (1 << c)]
    return out


def csr_from_table(h: np.ndarray, J: np.ndarray) -> sp.csr_matrix:
    n = h.shape[1]
    dim = 1 << n
    x = np.arange(dim, dtype=np.int64)
    s = _signs(n)
    rows, cols, vals = [x], [x], [diagonal(h, J).astype(complex)]
    for b in range(n):
        if h[0, b] != 0.0 or h[1, b] != 0.0:
            rows.append(x)
            cols.append(x ^ (1 << b))
            vals.append(0.5 * (h[0, b] - 1j * s[b] * h[1, b]))
    for b in range(n):
        for c in range(b + 1, n):
            if J[0, b, c] != 0.0 or J[1, b, c] != 0.0:
                rows.append(x)
                cols.append(x ^ (1 << b) ^ (1 << c))
                vals.append((0.25 * (J[0, b, c] - s[b] * s[c] * J[1, b, c])).astype(complex))
    M = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(dim, dim))
    M = M.tocsr()
    M.eliminate_zeros()
    return M


def observable_sums(psi: np.ndarray, sea_mask: int, rare_mask: int) -> Dict[str, float]:
    """The 7 normalised outputs of DE:671-679 plus the raw norm, from index arithmetic (appendix A.4)."""
    n = int(np.log2(psi.size))
    x = np.arange(psi.size, dtype=np.int64)
    p = np.abs(psi) ** 2
    nrm2 = float(p.sum())

    def site(b: int) -> Tuple[complex, float]:
        m = 1 << b
        lo = x[(x & m) == 0]
        c = np.sum(psi[lo].conj() * psi[lo | m])          # <I+> ; Ix = Re, Iy = Im
        z = 0.5 * float(np.sum((1.0 - 2.0 * ((x >> b) & 1)) * p))
        return c, z

    cs, zs = 0.0 + 0.0j, 0.0
    for b in range(n):
        if sea_mask >> b & 1:
            c, z = site(b)
            cs += c
            zs += z
    rb = int(rare_mask).bit_length() - 1
    cr, zr = site(rb)
    return {
        "Ix_sea": cs.real / nrm2, "Iy_sea": cs.imag / nrm2, "Iz_sea": zs / nrm2,
        "Iz_R": zr / nrm2, "Ix_R": cr.real / nrm2, "Iy_R": cr.imag / nrm2,
        "state_norm": 1.0, "raw_norm": float(np.sqrt(nrm2)),
    }
