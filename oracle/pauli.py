"""Weighted Pauli-string expectation values on host vectors.  TEST INFRASTRUCTURE ONLY.

numpy restatement of what ``qsb_propagate_expect`` / ``qsb_session_expect`` (include/qsb200.h) define: for terms
(x_mask, z_mask, weight, column), P = i^popcount(x_mask & z_mask) X^x_mask Z^z_mask (a bit in both masks carries
Y = i X Z), column c of the result is sum_t weight_t <psi|P_t|psi> / <psi|psi>, columns 6 / 7 the norm after / before
that normalisation.  The reference evaluates the same numbers as ``qt.expect`` of Kronecker-built operators on the
output-normalised ket (DE:573-587, DE:653-666).
"""
from __future__ import annotations

from typing import Sequence, Tuple

import numpy as np


def _parity(x: np.ndarray, mask: int) -> np.ndarray:
    par = np.zeros_like(x)
    b = 0
    while mask >> b:
        if (mask >> b) & 1:
            par ^= (x >> b) & 1
        b += 1
    return par


def expect_terms(states: np.ndarray, terms: Sequence[Tuple[int, int, float, int]]) -> np.ndarray:
    """states: (dim,) or (dim, nt) complex -> (8, nt) columns as the C ABI orders them."""
    st = np.asarray(states, dtype=complex)
    if st.ndim == 1:
        st = st[:, None]
    dim, nt = st.shape
    x = np.arange(dim, dtype=np.int64)
    out = np.zeros((8, nt))
    n2 = np.sum(np.abs(st) ** 2, axis=0)
    for xm, zm, w, col in terms:
        ny = bin(xm & zm).count("1")
        sign = 1.0 - 2.0 * _parity(x, zm)
        val = (1j ** ny) * np.sum(st[x ^ xm].conj() * (sign[:, None] * st), axis=0)
        out[col] += w * val.real
    out[:6] /= n2
    out[6] = 1.0
    out[7] = np.sqrt(n2)
    return out
