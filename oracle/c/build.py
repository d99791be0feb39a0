"""gcc build + ctypes binding of the C restatement (oracle/c/qsb_oracle.c).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "qsb_oracle.c")
LIB = os.path.join(HERE, "libqsb_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        subprocess.run(["gcc", "-O3", "-march=x86-64-v2", "-fopenmp", "-shared", "-fPIC", "-o", LIB, SRC, "-lm"], check=True)
    return LIB


def load():
    global _lib
    if _lib is None:
        build()
        lib = C.CDLL(LIB)
        dp, vp = C.POINTER(C.c_double), C.c_void_p
        lib.oc_max_threads.restype = C.c_int
        lib.oc_set_threads.argtypes = [C.c_int]
        lib.oc_diagonal.argtypes = [C.c_int, dp, dp, C.c_uint64, C.c_uint64, dp]
        lib.oc_apply_h_rows.argtypes = [C.c_int, dp, dp, dp, vp, vp, C.c_uint64, C.c_uint64]
        _lib = lib
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def max_threads() -> int:
    return load().oc_max_threads()


def set_threads(n: int) -> int:
    """Force the OpenMP team size (overrides an inherited OMP_NUM_THREADS); returns the resulting maximum."""
    load().oc_set_threads(int(n))
    return max_threads()


def diagonal(h: np.ndarray, J: np.ndarray, row0: int = 0, row1: int | None = None) -> np.ndarray:
    n = h.shape[1]
    row1 = (1 << n) if row1 is None else row1
    out = np.empty(row1 - row0, dtype=np.float64)
    load().oc_diagonal(n, _dp(np.ascontiguousarray(h)), _dp(np.ascontiguousarray(J)), row0, row1, _dp(out))
    return out


def apply_h(h: np.ndarray, J: np.ndarray, psi: np.ndarray, row0: int = 0, row1: int | None = None,
            diag: np.ndarray | None = None) -> np.ndarray:
    """(H psi)[row0:row1] - one CSR-matvec's worth of gather work on those rows."""
    n = h.shape[1]
    row1 = (1 << n) if row1 is None else row1
    h = np.ascontiguousarray(h, dtype=np.float64)
    J = np.ascontiguousarray(J, dtype=np.float64)
    psi = np.ascontiguousarray(psi, dtype=np.complex128)
    if diag is None:
        diag = diagonal(h, J, row0, row1)
    out = np.empty(row1 - row0, dtype=np.complex128)
    load().oc_apply_h_rows(n, _dp(h), _dp(J), _dp(diag), psi.ctypes.data, out.ctypes.data, row0, row1)
    return out
