"""ctypes binding of libqsb200.so (include/qsb200.h) and the Python-level hot-path calls built on it.

There is no CPU fallback anywhere in this module: if the shared library is missing, or no CUDA device
is present, the calls raise.  torch is used only as a device-buffer allocator (``use_torch_workspace``).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np

from .table import SpinTable

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libqsb200.so")

NOBS = 8
OBS_COLUMNS = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R", "state_norm", "raw_norm")
SMALL_MAX_QUBITS = 10
HBM_PROPAGATOR_MAX_QUBITS = 14
METHOD_DP5, METHOD_TAYLOR = 0, 1
_METHODS = {"dp5": METHOD_DP5, "taylor": METHOD_TAYLOR}

# QuTiP 5 defaults that apply when the reference passes options=None
DEFAULT_ATOL, DEFAULT_RTOL, DEFAULT_NSTEPS = 1e-8, 1e-6, 2500


class EngineError(RuntimeError):
    """Any non-zero qsb_status."""

    def __init__(self, code: int, msg: str):
        super().__init__(f"libqsb200 status {code}: {msg}")
        self.code = code


class IntegratorError(EngineError):
    """nsteps exhausted inside one output interval / step underflow (QuTiP raises IntegratorException here)."""


class _Table(C.Structure):
    _fields_ = [("n_qubits", C.c_int32), ("h", C.POINTER(C.c_double)), ("J", C.POINTER(C.c_double)),
                ("sea_mask", C.c_uint64), ("rare_mask", C.c_uint64)]


class _Opts(C.Structure):
    _fields_ = [("atol", C.c_double), ("rtol", C.c_double), ("max_step", C.c_double), ("nsteps", C.c_int64),
                ("method", C.c_int32), ("order", C.c_int32), ("first_step", C.c_double)]


class _Stats(C.Structure):
    _fields_ = [("n_accepted", C.c_int64), ("n_rejected", C.c_int64), ("n_rhs", C.c_int64), ("last_h", C.c_double),
                ("device_ms", C.c_double), ("path", C.c_int32), ("status", C.c_int32)]


class _PauliTerm(C.Structure):
    _fields_ = [("x_mask", C.c_uint64), ("z_mask", C.c_uint64), ("weight", C.c_double), ("column", C.c_int32),
                ("reserved", C.c_int32)]


MAX_PAULI_TERMS = 512

EXPORTS = (
    "qsb_version", "qsb_last_error", "qsb_device_count", "qsb_plan_create", "qsb_plan_destroy",
    "qsb_plan_workspace_bytes", "qsb_plan_set_workspace", "qsb_evolve", "qsb_evolve_batch", "qsb_apply_h",
    "qsb_apply_h_product", "qsb_apply_h_static",
    "qsb_profile", "qsb_profile_read", "qsb_debug_fp64_peak", "qsb_session_begin", "qsb_session_step", "qsb_session_advance", "qsb_session_observe", "qsb_session_state", "qsb_session_end",
    "qsb_shard_plan_create", "qsb_shard_export", "qsb_shard_connect", "qsb_shard_allgather", "qsb_shard_set_stream_barrier", "qsb_plan_local_dim", "qsb_debug_layouts", "qsb_debug_fusion",
    "qsb_local_group_create", "qsb_local_group_destroy", "qsb_local_group_member", "qsb_local_allgather", "qsb_local_stream_barrier",
    "qsb_propagate_batch", "qsb_propagate_hbm", "qsb_release_caches", "qsb_build_propagator", "qsb_debug_propagate_shape", "qsb_debug_squarings",
    "qsb_propagate_expect", "qsb_propagate_expect_batch", "qsb_propagate_hbm_expect", "qsb_session_expect",
)

SHARD_BLOB_BYTES = 96
# int (*qsb_allgather_fn)(void* ctx, const double* send, int32_t count, double* recv)
ALLGATHER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_double), C.c_int32, C.POINTER(C.c_double))
# int (*qsb_stream_barrier_fn)(void* ctx, void* cuda_stream)
STREAM_BARRIER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p)

_lib = None


def load_library() -> C.CDLL:
    """dlopen libqsb200.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing - build it with `python -m quantumsimulations_b200.build` "
            "(there is no CPU fallback for the time-evolution path)"
        )
    lib = C.CDLL(LIB_PATH)
    dp, vp = C.POINTER(C.c_double), C.c_void_p
    lib.qsb_version.restype = C.c_int
    lib.qsb_last_error.restype = C.c_char_p
    lib.qsb_device_count.argtypes = [C.POINTER(C.c_int)]
    lib.qsb_plan_create.argtypes = [C.POINTER(_Table), C.c_int, C.POINTER(vp)]
    lib.qsb_plan_destroy.argtypes = [vp]
    lib.qsb_plan_workspace_bytes.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_size_t)]
    lib.qsb_plan_set_workspace.argtypes = [vp, vp, C.c_size_t]
    lib.qsb_evolve.argtypes = [vp, C.c_uint64, dp, C.c_int64, dp, C.POINTER(_Opts), dp, C.POINTER(_Stats)]
    lib.qsb_evolve_batch.argtypes = [C.c_int32, C.POINTER(_Table), C.POINTER(C.c_uint64), C.c_int64, dp, C.c_int32,
                                     C.POINTER(_Opts), dp, C.POINTER(_Stats), C.c_int]
    lib.qsb_propagate_batch.argtypes = [C.c_int32, C.POINTER(_Table), C.POINTER(C.c_uint64), C.c_int64, dp, C.c_int32,
                                        C.POINTER(_Opts), C.c_int32, C.c_int32, dp, C.POINTER(_Stats), C.c_int]
    lib.qsb_propagate_hbm.argtypes = [C.POINTER(_Table), C.c_uint64, C.c_int64, C.c_double, C.POINTER(_Opts), C.c_int32,
                                      C.c_int32, dp, C.POINTER(_Stats), C.c_int]
    lib.qsb_build_propagator.argtypes = [C.POINTER(_Table), C.c_double, C.POINTER(_Opts), C.c_int32, dp,
                                         C.POINTER(_Stats), C.c_int]
    lib.qsb_debug_squarings.argtypes = [C.POINTER(_Table), C.c_double, C.POINTER(C.c_int32)]
    lib.qsb_propagate_expect.argtypes = [C.POINTER(_Table), C.c_uint64, C.c_int64, C.c_double, C.POINTER(_Opts), C.c_int32,
                                         C.POINTER(_PauliTerm), C.c_int32, C.c_int32, dp, C.POINTER(_Stats), C.c_int]
    lib.qsb_propagate_hbm_expect.argtypes = lib.qsb_propagate_expect.argtypes
    lib.qsb_propagate_expect_batch.argtypes = [C.c_int32, C.POINTER(_Table), C.POINTER(C.c_uint64), C.c_int64, dp, C.c_int32,
                                               C.POINTER(_Opts), C.c_int32, C.POINTER(_PauliTerm), C.c_int32, C.c_int32, dp,
                                               C.POINTER(_Stats), C.c_int]
    lib.qsb_session_expect.argtypes = [vp, C.POINTER(_PauliTerm), C.c_int32, dp]
    lib.qsb_debug_propagate_shape.argtypes = [C.c_int32, C.POINTER(C.c_int32)]
    lib.qsb_apply_h.argtypes = [vp, dp, dp]
    lib.qsb_apply_h_static.argtypes = [vp, dp, dp]
    lib.qsb_apply_h_product.argtypes = [vp, dp, C.POINTER(C.c_uint64), C.c_int64, dp]
    lib.qsb_session_begin.argtypes = [vp, C.c_uint64, dp, C.c_double, C.POINTER(_Opts)]
    lib.qsb_session_step.argtypes = [vp, C.c_int64, dp, C.POINTER(_Stats), dp]
    lib.qsb_session_advance.argtypes = [vp, C.c_double, dp, C.POINTER(_Stats)]
    lib.qsb_session_observe.argtypes = [vp, dp]
    lib.qsb_session_state.argtypes = [vp, dp]
    lib.qsb_session_end.argtypes = [vp]
    lib.qsb_shard_plan_create.argtypes = [C.POINTER(_Table), C.c_int, C.c_int32, C.c_int32, C.POINTER(vp)]
    lib.qsb_shard_export.argtypes = [vp, vp]
    lib.qsb_shard_connect.argtypes = [vp, vp, vp, vp]
    lib.qsb_shard_allgather.argtypes = [vp, dp, C.c_int32, dp]
    lib.qsb_shard_set_stream_barrier.argtypes = [vp, vp, vp]
    lib.qsb_plan_local_dim.argtypes = [vp, C.POINTER(C.c_uint64)]
    lib.qsb_debug_layouts.argtypes = [C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.qsb_local_group_create.argtypes = [C.c_int32, C.POINTER(vp)]
    lib.qsb_local_group_destroy.argtypes = [vp]
    lib.qsb_local_group_member.argtypes = [vp, C.c_int32]
    lib.qsb_local_group_member.restype = vp
    lib.qsb_local_allgather.argtypes = [vp, dp, C.c_int32, dp]
    lib.qsb_profile.argtypes = [C.c_int]
    lib.qsb_profile_read.argtypes = [dp, C.POINTER(C.c_int64), dp, C.c_int32]
    _lib = lib
    return lib


def _check(code: int) -> None:
    if code == 0:
        return
    msg = load_library().qsb_last_error().decode("utf-8", "replace")
    if code in (2, 3):
        raise IntegratorError(code, msg)
    if code == 1:
        raise ValueError(f"libqsb200: {msg}")
    raise EngineError(code, msg)


def profile(enable: bool) -> None:
    _check(load_library().qsb_profile(1 if enable else 0))


PROFILE_CLASSES = ("pass_first", "pass_mid", "pass_centre", "pass_last", "pass_first_comb", "error_sums",
                   "strided_fused", "pass_last_first")


def profile_read(with_bytes: bool = False):
    """({class: summed ms}, {class: launches}[, {class: bytes moved by design}]) of the tiled-path kernels since
    profile(True)."""
    n = len(PROFILE_CLASSES)
    ms = (C.c_double * n)()
    cnt = (C.c_int64 * n)()
    byt = (C.c_double * n)()
    _check(load_library().qsb_profile_read(ms, cnt, byt, n))
    out = dict(zip(PROFILE_CLASSES, list(ms))), dict(zip(PROFILE_CLASSES, list(cnt)))
    return out + (dict(zip(PROFILE_CLASSES, list(byt))),) if with_bytes else out


def tile_layouts(n_qubits: int, world: int = 1):
    """[(lk, gs, ge)] of the tiled path's layouts (host-only query; see qsb_debug_layouts)."""
    n = C.c_int32(0)
    out = (C.c_int32 * 18)()
    _check(load_library().qsb_debug_layouts(int(n_qubits), int(world), C.byref(n), out))
    return [(out[3 * k], out[3 * k + 1], out[3 * k + 2]) for k in range(n.value)]


def strided_fusion(n_qubits: int, n_sms: int = 148):
    """How the strided passes of an unsharded n-qubit plan are fused (host-only query; see qsb_debug_fusion).
    None when every pass is its own launch."""
    out = (C.c_int64 * 13)()
    lib = load_library()
    lib.qsb_debug_fusion.argtypes = [C.c_int32, C.c_int32, C.POINTER(C.c_int64)]
    _check(lib.qsb_debug_fusion(int(n_qubits), int(n_sms), out))
    if not out[0]:
        return None
    return {"f": out[1], "m": out[2], "fbits": out[3], "tiles_lb": out[4], "lag": out[5], "vmask": out[6],
            "fixed_vmask": [out[7 + k] for k in range(6)]}


def fp64_peak_tflops(device: int = 0) -> float:
    """Measured FP64 FMA rate of the device (qsb_debug_fp64_peak)."""
    lib = load_library()
    lib.qsb_debug_fp64_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
    v = C.c_double(0.0)
    _check(lib.qsb_debug_fp64_peak(int(device), C.byref(v)))
    return v.value


def release_caches(device: int = -1) -> None:
    """Free the device buffers the propagator path keeps between calls (qsb_release_caches)."""
    lib = load_library()
    lib.qsb_release_caches.argtypes = [C.c_int]
    _check(lib.qsb_release_caches(int(device)))


def device_count() -> int:
    n = C.c_int(0)
    _check(load_library().qsb_device_count(C.byref(n)))
    return n.value


def _dptr(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_double))


@dataclass
class SolverOptions:
    atol: float = DEFAULT_ATOL
    rtol: float = DEFAULT_RTOL
    max_step: float = 0.0
    nsteps: int = DEFAULT_NSTEPS
    method: str = "dp5"
    order: int = 0
    first_step: float = 0.0

    def to_c(self) -> _Opts:
        return _Opts(float(self.atol), float(self.rtol), float(self.max_step or 0.0), int(self.nsteps or 0),
                     _METHODS[self.method], int(self.order), float(self.first_step or 0.0))


@dataclass
class Stats:
    n_accepted: int
    n_rejected: int
    n_rhs: int
    last_h: float
    device_ms: float
    path: int
    status: int

    @classmethod
    def from_c(cls, s: _Stats) -> "Stats":
        return cls(s.n_accepted, s.n_rejected, s.n_rhs, s.last_h, s.device_ms, s.path, s.status)


class _TableHolder:
    """Keeps the numpy arrays behind a C qsb_table alive."""

    def __init__(self, table: SpinTable):
        bm = table.bit_major()
        self.h = np.ascontiguousarray(bm["h"], dtype=np.float64)
        self.J = np.ascontiguousarray(bm["J"], dtype=np.float64)
        self.c = _Table(table.n, _dptr(self.h), _dptr(self.J), table.sea_mask, table.rare_mask)


class Plan:
    """One compiled coupling table on one device (qsb_plan)."""

    def __init__(self, table: SpinTable, device: int = 0):
        self.lib = load_library()
        self.table = table
        self._holder = _TableHolder(table)
        self._h = C.c_void_p()
        self._ws = None
        self.local_dim = table.dim
        self._create(int(device))

    def _create(self, device: int) -> None:
        _check(self.lib.qsb_plan_create(C.byref(self._holder.c), device, C.byref(self._h)))

    def close(self) -> None:
        if self._h:
            self.lib.qsb_plan_destroy(self._h)
            self._h = C.c_void_p()
            self._ws = None

    def __del__(self):
        try:                       # interpreter teardown: the library object may already be gone
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def workspace_bytes(self, method: str = "dp5", order: int = 0) -> int:
        n = C.c_size_t(0)
        _check(self.lib.qsb_plan_workspace_bytes(self._h, _METHODS[method], order, C.byref(n)))
        return n.value

    def use_torch_workspace(self, method: str = "dp5", device: Optional[int] = None) -> None:
        """Allocate the HBM workspace through torch's caching allocator instead of cudaMalloc."""
        import torch

        need = self.workspace_bytes(method)
        if need == 0:
            return
        dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        self._ws = torch.empty(need + 256, dtype=torch.uint8, device=dev)
        ptr = (self._ws.data_ptr() + 255) & ~255
        _check(self.lib.qsb_plan_set_workspace(self._h, C.c_void_p(ptr), need))

    # ---- the hot path ----------------------------------------------------------------------------
    def evolve(self, t_out: np.ndarray, opts: SolverOptions, psi0: Optional[np.ndarray] = None):
        t_out = np.ascontiguousarray(t_out, dtype=np.float64)
        obs = np.empty((len(t_out), NOBS), dtype=np.float64)
        st = _Stats()
        copts = opts.to_c()
        psi_ptr = None
        if psi0 is not None:
            psi0 = np.ascontiguousarray(psi0, dtype=np.complex128)
            if psi0.shape != (self.local_dim,):
                raise ValueError("psi0 must have 2^n amplitudes (the rank's own slice for a sharded plan)")
            psi_ptr = psi0.view(np.float64).ctypes.data_as(C.POINTER(C.c_double))
        _check(self.lib.qsb_evolve(self._h, self.table.psi0_index, psi_ptr, len(t_out), _dptr(t_out),
                                   C.byref(copts), _dptr(obs), C.byref(st)))
        return obs, Stats.from_c(st)

    def apply_h(self, psi: np.ndarray) -> np.ndarray:
        psi = np.ascontiguousarray(psi, dtype=np.complex128)
        if psi.shape != (self.local_dim,):
            raise ValueError("psi must have 2^n amplitudes (the rank's own slice for a sharded plan)")
        out = np.empty_like(psi)
        _check(self.lib.qsb_apply_h(self._h, _dptr(psi.view(np.float64)), _dptr(out.view(np.float64))))
        return out

    def apply_h_static(self, psi: np.ndarray) -> np.ndarray:
        """``apply_h`` through qsb_apply_h_static: no device allocation (scratch in a static device array of
        36 KB, n_qubits <= 10); calls that overlap on one GPU race on that array."""
        psi = np.ascontiguousarray(psi, dtype=np.complex128)
        if psi.shape != (self.local_dim,):
            raise ValueError("psi must have 2^n amplitudes")
        out = np.empty_like(psi)
        _check(self.lib.qsb_apply_h_static(self._h, _dptr(psi.view(np.float64)), _dptr(out.view(np.float64))))
        return out

    def apply_h_product(self, phi: np.ndarray, rows: np.ndarray) -> np.ndarray:
        """Rows ``rows`` (indices into this plan's own slice) of H psi for the product state psi[x] = prod_b
        phi[b][bit_b(x)], formed and transformed on the device (qsb_apply_h_product; tiled path only).  ``phi`` is
        indexed by BIT, like everything in the C ABI."""
        phi = np.ascontiguousarray(phi, dtype=np.complex128)
        if phi.shape != (self.table.n, 2):
            raise ValueError("phi must be [n_qubits][2] complex")
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        out = np.empty(rows.shape[0], dtype=np.complex128)
        _check(self.lib.qsb_apply_h_product(self._h, _dptr(phi.view(np.float64)), rows.ctypes.data_as(C.POINTER(C.c_uint64)),
                                            rows.shape[0], _dptr(out.view(np.float64))))
        return out

    # ---- stepping session (tiled path) -------------------------------------------------------------
    def session_begin(self, opts: SolverOptions, t0: float = 0.0, psi0: Optional[np.ndarray] = None) -> None:
        copts = opts.to_c()
        ptr = None
        if psi0 is not None:
            psi0 = np.ascontiguousarray(psi0, dtype=np.complex128)
            if psi0.shape != (self.local_dim,):
                raise ValueError("psi0 must have 2^n amplitudes (the rank's own slice for a sharded plan)")
            ptr = _dptr(psi0.view(np.float64))
        _check(self.lib.qsb_session_begin(self._h, self.table.psi0_index, ptr, float(t0), C.byref(copts)))

    def session_step(self, n_attempts: int):
        t = C.c_double(0.0)
        ms = C.c_double(0.0)
        st = _Stats()
        _check(self.lib.qsb_session_step(self._h, int(n_attempts), C.byref(t), C.byref(st), C.byref(ms)))
        return t.value, Stats.from_c(st), ms.value

    def session_advance(self, t_target: float):
        """Integrate the open session to exactly ``t_target``; returns (t, Stats) with Stats.last_h = the controller's
        step size (what a resumed session passes as ``first_step``)."""
        t = C.c_double(0.0)
        st = _Stats()
        _check(self.lib.qsb_session_advance(self._h, float(t_target), C.byref(t), C.byref(st)))
        return t.value, Stats.from_c(st)

    def session_observe(self) -> np.ndarray:
        obs = np.empty(NOBS, dtype=np.float64)
        _check(self.lib.qsb_session_observe(self._h, _dptr(obs)))
        return obs

    def session_expect(self, terms) -> np.ndarray:
        """The NOBS columns of the session's current state with the six observables given as weighted Pauli strings
        ``(x_mask, z_mask, weight, column)`` (qsb_session_expect; what a local-dimension-4 table's trajectory reads
        instead of ``session_observe``)."""
        cterms = _pauli_terms(terms)
        obs = np.empty(NOBS, dtype=np.float64)
        _check(self.lib.qsb_session_expect(self._h, cterms, len(cterms), _dptr(obs)))
        return obs

    def session_state(self) -> np.ndarray:
        out = np.empty(self.local_dim, dtype=np.complex128)
        _check(self.lib.qsb_session_state(self._h, _dptr(out.view(np.float64))))
        return out

    def session_end(self) -> None:
        _check(self.lib.qsb_session_end(self._h))


class ShardedPlan(Plan):
    """One rank's share of a state vector sharded on its top log2(world) index bits (qsb_shard_plan_create).

    Every rank builds one from the SAME table, then ``export_blob()`` -> gather the blobs of all ranks ->
    ``connect(blobs, allgather, ctx)``; afterwards the ordinary Plan calls are collective: every rank makes the same
    calls, with state vectors holding only amplitudes [rank * local_dim, (rank + 1) * local_dim).
    See ``quantumsimulations_b200.shard`` for the torch.distributed and in-process (thread) connectors.
    """

    def __init__(self, table: SpinTable, rank: int, world: int, device: int = 0):
        self.rank, self.world = int(rank), int(world)
        self.device_index = int(device)
        self._hook = None
        self._stream_hook = None
        super().__init__(table, device)
        d = C.c_uint64(0)
        _check(self.lib.qsb_plan_local_dim(self._h, C.byref(d)))
        self.local_dim = int(d.value)

    def _create(self, device: int) -> None:
        _check(self.lib.qsb_shard_plan_create(C.byref(self._holder.c), device, self.rank, self.world, C.byref(self._h)))

    def export_blob(self) -> bytes:
        buf = C.create_string_buffer(SHARD_BLOB_BYTES)
        _check(self.lib.qsb_shard_export(self._h, buf))
        return buf.raw

    def connect(self, blobs: Sequence[bytes], allgather, ctx=None) -> None:
        """``allgather``: an ALLGATHER_FN instance or a raw function address (e.g. qsb_local_allgather)."""
        if len(blobs) != self.world or any(len(b) != SHARD_BLOB_BYTES for b in blobs):
            raise ValueError("need one blob of SHARD_BLOB_BYTES bytes per rank")
        joined = C.create_string_buffer(b"".join(blobs), SHARD_BLOB_BYTES * self.world)
        self._hook = allgather  # keep the ctypes callback alive as long as the plan
        fn = C.cast(allgather, C.c_void_p)
        _check(self.lib.qsb_shard_connect(self._h, joined, fn, C.c_void_p(ctx) if ctx is not None else None))

    def set_stream_barrier(self, hook, ctx=None) -> None:
        """Install (or, with None, remove) a stream-ordered rank barrier: a STREAM_BARRIER_FN instance.  With one, the
        H-apply overlaps the NVLink-bound centre pass of one branch with the HBM-bound passes of the other
        (qsb_shard_set_stream_barrier); only for ranks that each own their GPU."""
        self._stream_hook = hook   # keep the ctypes callback alive
        fn = C.cast(hook, C.c_void_p) if hook is not None else None
        _check(self.lib.qsb_shard_set_stream_barrier(self._h, fn, C.c_void_p(ctx) if ctx is not None else None))

    def allgather(self, values) -> np.ndarray:
        """[world][len(values)] doubles gathered over the plan's own hook (collective; host-side agreement only)."""
        send = np.ascontiguousarray(values, dtype=np.float64).ravel()
        recv = np.empty((self.world, send.size), dtype=np.float64)
        _check(self.lib.qsb_shard_allgather(self._h, _dptr(send), send.size, _dptr(recv)))
        return recv

    def local_slice(self) -> slice:
        return slice(self.rank * self.local_dim, (self.rank + 1) * self.local_dim)


def evolve_batch(tables: Sequence[SpinTable], t_out: np.ndarray, opts: Sequence[SolverOptions], device: int = 0):
    """Many n <= 10 trajectories in one launch (one CTA each).  t_out: (n_out,) shared or (n_traj, n_out)."""
    lib = load_library()
    n = len(tables)
    if n == 0:
        return np.empty((0, 0, NOBS)), []
    t_out = np.ascontiguousarray(t_out, dtype=np.float64)
    per_traj = 1 if t_out.ndim == 2 else 0
    n_out = t_out.shape[-1]
    holders = [_TableHolder(t) for t in tables]
    ctabs = (_Table * n)(*[h.c for h in holders])
    psi0 = (C.c_uint64 * n)(*[t.psi0_index for t in tables])
    copts = (_Opts * n)(*[o.to_c() for o in opts])
    cstats = (_Stats * n)()
    obs = np.empty((n, n_out, NOBS), dtype=np.float64)
    code = lib.qsb_evolve_batch(n, ctabs, psi0, n_out, _dptr(t_out), per_traj, copts, _dptr(obs), cstats, int(device))
    stats = [Stats.from_c(s) for s in cstats]
    _check(code)
    return obs, stats


def uniform_step(t_out: np.ndarray, rel: float = 1e-9) -> Optional[float]:
    """dt when ``t_out`` is a uniform grid (what np.linspace gives, DE:626), else None."""
    t_out = np.asarray(t_out, dtype=np.float64)
    if t_out.ndim != 1 or len(t_out) < 2:
        return None
    dt = (t_out[-1] - t_out[0]) / (len(t_out) - 1)
    if not dt > 0.0:
        return None
    want = t_out[0] + dt * np.arange(len(t_out))
    return float(dt) if np.max(np.abs(t_out - want)) <= rel * abs(t_out[-1] - t_out[0]) else None


def _max_squarings(value: Optional[int]) -> int:
    return int(os.environ.get("QSB200_PROP_SQUARINGS", "-1")) if value is None else int(value)


def propagate_batch(tables: Sequence[SpinTable], n_out: int, dt, opts: Sequence[SolverOptions], device: int = 0,
                    max_squarings: Optional[int] = None, by_column: bool = False):
    """Constant-H propagator path (qsb_propagate_batch): n <= 10, same n in every table, output k = state after k dt.
    ``dt``: one float or one per trajectory.  ``max_squarings``: cap on the scaling-and-squaring depth (None / < 0:
    automatic, 0: integrate the columns over the whole dt).  Returns (obs [n_traj, n_out, NOBS], [Stats]), or with
    ``by_column`` obs [n_traj, NOBS, n_out] (the device writes one contiguous trace per observable)."""
    lib = load_library()
    n = len(tables)
    if n == 0:
        return np.empty((0, 0, NOBS)), []
    dts = np.ascontiguousarray(np.atleast_1d(dt), dtype=np.float64)
    if dts.shape not in ((1,), (n,)):
        raise ValueError("dt must be a scalar or one value per trajectory")
    holders = [_TableHolder(t) for t in tables]
    ctabs = (_Table * n)(*[h.c for h in holders])
    psi0 = (C.c_uint64 * n)(*[t.psi0_index for t in tables])
    copts = (_Opts * n)(*[o.to_c() for o in opts])
    cstats = (_Stats * n)()
    obs = np.empty((n, NOBS, int(n_out)) if by_column else (n, int(n_out), NOBS), dtype=np.float64)
    code = lib.qsb_propagate_batch(n, ctabs, psi0, int(n_out), _dptr(dts), 1 if (dts.shape == (n,) and n > 1) else 0,
                                   copts, _max_squarings(max_squarings), 1 if by_column else 0, _dptr(obs), cstats,
                                   int(device))
    stats = [Stats.from_c(s) for s in cstats]
    _check(code)
    return obs, stats


def propagate_hbm(table: SpinTable, n_out: int, dt: float, opts: SolverOptions, device: int = 0,
                  max_squarings: Optional[int] = None, by_column: bool = False):
    """Constant-H propagator path for ONE trajectory of 11 <= n <= 14 qubits, U resident in HBM (qsb_propagate_hbm):
    output k = state after k dt.  Returns (obs [n_out, NOBS] or [NOBS, n_out] with ``by_column``, Stats)."""
    lib = load_library()
    holder = _TableHolder(table)
    copts = opts.to_c()
    st = _Stats()
    obs = np.empty((NOBS, int(n_out)) if by_column else (int(n_out), NOBS), dtype=np.float64)
    _check(lib.qsb_propagate_hbm(C.byref(holder.c), table.psi0_index, int(n_out), float(dt), C.byref(copts),
                                 _max_squarings(max_squarings), 1 if by_column else 0, _dptr(obs), C.byref(st), int(device)))
    return obs, Stats.from_c(st)


def _pauli_terms(terms):
    """ctypes array of qsb_pauli_term from (x_mask, z_mask, weight, column) tuples."""
    terms = list(terms)
    if not 1 <= len(terms) <= MAX_PAULI_TERMS:
        raise ValueError(f"need 1 .. {MAX_PAULI_TERMS} Pauli terms")
    return (_PauliTerm * len(terms))(*[_PauliTerm(int(xm), int(zm), float(w), int(col), 0) for xm, zm, w, col in terms])


def propagate_expect(table: SpinTable, n_out: int, dt: float, opts: SolverOptions, device: int = 0,
                     max_squarings: Optional[int] = None, by_column: bool = False, terms=None):
    """``propagate_batch`` / ``propagate_hbm`` for ONE trajectory whose six observables are weighted Pauli strings
    (``terms``, default ``table.pauli_terms``: a local-dimension-4 table).  n <= 10 bits - qsb_propagate_expect: U built
    as usual, the chain keeps every state on the device, one kernel reduces the strings; 11 <= n <= 14 -
    qsb_propagate_hbm_expect: U resident in HBM, the strings reduced on the chain's buffer after every matvec.
    Returns (obs [n_out, NOBS] or [NOBS, n_out] with ``by_column``, Stats)."""
    lib = load_library()
    terms = table.pauli_terms if terms is None else terms
    if terms is None:
        raise ValueError("the table carries no Pauli-string observables")
    cterms = _pauli_terms(terms)
    holder = _TableHolder(table)
    copts = opts.to_c()
    st = _Stats()
    obs = np.empty((NOBS, int(n_out)) if by_column else (int(n_out), NOBS), dtype=np.float64)
    fn = lib.qsb_propagate_expect if table.n <= SMALL_MAX_QUBITS else lib.qsb_propagate_hbm_expect
    _check(fn(C.byref(holder.c), table.psi0_index, int(n_out), float(dt), C.byref(copts), _max_squarings(max_squarings),
              cterms, len(cterms), 1 if by_column else 0, _dptr(obs), C.byref(st), int(device)))
    return obs, Stats.from_c(st)


def propagate_expect_batch(tables: Sequence[SpinTable], n_out: int, dt, opts: Sequence[SolverOptions], device: int = 0,
                           max_squarings: Optional[int] = None, by_column: bool = False, terms=None):
    """``propagate_expect`` for many trajectories of equal n <= 10 that share one list of Pauli strings (``terms``,
    default ``tables[0].pauli_terms``; a sweep's trajectories do) - qsb_propagate_expect_batch.  ``dt`` as in
    ``propagate_batch``.  Returns (obs [n_traj, n_out, NOBS] or [n_traj, NOBS, n_out] with ``by_column``, [Stats])."""
    lib = load_library()
    n = len(tables)
    if n == 0:
        return np.empty((0, 0, NOBS)), []
    if terms is None:
        terms = tables[0].pauli_terms
        if terms is None:
            raise ValueError("the tables carry no Pauli-string observables")
        if any(t.pauli_terms != terms for t in tables[1:]):
            raise ValueError("one call reduces ONE list of Pauli strings: group the tables by their pauli_terms")
    if any(t.n != tables[0].n for t in tables[1:]):
        raise ValueError("the tables of one call must have the same number of index bits")
    cterms = _pauli_terms(terms)
    dts = np.ascontiguousarray(np.atleast_1d(dt), dtype=np.float64)
    if dts.shape not in ((1,), (n,)):
        raise ValueError("dt must be a scalar or one value per trajectory")
    holders = [_TableHolder(t) for t in tables]
    ctabs = (_Table * n)(*[h.c for h in holders])
    psi0 = (C.c_uint64 * n)(*[t.psi0_index for t in tables])
    copts = (_Opts * n)(*[o.to_c() for o in opts])
    cstats = (_Stats * n)()
    obs = np.empty((n, NOBS, int(n_out)) if by_column else (n, int(n_out), NOBS), dtype=np.float64)
    code = lib.qsb_propagate_expect_batch(n, ctabs, psi0, int(n_out), _dptr(dts), 1 if (dts.shape == (n,) and n > 1) else 0,
                                          copts, _max_squarings(max_squarings), cterms, len(cterms), 1 if by_column else 0,
                                          _dptr(obs), cstats, int(device))
    stats = [Stats.from_c(s) for s in cstats]
    _check(code)
    return obs, stats


def build_propagator(table: SpinTable, dt: float, opts: SolverOptions, device: int = 0,
                     max_squarings: Optional[int] = None):
    """U = exp(-i H dt) as a (2^n, 2^n) complex matrix in the table's own index convention (bit-major), built
    column by column by the on-chip adaptive integrator (qsb_build_propagator).  Returns (U, Stats)."""
    lib = load_library()
    holder = _TableHolder(table)
    copts = opts.to_c()
    st = _Stats()
    ut = np.empty((table.dim, table.dim), dtype=np.complex128)  # ut[j, x] = <x|U|j>
    _check(lib.qsb_build_propagator(C.byref(holder.c), float(dt), C.byref(copts), _max_squarings(max_squarings),
                                    _dptr(ut.view(np.float64)), C.byref(st), int(device)))
    return np.ascontiguousarray(ut.T), Stats.from_c(st)


def propagator_squarings(table: SpinTable, dt: float) -> int:
    """Scaling-and-squaring depth the propagator build picks for this table and interval (host-only query)."""
    holder = _TableHolder(table)
    s = C.c_int32(0)
    _check(load_library().qsb_debug_squarings(C.byref(holder.c), float(dt), C.byref(s)))
    return s.value


def propagate_shape(n_qubits: int):
    """{threads, split, cached_columns, smem_bytes, ring} of the propagate kernel for n qubits (host-only query)."""
    out = (C.c_int32 * 5)()
    _check(load_library().qsb_debug_propagate_shape(int(n_qubits), out))
    return {"threads": out[0], "split": out[1], "cached_columns": out[2], "smem_bytes": out[3], "ring": out[4]}
