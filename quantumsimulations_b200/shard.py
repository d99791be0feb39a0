"""Sharded-state mode (BASELINE config C5): one trajectory whose state vector is split over several GPUs.

The amplitude index's top log2(world) bits are the rank.  Site k of the reference is bit N-1-k
(dipolar_ensemble_with_rare.py:43-45), so the first log2(world) sea spins are the "global" qubits.  All passes of the
H-apply are local except the centre pass, which works on the peers' transform scratch through P2P loads/stores over
NVLink (CUDA IPC between processes); see DESIGN.md "Sharded state".  Two connectors:

* ``connect_torch_distributed(plan)`` - one process per GPU (torchrun); blobs and scalar all-gathers travel over the
  default torch.distributed group (NCCL: tiny CUDA tensors; gloo: CPU tensors).
* ``LocalShardGroup`` - ``world`` ranks as threads of THIS process, each with its own plan, stream and (if there are
  enough) its own GPU; the all-gather is the library's in-process one.  This is also how the sharded path is tested
  on a one-GPU box.

torch is plumbing here (rendezvous + scalar exchange); the amplitudes never pass through it.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Callable, List, Optional, Sequence

import numpy as np

from . import engine
from .table import SpinTable


def shard_geometry(n_qubits: int, world: int):
    """(log2 world, local bits, local dim); raises like the C layer for unsupported splits."""
    wl = world.bit_length() - 1
    if world < 1 or world > 8 or (1 << wl) != world:
        raise ValueError("world must be 1, 2, 4 or 8")
    n_loc = n_qubits - wl
    if world > 1 and n_loc < 12:
        raise ValueError("a sharded plan needs n_qubits - log2(world) >= 12")
    return wl, n_loc, 1 << n_loc


def owner_of(index: int, n_qubits: int, world: int):
    """(rank, local index) of a computational-basis index (initial_state_rare's ket, DE:591-606)."""
    _, n_loc, dim = shard_geometry(n_qubits, world)
    return index >> n_loc, index & (dim - 1)


def global_sites(n_qubits: int, world: int) -> List[int]:
    """Reference site numbers whose qubits are the rank bits (site k = bit N-1-k)."""
    wl, _, _ = shard_geometry(n_qubits, world)
    return list(range(wl))


def make_torch_allgather(world: int, device=None):
    """An ALLGATHER_FN over the default torch.distributed group; rank-ordered, identical bits on every rank."""
    import torch
    import torch.distributed as dist

    backend = dist.get_backend()
    dev = torch.device("cpu")
    if "nccl" in str(backend):
        dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)

    def hook(_ctx, send, count, recv):
        try:
            if count == 0:
                dist.barrier()
                return 0
            mine = torch.from_numpy(np.ctypeslib.as_array(send, shape=(count,)).copy()).to(dev)
            out = torch.empty(world * count, dtype=torch.float64, device=dev)
            dist.all_gather_into_tensor(out, mine)
            np.ctypeslib.as_array(recv, shape=(world * count,))[:] = out.cpu().numpy()
            return 0
        except Exception as exc:  # never let an exception cross the C boundary
            print(f"[qsb200] all-gather hook failed: {exc!r}", flush=True)
            return 1

    return engine.ALLGATHER_FN(hook)


_BARRIER_GROUPS: dict = {}


def make_torch_stream_barrier(device=None, n_streams: int = 2):
    """A STREAM_BARRIER_FN over NCCL: a one-element all-reduce issued with the given CUDA stream as torch's current
    stream, so the collective is ordered after that stream's earlier work and the stream waits for it; the host does
    not block.  Every distinct stream gets its OWN process group (communicator): collectives of one communicator must
    not be in flight on two streams at once - their kernels could pair up across ranks in different orders, and a
    barrier that pairs with the wrong peer call is no barrier (measured: wrong amplitudes at N >= 24).  Collective:
    all ranks must call this (it creates the groups) and later issue their barriers in the same order per stream."""
    import torch
    import torch.distributed as dist

    dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
    key = (dist.get_world_size(), n_streams, dev.index)
    if key not in _BARRIER_GROUPS:              # one set of communicators per process, shared by every plan it makes
        groups = [dist.new_group(backend="nccl") for _ in range(n_streams)]
        tokens = [torch.zeros(1, dtype=torch.float32, device=dev) for _ in range(n_streams)]
        for g, tok in zip(groups, tokens):      # build the communicators now, not inside the first H-apply
            dist.all_reduce(tok, group=g)
        torch.cuda.synchronize(dev)
        _BARRIER_GROUPS[key] = (groups, tokens)
    groups, tokens = _BARRIER_GROUPS[key]
    slots = {}

    def hook(_ctx, stream_ptr):
        try:
            slot = slots.get(stream_ptr)
            if slot is None:
                if len(slots) >= n_streams:
                    raise RuntimeError("more streams than barrier groups")
                slot = slots[stream_ptr] = (len(slots), torch.cuda.ExternalStream(stream_ptr, device=dev))
            k, s = slot
            with torch.cuda.stream(s):
                dist.all_reduce(tokens[k], group=groups[k])
            return 0
        except Exception as exc:  # never let an exception cross the C boundary
            print(f"[qsb200] stream-barrier hook failed: {exc!r}", flush=True)
            return 1

    return engine.STREAM_BARRIER_FN(hook)


def connect_torch_distributed(plan: "engine.ShardedPlan", overlap: Optional[bool] = None) -> None:
    """Exchange the exchange-scratch descriptors over torch.distributed and map the peers (collective).
    ``overlap`` (default: on with the NCCL backend, QSB200_SHARD_OVERLAP=0 turns it off): install the stream-ordered
    barrier, i.e. run the centre pass of one branch under the HBM passes of the other."""
    import os

    import torch.distributed as dist

    if not dist.is_initialized():
        raise RuntimeError("connect_torch_distributed needs torch.distributed.init_process_group() first")
    if dist.get_world_size() != plan.world or dist.get_rank() != plan.rank:
        raise ValueError("plan rank/world do not match the torch.distributed group")
    blobs: List[Optional[bytes]] = [None] * plan.world
    dist.all_gather_object(blobs, plan.export_blob())
    plan.connect(blobs, make_torch_allgather(plan.world))
    if overlap is None:
        overlap = os.environ.get("QSB200_SHARD_OVERLAP", "1") != "0"
    if overlap and "nccl" in str(dist.get_backend()):
        plan.set_stream_barrier(make_torch_stream_barrier(plan.device_index))


def sharded_plan_from_env(table: SpinTable, device: Optional[int] = None) -> "engine.ShardedPlan":
    """Under torchrun: this rank's plan of a state sharded over the whole group, connected and ready."""
    import torch
    import torch.distributed as dist

    rank, world = dist.get_rank(), dist.get_world_size()
    if device is None:
        device = torch.cuda.current_device()
    plan = engine.ShardedPlan(table, rank, world, device)
    connect_torch_distributed(plan)
    return plan


class LocalShardGroup:
    """``world`` ranks as threads of this process.  ``run(fn)`` calls ``fn(plan, rank)`` on every rank concurrently
    (the engine's calls are collective, so they must overlap in time) and returns the per-rank results."""

    def __init__(self, table: SpinTable, world: int, devices: Optional[Sequence[int]] = None, pipeline: bool = False):
        """``pipeline``: install the host-blocking stream barrier (qsb_local_stream_barrier), i.e. run the chunk-pipelined
        H-apply of the multi-process mode - serialised, since these ranks may share a GPU - instead of the plain
        serialised passes."""
        shard_geometry(table.n, world)
        self.lib = engine.load_library()
        self.world = world
        n_dev = engine.device_count()
        self.devices = list(devices) if devices else [r % max(n_dev, 1) for r in range(world)]
        self._group = C.c_void_p()
        engine._check(self.lib.qsb_local_group_create(world, C.byref(self._group)))
        self.plans: List[engine.ShardedPlan] = [engine.ShardedPlan(table, r, world, self.devices[r]) for r in range(world)]
        blobs = [p.export_blob() for p in self.plans]
        fn = C.cast(self.lib.qsb_local_allgather, C.c_void_p).value
        self.run(lambda plan, r: plan.connect(blobs, fn, self.lib.qsb_local_group_member(self._group, r)))
        if pipeline:
            sb = C.cast(self.lib.qsb_local_stream_barrier, C.c_void_p).value
            for r, plan in enumerate(self.plans):
                plan.set_stream_barrier(sb, self.lib.qsb_local_group_member(self._group, r))

    def run(self, fn: Callable) -> List:
        out: List = [None] * self.world
        errs: List[BaseException] = []

        def work(r):
            try:
                out[r] = fn(self.plans[r], r)
            except BaseException as exc:
                errs.append(exc)

        threads = [threading.Thread(target=work, args=(r,)) for r in range(self.world)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        if errs:
            raise errs[0]
        return out

    # ---- whole-vector conveniences (host vectors are split / joined here) -----------------------------
    def apply_h(self, psi: np.ndarray) -> np.ndarray:
        psi = np.ascontiguousarray(psi, dtype=np.complex128)
        parts = self.run(lambda plan, r: plan.apply_h(psi[plan.local_slice()]))
        return np.concatenate(parts)

    def evolve(self, t_out: np.ndarray, opts: "engine.SolverOptions", psi0: Optional[np.ndarray] = None):
        """(obs, stats) of rank 0; every rank computes the same observables (checked)."""
        if psi0 is not None:
            psi0 = np.ascontiguousarray(psi0, dtype=np.complex128)
        res = self.run(lambda plan, r: plan.evolve(t_out, opts, None if psi0 is None else psi0[plan.local_slice()]))
        for obs, _ in res[1:]:
            if not np.array_equal(obs, res[0][0]):
                raise RuntimeError("ranks disagree on the observables (they must be bit-identical)")
        return res[0]

    def close(self) -> None:
        for p in self.plans:
            p.close()
        self.plans = []
        if self._group:
            self.lib.qsb_local_group_destroy(self._group)
            self._group = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
