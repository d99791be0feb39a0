"""Multi-GPU plumbing for the batched-sweep mode: independent trajectories, no data-path collective.

Two ways to use several B200s of one box ("batched sweep"):

* ``simulate_on_devices``  - ONE process drives several GPUs, one host thread per device (the C calls release
  the GIL); used by ``sweep.run_sweep_sea_detuning(devices=[...])``.
* ``simulate_distributed`` - one process per GPU under ``torchrun``; every rank integrates its share and the
  host-side results (7 x steps doubles per trajectory) are exchanged with ``all_gather_object``.  This is result
  plumbing over torch.distributed, not a collective inside the hot path.

Trajectory cost grows with ||H|| (about 4x between 0 and 150 kHz detuning), so shares are balanced
longest-first on a cheap norm bound rather than round-robin.
"""
from __future__ import annotations

import threading
from typing import Callable, List, Optional, Sequence

from .model import DipolarRareParams
from .table import compile_table


def estimate_cost(params: DipolarRareParams) -> float:
    """Relative integration cost: (bound on ||H||) x t_final x 2^N."""
    table = compile_table(params)
    return table.spectral_bound() * params.t_final * float(table.dim)


def partition(costs: Sequence[float], n_bins: int) -> List[List[int]]:
    """Longest-processing-time-first assignment of items to ``n_bins`` bins; indices inside a bin stay sorted."""
    if n_bins < 1:
        raise ValueError("n_bins must be >= 1")
    bins: List[List[int]] = [[] for _ in range(n_bins)]
    load = [0.0] * n_bins
    for i in sorted(range(len(costs)), key=lambda k: (-costs[k], k)):
        b = min(range(n_bins), key=lambda k: (load[k], k))
        bins[b].append(i)
        load[b] += costs[i]
    return [sorted(b) for b in bins]


def _run_batch(plist, device, **kw):
    from .api import simulate_rare_batch

    results, _ = simulate_rare_batch(plist, device=device, **kw)
    return results


def simulate_on_devices(plist: Sequence[DipolarRareParams], devices: Optional[Sequence[int]] = None, **kw) -> List:
    """[(t, obs)] in input order, trajectories spread over ``devices`` (default: device 0 only)."""
    devices = list(devices) if devices else [0]
    if len(devices) == 1 or len(plist) <= 1:
        return _run_batch(list(plist), devices[0], **kw)
    shares = partition([estimate_cost(p) for p in plist], len(devices))
    out: List = [None] * len(plist)
    errors: List[BaseException] = []

    def work(dev, idxs):
        try:
            if idxs:
                for i, r in zip(idxs, _run_batch([plist[i] for i in idxs], dev, **kw)):
                    out[i] = r
        except BaseException as exc:  # re-raised in the caller's thread
            errors.append(exc)

    threads = [threading.Thread(target=work, args=(d, s)) for d, s in zip(devices, shares)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    if errors:
        raise errors[0]
    return out


def simulate_distributed(plist: Sequence[DipolarRareParams], device: Optional[int] = None,
                         compute: Optional[Callable] = None, **kw) -> List:
    """Every rank returns the full result list.  Needs an initialised torch.distributed process group.

    ``compute(sub_list, device)`` defaults to the engine; tests inject a CPU stand-in to exercise the sharding
    and gathering logic with the gloo backend.
    """
    import torch.distributed as dist

    if not dist.is_initialized():
        raise RuntimeError("simulate_distributed needs torch.distributed.init_process_group() first")
    rank, world = dist.get_rank(), dist.get_world_size()
    shares = partition([estimate_cost(p) for p in plist], world)
    mine = shares[rank]
    run = compute or (lambda sub, dev: _run_batch(sub, dev, **kw))
    if device is None:
        import torch

        device = torch.cuda.current_device() if torch.cuda.is_available() else 0
    local = run([plist[i] for i in mine], device) if mine else []
    gathered: List = [None] * world
    dist.all_gather_object(gathered, list(zip(mine, local)))
    out: List = [None] * len(plist)
    for chunk in gathered:
        for i, r in chunk:
            out[i] = r
    return out
