#!/usr/bin/env python
"""Sharded-state leg of bench.py alone, for several per-GPU shard sizes (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
        scripts/time_shard.py 26 27 28 29 30

Prints one JSON line per size on rank 0: ms per H-apply, per-class kernel ms, NVLink GB/s of the centre pass.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def allmax(x):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


for bits in [int(a) for a in sys.argv[1:]] or [30]:
    args = argparse.Namespace(shard_local_bits=bits)
    out = bench.sharded_leg(args, rank, world, local, barrier, allmax)
    if rank == 0:
        keep = {k: out[k] for k in ("n_qubits", "world", "ms_per_happly", "layouts", "kernel_ms", "hbm_passes_ms_per_happly",
                                    "overlap", "parity_max_rel_err", "vs_unsharded", "strong")}
        keep["nvlink"] = {k: out["nvlink"][k] for k in ("centre_ms", "achieved_GBps_per_direction", "frac")} if world > 1 else None
        print(json.dumps(keep), flush=True)
    torch.cuda.empty_cache()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
