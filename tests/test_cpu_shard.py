"""Host-side logic of the sharded-state mode, no GPU: layouts, index ownership, the in-process all-gather, and the
torch.distributed all-gather hook under gloo with world_size 2 (the hook the C library calls once per rank barrier
and once per step for the error norm / observable sums)."""
import ctypes as C
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_layouts_cover_all_bits_and_centre_holds_rank_bits(world, monkeypatch):
    from quantumsimulations_b200 import engine

    wl = world.bit_length() - 1
    for gmax in (None, 2, 3, 5, 8):
        if gmax is None:
            monkeypatch.delenv("QSB200_MAX_GROUP", raising=False)
        else:
            monkeypatch.setenv("QSB200_MAX_GROUP", str(gmax))
        for n in range(12 + wl + (1 if world == 1 else 0), 35):
            lay = engine.tile_layouts(n, world)
            assert lay[0] == (12, 12, 12)
            covered = 12
            # groups of at most 8 bits (256-byte runs); unsharded plans may take 9 where that saves a layout
            widest = 9 if (world == 1 and gmax is None) else 8
            for lk, gs, ge in lay[1:]:
                assert gs == covered and 1 <= ge - gs <= widest and lk == 12 - (ge - gs)
                covered = ge
            if world == 1 and gmax is None:
                assert len(lay) - 1 == max(1, -(-(n - 12) // 9)) or n - 12 <= 16 and len(lay) - 1 == max(1, -(-(n - 12) // 8))
                assert (lay[-1][2] - lay[-1][1]) == min(ge - gs for _, gs, ge in lay[1:])   # the centre is the smallest group
            assert covered == n
            assert lay[-1][1] <= n - wl            # every rank bit is inside the centre group
            assert all(ge <= n - wl for _, _, ge in lay[1:-1])  # mid passes are local


def test_flagship_layouts():
    from quantumsimulations_b200 import engine

    assert engine.tile_layouts(28) == [(12, 12, 12), (4, 12, 20), (4, 20, 28)]
    # sharded: the centre group is capped at 5 bits where that adds no layout (NVLink throughput of the P2P centre
    # pass falls with the rows per tile, profiles/r1zz_centre_group_2gpu.log)
    assert engine.tile_layouts(33, 8) == [(12, 12, 12), (4, 12, 20), (4, 20, 28), (7, 28, 33)]
    assert engine.tile_layouts(31, 2) == [(12, 12, 12), (5, 12, 19), (5, 19, 26), (7, 26, 31)]
    assert engine.tile_layouts(28, 2) == [(12, 12, 12), (4, 12, 20), (4, 20, 28)]      # 5 bits would add a layout
    # unsharded: 9-bit groups where they save a layout (two HBM passes), the centre on the smallest group
    assert engine.tile_layouts(30, 1) == [(12, 12, 12), (3, 12, 21), (3, 21, 30)]
    assert engine.tile_layouts(29, 1) == [(12, 12, 12), (3, 12, 21), (4, 21, 29)]
    assert engine.tile_layouts(27, 1) == [(12, 12, 12), (4, 12, 20), (5, 20, 27)]
    with pytest.raises(ValueError):
        engine.tile_layouts(14, 8)
    with pytest.raises(ValueError):
        engine.tile_layouts(20, 3)


def test_owner_and_global_sites():
    from quantumsimulations_b200 import shard

    # initial_state_rare's ket (DE:591-606): index 2^N - 2 lives on the last rank
    assert shard.owner_of((1 << 33) - 2, 33, 8) == (7, (1 << 30) - 2)
    assert shard.owner_of(5, 20, 4) == (0, 5)
    assert shard.global_sites(33, 8) == [0, 1, 2]   # site k is bit N-1-k: the first sea spins are the rank bits
    assert shard.shard_geometry(28, 1) == (0, 28, 1 << 28)
    with pytest.raises(ValueError):
        shard.shard_geometry(13, 4)
    with pytest.raises(ValueError):
        shard.shard_geometry(20, 6)


def test_local_allgather_threads():
    from quantumsimulations_b200 import engine

    lib = engine.load_library()
    world = 4
    grp = C.c_void_p()
    assert lib.qsb_local_group_create(world, C.byref(grp)) == 0
    out = [None] * world

    def work(r):
        m = lib.qsb_local_group_member(grp, r)
        dp = C.POINTER(C.c_double)
        for it in range(50):
            send = np.arange(3, dtype=float) + 10 * r + it
            recv = np.zeros(3 * world)
            assert lib.qsb_local_allgather(m, send.ctypes.data_as(dp), 3, recv.ctypes.data_as(dp)) == 0
            assert lib.qsb_local_allgather(m, None, 0, None) == 0
        out[r] = recv

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join() for t in th]
    want = np.concatenate([np.arange(3, dtype=float) + 10 * r + 49 for r in range(world)])
    for r in range(world):
        assert np.array_equal(out[r], want)
    lib.qsb_local_group_destroy(grp)


def test_local_barrier_times_out_instead_of_hanging(monkeypatch):
    from quantumsimulations_b200 import engine

    monkeypatch.setenv("QSB200_BARRIER_TIMEOUT_S", "0.2")
    lib = engine.load_library()
    grp = C.c_void_p()
    assert lib.qsb_local_group_create(2, C.byref(grp)) == 0
    m = lib.qsb_local_group_member(grp, 0)
    assert lib.qsb_local_allgather(m, None, 0, None) == 1   # rank 1 never arrives
    assert lib.qsb_local_allgather(m, None, 0, None) == 1   # and the group stays broken
    lib.qsb_local_group_destroy(grp)


WORKER = r"""
import ctypes as C, os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np
import torch.distributed as dist
from quantumsimulations_b200 import shard
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
hook = shard.make_torch_allgather(world)
dp = C.POINTER(C.c_double)
send = np.array([rank + 0.5, 10.0 * rank, -1.0])
recv = np.zeros(3 * world)
assert hook(None, send.ctypes.data_as(dp), 3, recv.ctypes.data_as(dp)) == 0
want = np.concatenate([[r + 0.5, 10.0 * r, -1.0] for r in range(world)])
assert np.array_equal(recv, want), (recv, want)
assert hook(None, None, 0, None) == 0
blobs = [None] * world
dist.all_gather_object(blobs, bytes([rank]) * 96)
assert [b[0] for b in blobs] == list(range(world))
dist.destroy_process_group()
print("ok", rank)
"""


def test_torch_allgather_hook_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29731", str(script), ROOT]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=240, env=env)
    assert res.returncode == 0, res.stdout + res.stderr
    assert res.stdout.count("ok") == 2
