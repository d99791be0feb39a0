"""GPU parity of the sharded-state path (BASELINE config C5).

The ranks are threads of one process (LocalShardGroup), each with its own plan, stream and exchange scratch; with
one GPU on the box they all sit on cuda:0 and the "peer" loads/stores of the centre pass are plain device accesses
into the other ranks' buffers.  The rank barrier is host-side (stream drain + thread barrier), so nothing spins on
the device.  Checked: one H-apply against the numpy oracle of the table semantics, and whole trajectories against
the unsharded engine on the same table (same kernels, different partition: agreement to rounding).
"""
import numpy as np
import pytest

from conftest import exact_traces, params_from_meta, random_table

pytestmark = pytest.mark.gpu


def _rand_state(n, seed):
    rng = np.random.default_rng(seed)
    v = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    return v / np.linalg.norm(v)


@pytest.mark.parametrize("n,world", [(13, 2), (14, 2), (14, 4), (15, 8), (16, 4), (18, 8)])
def test_sharded_happly_random_table(n, world):
    from oracle import dense
    from quantumsimulations_b200 import shard

    table = random_table(n, 500 + n + world)
    bm = table.bit_major()
    psi = _rand_state(n, n * world)
    want = dense.apply_h(bm["h"], bm["J"], psi)
    with shard.LocalShardGroup(table, world) as grp:
        got = grp.apply_h(psi)
    err = np.max(np.abs(got - want)) / np.max(np.abs(want))
    assert err < 2e-13, f"n={n} world={world}: rel err {err:.3e}"


@pytest.mark.parametrize("n,world,gmax", [(17, 2, 3), (18, 4, 2), (19, 2, 4), (19, 8, 3)])
def test_sharded_happly_multi_group(n, world, gmax, monkeypatch):
    """Mid passes on local groups plus a centre group that mixes local and rank bits."""
    from oracle import dense
    from quantumsimulations_b200 import shard

    monkeypatch.setenv("QSB200_MAX_GROUP", str(gmax))
    table = random_table(n, 600 + n + world, dense_pairs=False)
    bm = table.bit_major()
    psi = _rand_state(n, 3 * n + world)
    want = dense.apply_h(bm["h"], bm["J"], psi)
    with shard.LocalShardGroup(table, world) as grp:
        got = grp.apply_h(psi)
    err = np.max(np.abs(got - want)) / np.max(np.abs(want))
    assert err < 2e-13, f"n={n} world={world} gmax={gmax}: rel err {err:.3e}"


@pytest.mark.parametrize("method,world", [("dp5", 2), ("taylor", 4), ("taylor", 8)])
def test_sharded_trajectory_matches_unsharded_and_exact(method, world, golden_cases):
    """N = 13 reference geometry (dataclass default, icosahedron) padded by nothing: 13 - 3 < 12, so use N = 15."""
    from bench import sweep_params
    from quantumsimulations_b200 import compile_table, engine, shard

    table = compile_table(sweep_params(14, 50e3, t_final=2e-5, steps=9))  # N = 15, Fibonacci shell, rare ON
    t = np.linspace(0.0, 2e-5, 9)
    opts = engine.SolverOptions(atol=1e-12, rtol=1e-11, max_step=1e-5, nsteps=100000, method=method,
                                order=16 if method == "taylor" else 0)
    with engine.Plan(table) as plan:
        ref, st_ref = plan.evolve(t, opts)
    with shard.LocalShardGroup(table, world) as grp:
        got, st = grp.evolve(t, opts)
    assert st.n_accepted == st_ref.n_accepted and st.n_rejected == st_ref.n_rejected
    assert np.max(np.abs(got - ref)) < 1e-11
    want = exact_traces(table, t) if table.n <= 13 else None
    if want is not None:
        for i, k in enumerate(("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")):
            assert np.max(np.abs(got[:, i] - want[k])) < 1e-6


def test_sharded_exact_n13_world2(golden_cases):
    """The reference's dataclass-default geometry (N = 13, DE:332-384) split over 2 ranks vs exact diagonalisation."""
    from quantumsimulations_b200 import compile_table, engine, shard

    meta, _ = golden_cases
    table = compile_table(params_from_meta(meta["n13_dataclass_default"]))
    t = np.linspace(0.0, 2e-3, 6)
    opts = engine.SolverOptions(atol=1e-11, rtol=1e-10, nsteps=100000, method="taylor", order=20)
    with shard.LocalShardGroup(table, 2) as grp:
        got, _ = grp.evolve(t, opts)
    want = exact_traces(table, t)
    for i, k in enumerate(("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")):
        assert np.max(np.abs(got[:, i] - want[k])) < 1e-6, k


def test_sharded_state_roundtrip_and_host_psi0():
    from quantumsimulations_b200 import engine, shard

    n, world = 14, 4
    table = random_table(n, 77)
    psi0 = _rand_state(n, 5)
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, method="dp5")
    with engine.Plan(table) as plan:
        plan.session_begin(opts, psi0=psi0)
        plan.session_step(3)
        want = plan.session_state()
        wobs = plan.session_observe()
        plan.session_end()
    with shard.LocalShardGroup(table, world) as grp:
        def work(plan, r):
            plan.session_begin(opts, psi0=psi0[plan.local_slice()])
            plan.session_step(3)
            out = plan.session_state(), plan.session_observe()
            plan.session_end()
            return out
        res = grp.run(work)
    got = np.concatenate([r[0] for r in res])
    assert np.max(np.abs(got - want)) < 1e-13
    for r in res:
        assert np.max(np.abs(r[1] - wobs)) < 1e-12


def test_unconnected_sharded_plan_refuses():
    from quantumsimulations_b200 import engine

    table = random_table(13, 9)
    with engine.ShardedPlan(table, 0, 2) as plan:
        with pytest.raises(ValueError):
            plan.session_begin(engine.SolverOptions())
    with pytest.raises((ValueError, engine.EngineError)):
        engine.ShardedPlan(random_table(12, 9), 0, 2)


def test_multi_process_parity_two_gpus():
    """One process per GPU over CUDA IPC + NCCL (needs >= 2 GPUs on the box; skipped otherwise)."""
    import os
    import subprocess
    import sys

    from quantumsimulations_b200 import engine

    if engine.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29733", os.path.join(root, "scripts", "shard_parity.py"), "--qubits", "18"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert '"ok": true' in res.stdout


@pytest.mark.parametrize("n,world", [(16, 2), (19, 4), (21, 8)])
def test_ring_centre_pass_is_bitwise_the_register_staged_one(n, world):
    """The TMA-ring version of the sharded centre pass (QSB200_SHARD_RING=1, read once per process) gives the same
    observables, bit for bit, as the default P2P load/store kernel (ranks = threads on one GPU)."""
    import json
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for ring in ("0", "1"):
        env = dict(os.environ, QSB200_SHARD_RING=ring)
        res = subprocess.run([sys.executable, os.path.join(root, "scripts", "shard_ring_check.py"), "--n", str(n), "--world", str(world)],
                             capture_output=True, text=True, timeout=300, env=env)
        assert res.returncode == 0, res.stdout[-1500:] + res.stderr[-1500:]
        outs.append(json.loads(res.stdout.strip().splitlines()[-1]))
    assert outs[0]["ring"] == "0" and outs[1]["ring"] == "1"
    assert outs[0]["obs"] == outs[1]["obs"] and outs[0]["rhs"] == outs[1]["rhs"]


@pytest.mark.parametrize("n,world,gmax", [(14, 2, None), (16, 4, None), (18, 8, None), (19, 2, 4), (19, 8, 3), (21, 4, None)])
def test_pipelined_centre_kernel(n, world, gmax, monkeypatch):
    """pass_centre_pipe - the persistent, double-buffered (cp.async) centre pass the overlapped multi-GPU mode runs -
    forced into the serialised path (QSB200_CENTRE_PIPE=2) so that a one-GPU box exercises it: H-apply against the
    numpy oracle, and the observable sums it collects against the default kernel's, bit for bit."""
    from oracle import dense
    from quantumsimulations_b200 import engine, shard

    if gmax is not None:
        monkeypatch.setenv("QSB200_MAX_GROUP", str(gmax))
    table = random_table(n, 800 + n + world, dense_pairs=False)
    bm = table.bit_major()
    psi = _rand_state(n, 7 * n + world)
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, method="taylor", order=6)

    def run():
        with shard.LocalShardGroup(table, world) as grp:
            got = grp.apply_h(psi)
            obs, st = grp.evolve(np.linspace(0.0, 1e-3, 3), opts, psi)
        return got, obs, st

    monkeypatch.setenv("QSB200_CENTRE_PIPE", "0")
    ref_h, ref_obs, ref_st = run()
    monkeypatch.setenv("QSB200_CENTRE_PIPE", "2")
    got_h, got_obs, got_st = run()
    want = dense.apply_h(bm["h"], bm["J"], psi)
    assert np.max(np.abs(got_h - want)) / np.max(np.abs(want)) < 2e-13
    assert np.array_equal(got_h, ref_h) and np.array_equal(got_obs, ref_obs)
    assert (got_st.n_accepted, got_st.n_rejected) == (ref_st.n_accepted, ref_st.n_rejected)


@pytest.mark.parametrize("n,world,gmax,pipe", [(14, 4, None, 3), (16, 2, None, 3), (18, 8, None, 2), (19, 2, 4, 3), (20, 4, 3, 4),
                                               (21, 8, None, 1), (22, 2, None, 3)])
def test_pipelined_strided_passes(n, world, gmax, pipe, monkeypatch):
    """The chunk-pipelined H-apply of the one-process-per-GPU mode (qsb_shard_set_stream_barrier: chunk-group tile maps,
    persistent centre pass, two streams + events) with ranks as threads and the HOST-blocking barrier
    qsb_local_stream_barrier, so that a one-GPU box runs that code path: H psi against the numpy oracle and, bit for
    bit, against the serialised passes; trajectories (observable sums collected by the chunked centre passes) too."""
    from oracle import dense
    from quantumsimulations_b200 import engine, shard

    if gmax is not None:
        monkeypatch.setenv("QSB200_MAX_GROUP", str(gmax))
    monkeypatch.setenv("QSB200_SHARD_PIPE", str(pipe))
    table = random_table(n, 880 + n + world, dense_pairs=False)
    bm = table.bit_major()
    psi = _rand_state(n, 11 * n + world)
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, method="dp5")
    t = np.linspace(0.0, 1e-3, 3)
    with shard.LocalShardGroup(table, world) as grp:
        ref_h = grp.apply_h(psi)
        ref_obs, ref_st = grp.evolve(t, opts, psi)
    with shard.LocalShardGroup(table, world, pipeline=True) as grp:
        got_h = grp.apply_h(psi)
        got_obs, got_st = grp.evolve(t, opts, psi)
    want = dense.apply_h(bm["h"], bm["J"], psi)
    assert np.max(np.abs(got_h - want)) / np.max(np.abs(want)) < 2e-13
    assert np.array_equal(got_h, ref_h) and np.array_equal(got_obs, ref_obs)
    assert (got_st.n_accepted, got_st.n_rejected) == (ref_st.n_accepted, ref_st.n_rejected)


@pytest.mark.parametrize("n,world,gmax,pipeline", [(16, 2, 2, False), (17, 4, 3, False), (19, 8, 3, False), (19, 2, 4, True),
                                                    (20, 4, 3, True), (21, 8, 5, True), (22, 2, None, True), (14, 4, None, False)])
def test_push_only_exchange(n, world, gmax, pipeline, monkeypatch):
    """QSB200_SHARD_PUSH=1: the last forward mid pass stores every amplitude into the receive buffer of the rank that
    owns its centre tile, the centre pass loads locally (pass_mid<.., PUSH>, pass_centre<.., PUSHED>); serialised and
    chunk-pipelined passes, ranks as threads.  H psi against the numpy oracle and, bit for bit, against the pulling
    exchange; trajectories too.  (n = 14 x 4 has no mid pass: it must fall back to pulling.)"""
    from oracle import dense
    from quantumsimulations_b200 import engine, shard

    if gmax is not None:
        monkeypatch.setenv("QSB200_MAX_GROUP", str(gmax))
    table = random_table(n, 990 + n + world, dense_pairs=False)
    bm = table.bit_major()
    psi = _rand_state(n, 13 * n + world)
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, method="taylor", order=6)
    t = np.linspace(0.0, 1e-3, 3)

    def run():
        with shard.LocalShardGroup(table, world, pipeline=pipeline) as grp:
            return grp.apply_h(psi), grp.evolve(t, opts, psi)

    monkeypatch.setenv("QSB200_SHARD_PUSH", "0")
    ref_h, (ref_obs, ref_st) = run()
    monkeypatch.setenv("QSB200_SHARD_PUSH", "1")
    got_h, (got_obs, got_st) = run()
    want = dense.apply_h(bm["h"], bm["J"], psi)
    assert np.max(np.abs(got_h - want)) / np.max(np.abs(want)) < 2e-13
    assert np.array_equal(got_h, ref_h) and np.array_equal(got_obs, ref_obs)
    assert (got_st.n_accepted, got_st.n_rejected) == (ref_st.n_accepted, ref_st.n_rejected)
