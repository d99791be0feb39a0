"""GPU parity of whole trajectories against reference-generated goldens and the oracle solvers."""
import numpy as np
import pytest

from conftest import params_from_meta

pytestmark = pytest.mark.gpu
KEYS = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")
TOL = 1e-6  # north_star: <= 1e-6 absolute on the normalised magnetisation traces


def _max_dev(out, arrs, name):
    return max(float(np.max(np.abs(out[k] - arrs[f"{name}/{k}"]))) for k in KEYS)


@pytest.mark.parametrize("method", ["dp5", "taylor"])
def test_n7_sweep_points_tight_tolerance(golden_cases, method):
    """Engine vs exact (reference-built H, exact propagation) at tightened tolerances: gate 1."""
    import dataclasses
    from quantumsimulations_b200 import simulate_rare_batch

    meta, arrs = golden_cases
    names = [k for k in meta if k.startswith("n7_d") and meta[k]["backend"] == "exact"]
    plist = [dataclasses.replace(params_from_meta(meta[k]), solver_atol=1e-13, solver_rtol=1e-12) for k in names]
    res, stats = simulate_rare_batch(plist, method=method)
    for name, (t, out), st in zip(names, res, stats):
        assert np.array_equal(t, arrs[f"{name}/t"])
        dev = _max_dev(out, arrs, name)
        assert dev < TOL, f"{name} [{method}]: max |E-X| = {dev:.2e} ({st})"
        assert np.all(out["state_norm"] == 1.0)


def test_reference_tolerances_not_worse_than_zvode(golden_cases):
    """At the sweep's own atol/rtol the engine must be at least as close to exact as the reference-equivalent Q."""
    from quantumsimulations_b200 import simulate_rare

    meta, arrs = golden_cases
    x, q = "n7_d050k_center_on", "q_n7_d050k_center_on_30ms"
    p = params_from_meta(meta[x])
    t, out = simulate_rare(p)
    q_dev = max(float(np.max(np.abs(arrs[f"{q}/{k}"] - arrs[f"{x}/{k}"]))) for k in KEYS)
    e_dev = _max_dev(out, arrs, x)
    assert e_dev <= q_dev, f"engine {e_dev:.2e} vs zvode {q_dev:.2e}"


def test_short_window_matches_zvode(golden_cases):
    """Gate 2a: same solver settings, short window.  Q itself is ~1.5e-6 from exact there, so the bound is
    |E-Q| <= 1e-6 + |Q-X| with |Q-X| taken from the goldens."""
    from quantumsimulations_b200 import simulate_rare

    meta, arrs = golden_cases
    q, x = "q_n7_d050k_center_on_short", "x_n7_d050k_center_on_short"
    t, out = simulate_rare(params_from_meta(meta[q]))
    q_dev = max(float(np.max(np.abs(arrs[f"{q}/{k}"] - arrs[f"{x}/{k}"]))) for k in KEYS)
    assert _max_dev(out, arrs, x) < TOL
    assert _max_dev(out, arrs, q) < TOL + q_dev


def test_converged_zvode_matches(golden_cases):
    """Gate 2b: where the reference-equivalent integrator is converged (atol 1e-14 / rtol 1e-13), E and Q agree
    to 1e-6 over the whole 0.03 s window at the same settings."""
    from quantumsimulations_b200 import simulate_rare

    meta, arrs = golden_cases
    name = "q_n7_d050k_center_on_30ms_tight"
    t, out = simulate_rare(params_from_meta(meta[name]))
    assert _max_dev(out, arrs, name) < TOL


@pytest.mark.parametrize("name", ["n4_fib3_default_opts", "n5_tetra", "n6_fib5_shell_off", "n9_cube", "n7_spin_up_phases"])
def test_other_geometries(golden_cases, name):
    import dataclasses
    from quantumsimulations_b200 import simulate_rare

    meta, arrs = golden_cases
    p = params_from_meta(meta[name])
    tight = dataclasses.replace(p, solver_atol=1e-13, solver_rtol=1e-12) if p.solver_atol is not None else p
    t, out = simulate_rare(tight)
    tol = TOL if p.solver_atol is not None else 2e-4   # QuTiP-default tolerances (1e-8/1e-6) are loose
    assert _max_dev(out, arrs, name) < tol
    if p.solver_atol is None:
        assert out["state_norm"].size == 0      # DE:669 with options=None


def test_n13_dataclass_default_tiled_path(golden_cases):
    """C1 (ii): dataclass defaults (N=13) run on the tiled HBM path; compare with exact at tight tolerance."""
    import dataclasses
    from quantumsimulations_b200 import simulate_rare

    meta, arrs = golden_cases
    name = "n13_dataclass_default"
    p = dataclasses.replace(params_from_meta(meta[name]), solver_atol=1e-12, solver_rtol=1e-11, solver_nsteps=100000)
    for method in ("dp5", "taylor"):
        t, out, st = simulate_rare(p, method=method, return_stats=True)
        assert st.path == 1
        dev = _max_dev(out, arrs, name)
        assert dev < TOL, f"{method}: {dev:.2e} {st}"


def test_nsteps_exhaustion_raises(golden_cases):
    import dataclasses
    from quantumsimulations_b200 import simulate_rare
    from quantumsimulations_b200.engine import IntegratorError

    meta, _ = golden_cases
    p = dataclasses.replace(params_from_meta(meta["n7_d050k_center_on"]), solver_nsteps=3, steps=5)
    with pytest.raises(IntegratorError):
        simulate_rare(p)


def test_bad_grid_raises():
    from quantumsimulations_b200 import DipolarRareParams, simulate_rare

    with pytest.raises(ValueError, match="Bad time grid"):
        simulate_rare(DipolarRareParams(n_sea=6, is_spin_three_half=False, steps=1))


def test_deterministic(golden_cases):
    from quantumsimulations_b200 import simulate_rare

    meta, _ = golden_cases
    p = params_from_meta(meta["n7_d100k_center_on"])
    a = simulate_rare(p)[1]
    b = simulate_rare(p)[1]
    for k in KEYS:
        assert np.array_equal(a[k], b[k])


def _sweep_like(n_sea, delta, variant="center_on", t_final=2e-4, steps=21, **kw):
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench

    import dataclasses
    return dataclasses.replace(bench.sweep_params(n_sea, delta, variant, t_final, steps), **kw)


@pytest.mark.parametrize("n_sea,method", [(9, "dp5"), (9, "taylor"), (10, "dp5"), (11, "taylor")])
def test_boundary_sizes_vs_oracle(n_sea, method):
    """N = 10 (1024-thread on-chip CTA), N = 11 (padded to the 12-bit tile), N = 12 (single-tile tiled path)."""
    from conftest import exact_traces
    from quantumsimulations_b200 import compile_table, simulate_rare

    p = _sweep_like(n_sea, 50e3, solver_atol=1e-12, solver_rtol=1e-11)
    t, out, st = simulate_rare(p, method=method, return_stats=True)
    want = exact_traces(compile_table(p), t)
    dev = max(float(np.max(np.abs(out[k] - want[k]))) for k in KEYS)
    assert dev < TOL, f"N={n_sea+1} {method}: {dev:.2e} {st}"
    assert st.path == (0 if n_sea + 1 <= 10 else 1)


def test_large_batch_matches_single_runs():
    """A batch big enough to change the lanes-per-amplitude choice must give the same traces (to rounding)."""
    from quantumsimulations_b200 import simulate_rare, simulate_rare_batch

    dets = np.linspace(0.0, 150e3, 100)
    plist = [_sweep_like(6, d, v) for d in dets for v in ("center_off", "center_on", "shell_off")]
    res, stats = simulate_rare_batch(plist)
    assert len(res) == 300
    for i in (0, 17, 151, 299):
        t, single = simulate_rare(plist[i])
        for k in KEYS:
            assert np.max(np.abs(res[i][1][k] - single[k])) < 1e-9


def test_full_sweep_matches_reference_sweep_metrics(tmp_path):
    """C2 (CI-sized window): the engine's batched sweep vs the metrics the reference's unchanged sweep driver
    produced on exact propagation (tests/golden/sweep_small.json).  Gate 3 of SURVEY 7.3."""
    import json, os
    from conftest import GOLDEN
    from quantumsimulations_b200.sweep import run_sweep_sea_detuning

    with open(os.path.join(GOLDEN, "sweep_small.json")) as f:
        gold = json.load(f)
    gp = gold["summary"]["global_params"]
    base = run_sweep_sea_detuning(
        f_Az=gp["f_Az_Hz"], f1A=gp["f1A_Hz"], target_sea_detuning=gp["target_sea_detuning"],
        gamma_sea=gp["gamma_sea"], gamma_rare=gp["gamma_rare"], sea_detunings_Hz=gp["sea_detunings_Hz"],
        n_sea=gp["n_sea"], t_final=gp["t_final_s"], steps=gp["steps"], phi_sea=gp["phi_sea_rad"],
        phi_rare=gp["phi_rare_rad"], out_root=str(tmp_path), is_spin_three_half=False,
        solver_atol=gp["solver_atol"], solver_rtol=gp["solver_rtol"], solver_nsteps=gp["solver_nsteps"],
        solver_max_step=gp["solver_max_step"], coarse_window=gp["coarse_window"], timestamp="gpu", verbose=False)
    with open(os.path.join(base, "summary.json")) as f:
        mine = json.load(f)
    assert len(mine["sweep_results"]) == 13
    for got, want in zip(mine["sweep_results"], gold["summary"]["sweep_results"]):
        for k in want:
            if k.startswith("I_z_slope"):
                assert abs(got[k] - want[k]) < 2e-6, (k, got[k], want[k])       # traces agree to 1e-6
            elif k.startswith("t_") or k.startswith("R_"):
                assert got[k] == pytest.approx(want[k], rel=2e-3, abs=2e-4), k
            elif k.startswith("contrast"):
                # discontinuous at |t| = 1 (SW:308-309): only compare when neither t is near the gate
                ts = [want[x] for x in ("t_on_center", "t_off_center", "t_off_sea_center")]
                if all(abs(abs(t) - 1.0) > 0.01 for t in ts):
                    assert got[k] == pytest.approx(want[k], abs=2e-3), k
            else:
                assert got[k] == pytest.approx(want[k], rel=1e-9, abs=1e-12), k


@pytest.mark.parametrize("method", ["dp5", "taylor"])
def test_checkpointed_trajectory_resumes_bitwise(tmp_path, method):
    """8(f) rank 3: a tiled-path trajectory interrupted after some output times and resumed from its checkpoint file
    (state, time, controller step size) produces exactly the traces and step counts of the uninterrupted run."""
    from bench import sweep_params
    from quantumsimulations_b200 import compile_table, engine
    from quantumsimulations_b200.checkpoint import evolve_checkpointed

    table = compile_table(sweep_params(13, 50e3))      # N = 14: tiled path
    t_out = np.linspace(0.0, 4.0e-5, 9)
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, max_step=1e-5, nsteps=100000, method=method, order=12)
    with engine.Plan(table) as plan:
        want, wst = plan.evolve(t_out, opts)
    path = str(tmp_path / "traj.npz")
    with engine.Plan(table) as plan:
        part, _, done = evolve_checkpointed(plan, t_out, opts, path, every=2, stop_after=3)
    assert not done and np.array_equal(part[:4], want[:4]) and np.all(np.isnan(part[4:]))
    with engine.Plan(table) as plan:               # a fresh plan, as after a crash
        part, _, done = evolve_checkpointed(plan, t_out, opts, path, every=2, stop_after=2)
    assert not done and np.array_equal(part[:6], want[:6])
    with engine.Plan(table) as plan:
        got, gst, done = evolve_checkpointed(plan, t_out, opts, path, every=100)
    assert done and np.array_equal(got, want)
    assert (gst.n_accepted, gst.n_rejected) == (wst.n_accepted, wst.n_rejected)
    with engine.Plan(table) as plan:               # finished checkpoint: nothing left to integrate
        again, _, done = evolve_checkpointed(plan, t_out, opts, path)
    assert done and np.array_equal(again, want)
    with pytest.raises(ValueError):
        with engine.Plan(table) as plan:
            evolve_checkpointed(plan, t_out[:5], opts, path)


def test_checkpoint_refuses_another_trajectory(tmp_path):
    """A checkpoint is tied to the Hamiltonian table, the initial state and max_step / nsteps, not only to the time
    grid: resuming it with another detuning (same grid, same tolerances) must raise instead of silently continuing the
    old state under a different H."""
    import dataclasses

    from bench import sweep_params
    from quantumsimulations_b200 import compile_table, engine
    from quantumsimulations_b200.checkpoint import evolve_checkpointed

    t_out = np.linspace(0.0, 2.0e-5, 5)
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, max_step=1e-5, nsteps=100000, method="taylor", order=12)
    path = str(tmp_path / "traj.npz")
    with engine.Plan(compile_table(sweep_params(13, 50e3))) as plan:
        _, _, done = evolve_checkpointed(plan, t_out, opts, path, stop_after=2)
    assert not done
    with pytest.raises(ValueError, match="different trajectory"):
        with engine.Plan(compile_table(sweep_params(13, 62.5e3))) as plan:          # another detuning
            evolve_checkpointed(plan, t_out, opts, path)
    with pytest.raises(ValueError, match="different trajectory"):
        with engine.Plan(compile_table(sweep_params(13, 50e3))) as plan:            # another max_step
            evolve_checkpointed(plan, t_out, dataclasses.replace(opts, max_step=2e-5), path)
    with engine.Plan(compile_table(sweep_params(13, 50e3))) as plan:                # the right one still resumes
        _, _, done = evolve_checkpointed(plan, t_out, opts, path)
    assert done


def test_sharded_checkpoint_ranks_must_agree(tmp_path):
    """Sharded resume: every rank writes its own file; ranks whose files were taken at different output times (a crash
    between two ranks' writes) must refuse to resume instead of starting at different times."""
    import shutil

    from bench import sweep_params
    from quantumsimulations_b200 import compile_table, engine, shard
    from quantumsimulations_b200.checkpoint import evolve_checkpointed

    table = compile_table(sweep_params(13, 50e3))
    t_out = np.linspace(0.0, 2.0e-5, 6)
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, max_step=1e-5, nsteps=100000, method="taylor", order=12)
    with engine.Plan(table) as plan:
        want, _ = plan.evolve(t_out, opts)
    paths = [str(tmp_path / f"r{r}.npz") for r in range(2)]
    with shard.LocalShardGroup(table, 2) as grp:
        grp.run(lambda plan, r: evolve_checkpointed(plan, t_out, opts, paths[r], stop_after=2))
        shutil.copy(paths[1], str(tmp_path / "r1_at2.npz"))
        grp.run(lambda plan, r: evolve_checkpointed(plan, t_out, opts, paths[r], stop_after=1))
        good = [str(tmp_path / f"good{r}.npz") for r in range(2)]
        for r in range(2):
            shutil.copy(paths[r], good[r])
        shutil.copy(str(tmp_path / "r1_at2.npz"), paths[1])              # rank 1 is one output time behind
        with pytest.raises(ValueError, match="different output times"):
            grp.run(lambda plan, r: evolve_checkpointed(plan, t_out, opts, paths[r]))
        res = grp.run(lambda plan, r: evolve_checkpointed(plan, t_out, opts, good[r]))
    assert res[0][2] and np.max(np.abs(res[0][0] - want)) < 1e-11


def test_dense_output_of_taylor_steps(golden_cases, monkeypatch):
    """Tiled path, output grid finer than the controller's step (C1 (ii): N = 13, 2000 outputs over 0.02 s): ONE
    unclipped Taylor step serves every output time inside it from its own polynomial.  Same traces as stepping onto
    every output (and as exact diagonalisation, on the golden's coarser grid), for a fraction of the H-applies."""
    import dataclasses
    from quantumsimulations_b200 import simulate_rare

    meta, arrs = golden_cases
    name = "n13_dataclass_default"
    base = params_from_meta(meta[name])
    assert base.steps == 201
    p = dataclasses.replace(base, steps=2001, solver_atol=1e-11, solver_rtol=1e-10, solver_nsteps=100000)
    t, dense, sd = simulate_rare(p, method="taylor", return_stats=True)
    monkeypatch.setenv("QSB200_DENSE_OUTPUT", "0")
    _, stepped, ss = simulate_rare(p, method="taylor", return_stats=True)
    assert sd.path == 1 and ss.path == 1
    assert ss.n_accepted >= 2000                       # stepping lands on every output time
    assert sd.n_accepted + sd.n_rejected <= ss.n_accepted // 3, (sd, ss)
    for k in KEYS:
        assert np.max(np.abs(dense[k] - stepped[k])) < 1e-8, k
        assert np.max(np.abs(dense[k][::10] - arrs[f"{name}/{k}"])) < TOL, k
    assert np.max(np.abs(dense["state_norm"] - 1.0)) < 1e-12
