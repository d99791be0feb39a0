"""GPU parity of the constant-H propagator path (qsb_propagate_batch / qsb_build_propagator) against the exact
oracle, the reference-generated goldens and the engine's own stepping integrators."""
import dataclasses

import numpy as np
import pytest

from conftest import exact_traces, params_from_meta, random_table

pytestmark = pytest.mark.gpu
KEYS = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")
TOL = 1e-6  # north_star: <= 1e-6 absolute on the normalised magnetisation traces


def _tight(method="taylor"):
    from quantumsimulations_b200.engine import SolverOptions

    return SolverOptions(atol=1e-13, rtol=1e-12, nsteps=1000000, method=method, order=20 if method == "taylor" else 0)


def _dense_h(table):
    from oracle import dense

    bm = table.bit_major()
    return np.asarray(dense.csr_from_table(bm["h"], bm["J"]).todense())


@pytest.mark.parametrize("n", [2, 3, 5, 7, 8, 9, 10])
@pytest.mark.parametrize("method", ["taylor", "dp5"])
def test_propagator_is_exp_minus_iHdt(n, method):
    """U built column by column on the GPU vs exp(-i H dt) from a dense eigendecomposition; U must be unitary."""
    from quantumsimulations_b200 import engine

    if method == "dp5" and n > 7:
        pytest.skip("one integrator is enough for the wide tables")
    table = random_table(n, 100 + n)
    H = _dense_h(table)
    w, v = np.linalg.eigh(H)
    dt = 4.0 / max(1.0, float(np.max(np.abs(w))))
    want = (v * np.exp(-1j * w * dt)) @ v.conj().T
    U, st = engine.build_propagator(table, dt, _tight(method), max_squarings=0)
    assert st.path == 2 and st.n_rhs > 0 and st.status == 0
    assert np.max(np.abs(U - want)) < 1e-9
    assert np.max(np.abs(U.conj().T @ U - np.eye(table.dim))) < 1e-9


@pytest.mark.parametrize("n", [2, 4, 6, 7, 8, 9, 10])
def test_scaling_and_squaring(n):
    """||H|| dt ~ 300: the columns are integrated over dt / 2^s and the matrix squared s times on chip; the result must
    equal exp(-i H dt) and the propagator integrated over the whole interval."""
    from quantumsimulations_b200 import engine

    table = random_table(n, 300 + n)
    w, v = np.linalg.eigh(_dense_h(table))
    dt = 300.0 / float(np.max(np.abs(w)))
    want = (v * np.exp(-1j * w * dt)) @ v.conj().T
    s = engine.propagator_squarings(table, dt)
    assert 5 <= s <= 12
    U, st = engine.build_propagator(table, dt, _tight())
    assert np.max(np.abs(U - want)) < 1e-9
    for cap in (1, 4) if n <= 7 else ():   # odd and even depths exercise both buffer parities
        Uc, stc = engine.build_propagator(table, dt, _tight(), max_squarings=cap)
        assert np.max(np.abs(Uc - want)) < 1e-9
        assert stc.n_rhs > st.n_rhs
    if n <= 7:
        U0, st0 = engine.build_propagator(table, dt, _tight(), max_squarings=0)
        assert np.max(np.abs(U0 - U)) < 1e-10
        assert st0.n_rhs > 2 * st.n_rhs


@pytest.mark.parametrize("n", [2, 4, 6, 7, 8, 9, 10])
def test_random_tables_match_exact(n):
    """Traces of psi_k = U^k psi0 (all three axes populated, random basis state) vs oracle X."""
    from quantumsimulations_b200 import engine

    tables = [random_table(n, 7 * n + s) for s in range(3)]
    n_out, dt = 41, 0.11
    obs, stats = engine.propagate_batch(tables, n_out, dt, [_tight()] * len(tables))
    t = dt * np.arange(n_out)
    for table, o, st in zip(tables, obs, stats):
        want = exact_traces(table, t)
        for i, k in enumerate(KEYS):
            assert np.max(np.abs(o[:, i] - want[k])) < 1e-8, (n, k)
        assert np.all(o[:, 6] == 1.0)
        assert np.max(np.abs(o[:, 7] - 1.0)) < 1e-9   # raw norm: U is unitary to the integrator's tolerance
        assert st.path == 2


@pytest.mark.parametrize("method", ["propagator", "propagator-dp5"])
def test_n7_sweep_points_vs_reference_goldens(golden_cases, method):
    """The sweep's own N = 7 points (reference-built H, exact propagation) through simulate_rare_batch."""
    from quantumsimulations_b200 import simulate_rare_batch

    meta, arrs = golden_cases
    names = [k for k in meta if k.startswith("n7_d") and meta[k]["backend"] == "exact"]
    plist = [params_from_meta(meta[k]) for k in names]
    if method.endswith("dp5"):
        plist = [dataclasses.replace(p, solver_atol=1e-13, solver_rtol=1e-12) for p in plist]
    res, stats = simulate_rare_batch(plist, method=method)
    for name, (t, out), st in zip(names, res, stats):
        assert np.array_equal(t, arrs[f"{name}/t"])
        dev = max(float(np.max(np.abs(out[k] - arrs[f"{name}/{k}"]))) for k in KEYS)
        assert dev < TOL, f"{name} [{method}]: {dev:.2e} ({st})"
        assert np.all(out["state_norm"] == 1.0)
        assert st.path == 2


@pytest.mark.parametrize("name", ["n4_fib3_default_opts", "n5_tetra", "n6_fib5_shell_off", "n9_cube", "n7_spin_up_phases"])
def test_other_geometries(golden_cases, name):
    from quantumsimulations_b200 import simulate_rare

    meta, arrs = golden_cases
    p = params_from_meta(meta[name])
    t, out = simulate_rare(p, method="propagator")
    dev = max(float(np.max(np.abs(out[k] - arrs[f"{name}/{k}"]))) for k in KEYS)
    assert dev < (TOL if p.solver_atol is not None else 2e-4)
    if p.solver_atol is None:
        assert out["state_norm"].size == 0


def test_agrees_with_stepping_integrator(golden_cases):
    """Same parameters through the stepping Taylor integrator and through the propagator: <= 1e-8 apart."""
    from quantumsimulations_b200 import simulate_rare

    meta, _ = golden_cases
    p = params_from_meta(meta["n7_d050k_center_on"])
    _, a = simulate_rare(p, method="taylor")
    _, b, st = simulate_rare(p, method="propagator", return_stats=True)
    assert max(float(np.max(np.abs(a[k] - b[k]))) for k in KEYS) < 1e-8
    assert st.path == 2


def test_full_size_reference_trajectory(golden_cases):
    """One trajectory at the reference's FULL size (SW:1201-1252: t_final = 30 s, 20000 outputs, atol 1e-10,
    rtol 1e-9, max_step 1e-5, nsteps 1e7) vs oracle X; ~10^9 integrator steps for QuTiP, 128 short integrations +
    19999 matvecs here."""
    from quantumsimulations_b200 import compile_table, simulate_rare

    meta, _ = golden_cases
    p = dataclasses.replace(params_from_meta(meta["n7_d050k_center_on"]), t_final=30.0, steps=20000,
                            solver_atol=1e-10, solver_rtol=1e-9, solver_nsteps=10_000_000, solver_max_step=1e-5)
    t, out, st = simulate_rare(p, method="propagator", return_stats=True, raw_norm=True)
    want = exact_traces(compile_table(p), t)
    dev = max(float(np.max(np.abs(out[k] - want[k]))) for k in KEYS)
    assert dev < TOL, f"{dev:.2e} {st}"
    assert np.max(np.abs(out["raw_norm"] - 1.0)) < 1e-7


def test_per_trajectory_dt_and_determinism(monkeypatch):
    from quantumsimulations_b200 import engine

    tables = [random_table(6, 40 + s) for s in range(4)]
    dts = np.array([0.05, 0.1, 0.2, 0.07])
    tables[1] = dataclasses.replace(tables[1], hz=tables[1].hz * 40.0)   # different squaring depths inside one batch
    assert engine.propagator_squarings(tables[1], 0.1) > engine.propagator_squarings(tables[0], 0.05)
    a, _ = engine.propagate_batch(tables, 25, dts, [_tight()] * 4)
    b, _ = engine.propagate_batch(tables, 25, dts, [_tight()] * 4)
    assert np.array_equal(a, b)
    for chunk in ("3", "1"):   # several passes over the batch (what a grid too large for one pass gets)
        monkeypatch.setenv("QSB200_PROP_CHUNK", chunk)
        c, cst = engine.propagate_batch(tables, 25, dts, [_tight()] * 4)
        assert np.array_equal(a, c) and all(s.path == 2 and s.n_rhs > 0 for s in cst)
    monkeypatch.delenv("QSB200_PROP_CHUNK")
    cc, _ = engine.propagate_batch(tables, 25, dts, [_tight()] * 4, by_column=True)   # one contiguous trace per observable
    assert cc.shape == (4, 8, 25) and np.array_equal(cc.transpose(0, 2, 1), a)
    for i in range(4):
        one, _ = engine.propagate_batch([tables[i]], 25, float(dts[i]), [_tight()])
        assert np.array_equal(one[0], a[i])
        want = exact_traces(tables[i], dts[i] * np.arange(25))
        assert max(float(np.max(np.abs(a[i][:, j] - want[k]))) for j, k in enumerate(KEYS)) < 1e-8


def test_errors():
    from quantumsimulations_b200 import DipolarRareParams, engine, simulate_rare
    from quantumsimulations_b200.engine import IntegratorError, SolverOptions

    with pytest.raises(ValueError, match="same n_qubits"):
        engine.propagate_batch([random_table(4, 1), random_table(5, 2)], 5, 0.1, [_tight()] * 2)
    with pytest.raises(ValueError, match="dt must be"):
        engine.propagate_batch([random_table(4, 1)], 5, -1.0, [_tight()])
    with pytest.raises(IntegratorError):
        engine.propagate_batch([random_table(6, 3)], 5, 50.0, [SolverOptions(atol=1e-12, rtol=1e-11, nsteps=3)],
                               max_squarings=0)
    with pytest.raises(ValueError, match="N <= 14"):   # beyond the HBM-resident propagator (11 <= N <= 14) as well
        simulate_rare(DipolarRareParams(n_sea=14, is_spin_three_half=False), method="propagator")
    with pytest.raises(engine.EngineError):            # the C entry point refuses sizes outside its range itself
        engine.propagate_hbm(random_table(10, 4), 5, 0.1, _tight())


def test_full_size_sweep_reproduces_reference_metrics(tmp_path):
    """north_star target: the sweep at the reference's own __main__ size (SW:1199-1252: 13 detunings x 3 variants,
    30 s, 20000 outputs, coarse_window 100) through the engine's driver with the API default method, against the
    metrics rows and trace samples the reference's unchanged driver produced on exact propagation
    (tests/golden/sweep_full.json, oracle/make_golden_full_sweep.py)."""
    import json
    import os

    from conftest import GOLDEN
    from quantumsimulations_b200.sweep import run_sweep_sea_detuning

    with open(os.path.join(GOLDEN, "sweep_full.json")) as f:
        gold = json.load(f)
    gp = gold["summary"]["global_params"]
    base = run_sweep_sea_detuning(
        f_Az=gp["f_Az_Hz"], f1A=gp["f1A_Hz"], target_sea_detuning=gp["target_sea_detuning"],
        gamma_sea=gp["gamma_sea"], gamma_rare=gp["gamma_rare"], sea_detunings_Hz=gp["sea_detunings_Hz"],
        n_sea=gp["n_sea"], t_final=gp["t_final_s"], steps=gp["steps"], phi_sea=gp["phi_sea_rad"],
        phi_rare=gp["phi_rare_rad"], out_root=str(tmp_path), is_spin_three_half=False,
        solver_atol=gp["solver_atol"], solver_rtol=gp["solver_rtol"], solver_nsteps=gp["solver_nsteps"],
        solver_max_step=gp["solver_max_step"], coarse_window=gp["coarse_window"], timestamp="full", verbose=False)
    for key, s in gold["samples"].items():
        d, tag = key.split("/")
        with np.load(os.path.join(base, d, f"time_and_obs_{tag}.npz")) as z:
            assert z["t"].shape == (20000,) and z["state_norm"].shape == (20000,)
            for k in ("Iz_sea", "Iz_R"):
                assert np.max(np.abs(z[k][s["index"]] - np.array(s[k]))) < TOL, (key, k)
    with open(os.path.join(base, "summary.json")) as f:
        mine = json.load(f)
    assert len(mine["sweep_results"]) == 13
    for got, want in zip(mine["sweep_results"], gold["summary"]["sweep_results"]):
        for k in want:
            if k.startswith("I_z_slope"):
                assert abs(got[k] - want[k]) < 1e-7, (k, got[k], want[k])     # traces within 1e-6 over a 30 s baseline
            elif k.startswith("t_") or k.startswith("R_"):
                assert got[k] == pytest.approx(want[k], rel=2e-3, abs=2e-4), k
            elif k.startswith("contrast"):
                ts = [want[x] for x in ("t_on_center", "t_off_center", "t_off_sea_center")]
                if all(abs(abs(t) - 1.0) > 0.01 for t in ts):   # discontinuous at |t| = 1 (SW:308-309)
                    assert got[k] == pytest.approx(want[k], abs=2e-3), k
            else:
                assert got[k] == pytest.approx(want[k], rel=1e-9, abs=1e-12), k


@pytest.mark.parametrize("n", [3, 6, 7, 9])
def test_time_split_chain_matches_unsplit(n, monkeypatch):
    """Chains of >= 256 outputs are cut into segments that run as separate CTAs from seeds (U^L)^c psi_0.  The result
    must equal the single sequential chain to rounding, for any segment length, with a ragged last segment, in both
    observable layouts, and must not depend on the batch a trajectory is part of."""
    from quantumsimulations_b200 import engine

    tables = [random_table(n, 900 + 10 * n + s) for s in range(3)]
    n_out, dt = 1000, 0.05          # default: 64-output segments -> 15 full ones + 40 outputs
    monkeypatch.delenv("QSB200_PROP_SEGMENT", raising=False)
    a, _ = engine.propagate_batch(tables, n_out, dt, [_tight()] * 3)
    one, _ = engine.propagate_batch(tables[1:2], n_out, dt, [_tight()])
    assert np.array_equal(one[0], a[1])
    col, _ = engine.propagate_batch(tables, n_out, dt, [_tight()] * 3, by_column=True)
    assert np.array_equal(col.transpose(0, 2, 1), a)
    monkeypatch.setenv("QSB200_PROP_SEGMENT", "0")
    whole, _ = engine.propagate_batch(tables, n_out, dt, [_tight()] * 3)
    assert np.max(np.abs(a - whole)) < 1e-10 and not np.array_equal(a, whole)
    for lb in ("7", "9"):
        monkeypatch.setenv("QSB200_PROP_SEGMENT", lb)
        other, _ = engine.propagate_batch(tables, n_out, dt, [_tight()] * 3)
        assert np.max(np.abs(other - whole)) < 1e-10
    want = exact_traces(tables[0], dt * np.arange(n_out))
    assert max(float(np.max(np.abs(a[0][:, j] - want[k]))) for j, k in enumerate(KEYS)) < 1e-8


@pytest.mark.parametrize("n_sea,steps,variant", [(10, 600, "center_on"), (10, 40, "shell_off"), (11, 300, "center_on")])
def test_hbm_resident_propagator_matches_exact(n_sea, steps, variant):
    """11 <= N <= 14: U = exp(-i H dt) resident in HBM (qsb_propagate_hbm): all columns integrated at once by the tiled
    path, cuBLAS squarings and chain.  Traces against exact diagonalisation (<= 1e-6), time-split (>= 256 outputs) and
    single-segment chains, centre-rare and sea-centre variants; the raw norm stays at 1."""
    from bench import sweep_params
    from conftest import exact_traces
    from quantumsimulations_b200 import compile_table, simulate_rare

    p = sweep_params(n_sea, 50e3, variant, t_final=3.0e-2 * steps / 2000.0, steps=steps)
    t, out, st = simulate_rare(p, method="propagator", return_stats=True, raw_norm=True)
    assert st.path == 3 and st.n_rhs > 0
    want = exact_traces(compile_table(p), t)
    for k in ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R"):
        assert np.max(np.abs(out[k] - want[k])) < 1e-6, k
    assert np.max(np.abs(out["raw_norm"] - 1.0)) < 1e-7
    assert np.max(np.abs(out["state_norm"] - 1.0)) < 1e-12


@pytest.mark.parametrize("n_sea", [12, 13])
def test_hbm_propagator_agrees_with_stepping_n13_n14(n_sea):
    """N = 13 (the reference sweep function's default n_sea = 12) and N = 14 (the path's limit): a short window both ways."""
    from bench import sweep_params
    from quantumsimulations_b200 import simulate_rare

    p = sweep_params(n_sea, 50e3, t_final=3.0e-4, steps=21)
    _, a, sa = simulate_rare(p, method="propagator", return_stats=True)
    _, b, sb = simulate_rare(p, method="taylor", return_stats=True)
    assert sa.path == 3 and sb.path == 1
    for k in ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R"):
        assert np.max(np.abs(a[k] - b[k])) < 1e-7, k
