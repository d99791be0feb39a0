"""Local dimension 4 on the GPU: the J = 3/2 rare spin of DE:22-25 / DE:499-501 on two index bits.

Fixtures: tests/golden/spin32_cases.* - traces the reference's own simulate_rare produced with DE:486 repaired
(oracle/make_goldens_spin32.py).  The time evolution runs through the spin-1/2 kernels (rotated frame); the
observables through qsb_propagate_expect (n_sea + 2 <= 10 bits) or qsb_session_expect (tiled path).
"""
import numpy as np
import pytest

from conftest import params_from_meta, random_table

pytestmark = pytest.mark.gpu
KEYS = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")


def standard_terms(table):
    """The six mask-based observables of an ordinary (spin-1/2) table as weighted Pauli strings."""
    terms = []
    for b in range(table.n):
        m = 1 << b
        if table.sea_mask & m:
            terms += [(m, 0, 0.5, 0), (m, m, 0.5, 1), (0, m, 0.5, 2)]
    r = table.rare_mask
    return terms + [(0, r, 0.5, 3), (r, 0, 0.5, 4), (r, r, 0.5, 5)]


@pytest.mark.parametrize("n", [3, 6, 7, 10])
def test_string_observables_equal_the_fused_ones_on_a_spin_half_table(n):
    """qsb_propagate_expect against qsb_propagate_batch on a random XYZ table: same U, same chain, the six standard
    observables once as Pauli strings on the stored states and once as the mask-based sums fused into the chain."""
    from quantumsimulations_b200 import engine

    table = random_table(n, 4100 + n)
    opts = engine.SolverOptions(atol=1e-12, rtol=1e-11, method="taylor", order=20, nsteps=0)
    n_out, dt = 200, 0.05      # < 256 outputs: neither chain is time-split, so the states are bitwise the same
    want, st_w = engine.propagate_batch([table], n_out, dt, [opts])
    got, st_g = engine.propagate_expect(table, n_out, dt, opts, terms=standard_terms(table))
    dev = float(np.max(np.abs(got - want[0])))
    print(f"n={n}: strings vs fused sums max dev {dev:.2e}; path {st_g.path}, rhs {st_g.n_rhs} vs {st_w[0].n_rhs}")
    assert dev < 1e-11 and st_g.path == 4 and st_g.n_rhs == st_w[0].n_rhs
    by_col, _ = engine.propagate_expect(table, n_out, dt, opts, terms=standard_terms(table), by_column=True)
    assert np.array_equal(by_col, got.T)
    assert np.array_equal(got[:, 6], np.ones(n_out)) and abs(got[0, 7] - 1.0) < 1e-15


@pytest.mark.parametrize("n", [11, 13])
def test_session_expect_equals_session_observe(n):
    from quantumsimulations_b200 import engine

    table = random_table(n, 4200 + n, dense_pairs=False)
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, method="taylor", order=12)
    rng = np.random.default_rng(n)
    psi = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    with engine.Plan(table) as plan:
        plan.session_begin(opts, psi0=psi)
        for k in range(2):
            want = plan.session_observe()
            got = plan.session_expect(standard_terms(table))
            dev = float(np.max(np.abs(got - want) / np.maximum(1.0, np.abs(want))))
            print(f"n={n} step {k}: session_expect vs session_observe rel dev {dev:.2e}")
            assert dev < 1e-12
            plan.session_advance(1e-3 * (k + 1))
        plan.session_end()


@pytest.mark.parametrize("name", ["s32_n3_default_opts", "s32_n5_phases_detuned", "s32_n6_center_on", "s32_n6_center_off",
                                  "s32_n8_cube", "s32_n9_tiled", "s32_n10_tiled", "s32_dataclass_default"])
def test_simulate_rare_spin_three_half_vs_repaired_reference(golden_spin32, name):
    """simulate_rare(is_spin_three_half=True) against the reference's own traces, <= 1e-6 on all six (the bar
    north_star states for the spin-1/2 trajectories)."""
    import quantumsimulations_b200 as q

    meta, gold = golden_spin32
    p = params_from_meta(meta[name])
    assert p.is_spin_three_half
    t, out, st = q.simulate_rare(p, return_stats=True, raw_norm=True)
    assert np.array_equal(t, gold[f"{name}/t"])
    devs = {k: float(np.max(np.abs(out[k] - gold[f"{name}/{k}"]))) for k in KEYS}
    print(f"{name}: path {st.path}, {st.n_accepted} steps, {st.n_rhs} H-applies, max dev {max(devs.values()):.2e} {devs}")
    assert max(devs.values()) < 1e-6, devs
    assert st.path == (4 if p.n_sea + 2 <= 10 else 1)
    assert out["state_norm"].shape == gold[f"{name}/state_norm"].shape
    assert np.max(np.abs(out["raw_norm"] - 1.0)) < 1e-6


def test_dp5_columns_and_a_mixed_batch(golden_spin32, golden_cases):
    """method="dp5" integrates the propagator's columns with DP5; simulate_rare_batch takes spin-3/2 records next to
    spin-1/2 ones."""
    import quantumsimulations_b200 as q

    meta, gold = golden_spin32
    p3 = params_from_meta(meta["s32_n3_default_opts"])
    _, out = q.simulate_rare(p3, method="dp5", tol_scale=1e-4)
    dev = max(float(np.max(np.abs(out[k] - gold[f"s32_n3_default_opts/{k}"]))) for k in KEYS)
    print(f"dp5 columns: max dev {dev:.2e}")
    assert dev < 1e-6
    meta_h, gold_h = golden_cases
    half = params_from_meta(meta_h["n5_tetra"])
    p5 = params_from_meta(meta["s32_n5_phases_detuned"])
    res, stats = q.simulate_rare_batch([half, p5, p3])
    for (name, g), (_, o) in zip((("n5_tetra", gold_h), ("s32_n5_phases_detuned", gold), ("s32_n3_default_opts", gold)), res):
        dev = max(float(np.max(np.abs(o[k] - g[f"{name}/{k}"]))) for k in KEYS)
        print(f"batch {name}: max dev {dev:.2e}")
        assert dev < 1e-6
    assert [s.path for s in stats] == [2, 4, 4]


def test_lab_frame_h_apply_of_a_spin_three_half_operator():
    """SpinOperator.apply of a local-dimension-4 H (GPU H-apply in the rotated frame, diagonal frame change on the
    host) against the literal Kronecker oracle with the 4 x 4 J matrices."""
    import quantumsimulations_b200 as q
    from oracle import kron

    p = q.DipolarRareParams(n_sea=5, phi_rare=2.3, phi_sea=-0.4, drive_sea=True, drive_rare=True, B1_sea=1e-3, B1_rare=2e-3,
                            omega_rf_sea=8.1e7 * 2.999, omega_rf_rare=6.9e7 * 3.001)
    H, _ = q.build_hamiltonian_rare(p)
    Hk, _, _ = kron.build_spin32(p, q.get_derived_frequencies,
                                 lambda n, r: q.shell_positions_with_rare_center(n_sea=n, radius=r),
                                 q.dipolar_couplings_from_positions)
    rng = np.random.default_rng(2)
    psi = rng.normal(size=H.shape[0]) + 1j * rng.normal(size=H.shape[0])
    got, want = H.apply(psi), Hk @ psi
    rel = float(np.max(np.abs(got - want)) / np.max(np.abs(want)))
    print(f"lab-frame H-apply rel err {rel:.2e}")
    assert rel < 1e-13


@pytest.mark.parametrize("n", [5, 7, 8])
def test_time_split_state_storing_chain(n, monkeypatch):
    """>= 256 outputs: qsb_propagate_expect cuts the chain into segments seeded with powers of U^seg_len, like
    qsb_propagate_batch.  Against that call (same segments: the states are the same numbers) and against its own
    unsplit chain (QSB200_PROP_SEGMENT=0: one CTA walks all outputs)."""
    from quantumsimulations_b200 import engine

    table = random_table(n, 4300 + n)
    opts = engine.SolverOptions(atol=1e-12, rtol=1e-11, method="taylor", order=20, nsteps=0)
    n_out, dt = 1500, 0.01
    want, _ = engine.propagate_batch([table], n_out, dt, [opts])
    got, st = engine.propagate_expect(table, n_out, dt, opts, terms=standard_terms(table))
    monkeypatch.setenv("QSB200_PROP_SEGMENT", "0")
    flat, _ = engine.propagate_expect(table, n_out, dt, opts, terms=standard_terms(table))
    d_split, d_flat = float(np.max(np.abs(got - want[0]))), float(np.max(np.abs(got - flat)))
    print(f"n={n}: split strings vs split fused {d_split:.2e}, split vs unsplit {d_flat:.2e}, {st.device_ms:.2f} ms")
    assert d_split < 1e-12 and d_flat < 1e-10


def test_sweep_driver_with_a_spin_three_half_rare_spin(tmp_path):
    """run_sweep_sea_detuning(is_spin_three_half=True) (SW:370, SW:651) at the function-default grid against the
    metrics the reference's own unmodified driver produced on top of its model with the dims repaired
    (tests/golden/sweep_small_spin32.json): 26 trajectories with the four-level rare site + 13 sea-centre controls."""
    import json
    import os

    from conftest import GOLDEN
    from quantumsimulations_b200.sweep import run_sweep_sea_detuning

    with open(os.path.join(GOLDEN, "sweep_small_spin32.json")) as f:
        gold = json.load(f)
    gp = gold["summary"]["global_params"]
    assert gp["rare_spin_type"] == "3/2"
    base = run_sweep_sea_detuning(
        f_Az=gp["f_Az_Hz"], f1A=gp["f1A_Hz"], target_sea_detuning=gp["target_sea_detuning"],
        gamma_sea=gp["gamma_sea"], gamma_rare=gp["gamma_rare"], sea_detunings_Hz=gp["sea_detunings_Hz"],
        n_sea=gp["n_sea"], t_final=gp["t_final_s"], steps=gp["steps"], phi_sea=gp["phi_sea_rad"],
        phi_rare=gp["phi_rare_rad"], out_root=str(tmp_path), is_spin_three_half=True,
        solver_atol=gp["solver_atol"], solver_rtol=gp["solver_rtol"], solver_nsteps=gp["solver_nsteps"],
        solver_max_step=gp["solver_max_step"], coarse_window=gp["coarse_window"], timestamp="gpu", verbose=False)
    with open(os.path.join(base, "summary.json")) as f:
        mine = json.load(f)
    assert mine["global_params"]["rare_spin_type"] == "3/2" and len(mine["sweep_results"]) == 13
    worst = 0.0
    for got, want in zip(mine["sweep_results"], gold["summary"]["sweep_results"]):
        for k in want:
            if k.startswith("I_z_slope"):
                worst = max(worst, abs(got[k] - want[k]))
                assert abs(got[k] - want[k]) < 2e-6, (k, got[k], want[k])       # traces agree to 1e-6
            elif k.startswith("t_") or k.startswith("R_"):
                assert got[k] == pytest.approx(want[k], rel=2e-3, abs=2e-4), k
            elif k.startswith("contrast"):
                ts = [want[x] for x in ("t_on_center", "t_off_center", "t_off_sea_center")]
                if all(abs(abs(t) - 1.0) > 0.01 for t in ts):                   # discontinuous at |t| = 1 (SW:308-309)
                    assert got[k] == pytest.approx(want[k], abs=2e-3), k
            else:
                assert got[k] == pytest.approx(want[k], rel=1e-9, abs=1e-12), k
    print(f"spin-3/2 sweep: worst slope deviation {worst:.2e}")


def test_full_size_spin_three_half_trajectory_vs_exact():
    """The reference's full trajectory size (30 s, 20000 outputs, SW:1223-1250) with the J = 3/2 rare spin, n_sea = 6:
    all six traces against exact diagonalisation (oracle X on the compiled table + Pauli strings, which the CPU suite
    pins to the repaired reference) <= 1e-6 - the bar of the spin-1/2 full-size test."""
    import dataclasses

    import quantumsimulations_b200 as q
    from bench import sweep_params
    from conftest import exact_traces

    p = dataclasses.replace(sweep_params(6, 50e3, t_final=30.0, steps=20000), is_spin_three_half=True)
    t, out, st = q.simulate_rare(p, return_stats=True, raw_norm=True)
    want = exact_traces(q.compile_table(p), t)
    devs = {k: float(np.max(np.abs(out[k] - want[k]))) for k in KEYS}
    drift = float(np.max(np.abs(out["raw_norm"] - 1.0)))
    print(f"full-size spin-3/2 trajectory: {st.device_ms:.1f} ms on the device, max dev {max(devs.values()):.2e}, norm drift {drift:.1e}")
    assert max(devs.values()) < 1e-6 and drift < 1e-6 and st.path == 4


@pytest.mark.parametrize("n_out", [40, 300])
def test_hbm_propagator_string_observables_equal_the_fused_ones(n_out):
    """qsb_propagate_hbm_expect against qsb_propagate_hbm on a random 11-bit spin-1/2 table (same U, same chain; the
    strings reduced by launch_expect on the chain buffer instead of observe_columns_kernel), unsplit and time-split."""
    from quantumsimulations_b200 import engine

    table = random_table(11, 4411, dense_pairs=False)
    opts = engine.SolverOptions(atol=1e-12, rtol=1e-11, method="taylor", order=20, nsteps=0)
    want, st_w = engine.propagate_hbm(table, n_out, 0.02, opts)
    got, st_g = engine.propagate_expect(table, n_out, 0.02, opts, terms=standard_terms(table))
    by_col, _ = engine.propagate_expect(table, n_out, 0.02, opts, terms=standard_terms(table), by_column=True)
    dev = float(np.max(np.abs(got - want)))
    print(f"n_out={n_out}: HBM strings vs fused sums max dev {dev:.2e}; path {st_g.path}, {st_g.device_ms:.1f} ms")
    assert dev < 1e-12 and st_g.path == 3 and st_g.n_rhs == st_w.n_rhs
    assert np.array_equal(by_col, got.T)


@pytest.mark.parametrize("name", ["s32_n9_tiled", "s32_n10_tiled"])
def test_spin_three_half_through_the_hbm_propagator(golden_spin32, name):
    """method="propagator" at 11 / 12 index bits: the HBM-resident propagator with the spin-3/2 strings, against the
    repaired reference's traces (the stepped session of the same records is in the test above)."""
    import quantumsimulations_b200 as q

    meta, gold = golden_spin32
    p = params_from_meta(meta[name])
    _, out, st = q.simulate_rare(p, method="propagator", return_stats=True)
    dev = max(float(np.max(np.abs(out[k] - gold[f"{name}/{k}"]))) for k in KEYS)
    print(f"{name} via the HBM propagator: path {st.path}, max dev {dev:.2e}, {st.device_ms:.1f} ms")
    assert dev < 1e-6 and st.path == 3


def test_batched_string_trajectories_equal_single_calls(monkeypatch):
    """qsb_propagate_expect_batch (what a spin-3/2 sweep runs: one pass for all trajectories that share bits, grid length
    and strings) against one qsb_propagate_expect call per trajectory: bitwise, also when the batch is cut into chunks."""
    import dataclasses

    import quantumsimulations_b200 as q
    from bench import sweep_params
    from quantumsimulations_b200 import api, engine

    plist = [dataclasses.replace(sweep_params(4, d, variant=v, t_final=6e-3, steps=400), is_spin_three_half=True)
             for d in (0.0, 40e3, 90e3) for v in ("center_on", "center_off")][:5]
    tabs = [q.compile_table(p) for p in plist]
    opts = [api._options_from(p, None)[0] for p in plist]
    dts = [p.t_final / (p.steps - 1) * (1.0 + 0.01 * i) for i, p in enumerate(plist)]
    single = np.stack([engine.propagate_expect(tb, 400, dt, o)[0] for tb, dt, o in zip(tabs, dts, opts)])
    batch, stats = engine.propagate_expect_batch(tabs, 400, dts, opts)
    by_col, _ = engine.propagate_expect_batch(tabs, 400, dts, opts, by_column=True)
    monkeypatch.setenv("QSB200_PROP_CHUNK", "2")
    chunked, _ = engine.propagate_expect_batch(tabs, 400, dts, opts)
    print(f"batch of 5: max |batch - single| = {float(np.max(np.abs(batch - single))):.1e}, {stats[0].device_ms:.2f} ms")
    assert np.array_equal(batch, single) and np.array_equal(chunked, single)
    assert np.array_equal(by_col, np.swapaxes(single, 1, 2)) and [s.path for s in stats] == [4] * 5
