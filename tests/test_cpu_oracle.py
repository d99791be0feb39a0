"""CPU tests: the oracle restatements against the reference-generated goldens (and against the reference itself
while /root/reference is present).  No GPU, no product kernels."""
import json
import os

import numpy as np
import pytest
import scipy.sparse as sp

from conftest import GOLDEN, exact_traces, params_from_meta

KEYS = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")
HAVE_REF = os.path.isfile("/root/reference/dipolar_ensemble_with_rare.py")
needs_ref = pytest.mark.skipif(not HAVE_REF, reason="/root/reference not present (GPU box)")


def _golden_csr(hams, name):
    dim = len(hams[f"{name}/indptr"]) - 1
    return sp.csr_matrix((hams[f"{name}/data"], hams[f"{name}/indices"], hams[f"{name}/indptr"]), shape=(dim, dim))


def _names_with_h(hams):
    return sorted({k.split("/")[0] for k in hams.files})


def test_kron_restatement_matches_reference_built_hamiltonians(golden_cases, golden_hams):
    """oracle.kron (K) vs the CSR the reference's own build_hamiltonian_rare produced (goldens)."""
    from oracle import kron
    from quantumsimulations_b200 import model

    meta, _ = golden_cases
    for name in _names_with_h(golden_hams):
        p = params_from_meta(meta[name])
        H, obs, psi0 = kron.build(p, model.get_derived_frequencies, model.shell_positions_with_rare_center,
                                  model.dipolar_couplings_from_positions)
        G = _golden_csr(golden_hams, name)
        scale = np.abs(G.data).max()
        assert abs(H - G).max() / scale < 1e-14, name
        assert int(np.argmax(np.abs(psi0))) == int(golden_hams[f"{name}/psi0_index"]), name


def test_table_semantics_match_reference_built_hamiltonians(golden_cases, golden_hams):
    """Product table compiler + oracle.dense (index arithmetic) vs the reference-built CSR."""
    from oracle import dense
    from quantumsimulations_b200 import compile_table

    meta, _ = golden_cases
    for name in _names_with_h(golden_hams):
        t = compile_table(params_from_meta(meta[name]))
        bm = t.bit_major()
        H = dense.csr_from_table(bm["h"], bm["J"])
        G = _golden_csr(golden_hams, name)
        assert abs(H - G).max() / np.abs(G.data).max() < 1e-14, name
        assert t.psi0_index == int(golden_hams[f"{name}/psi0_index"])
        # matrix-free apply agrees with the CSR
        rng = np.random.default_rng(1)
        v = rng.normal(size=t.dim) + 1j * rng.normal(size=t.dim)
        assert np.max(np.abs(dense.apply_h(bm["h"], bm["J"], v) - G @ v)) / np.abs(G.data).max() < 1e-12


@pytest.mark.parametrize("name", ["n7_d050k_center_on", "n7_d150k_shell_off", "n5_tetra", "n4_fib3_default_opts",
                                  "n7_spin_up_phases"])
def test_exact_solver_matches_golden_traces(golden_cases, name):
    """Oracle X on the product-compiled table vs traces the reference's simulate_rare returned (exact backend)."""
    from quantumsimulations_b200 import compile_table

    meta, arrs = golden_cases
    p = params_from_meta(meta[name])
    t = np.linspace(0.0, p.t_final, p.steps)
    assert np.array_equal(t, arrs[f"{name}/t"])
    got = exact_traces(compile_table(p), t[:400])
    for k in KEYS:
        assert np.max(np.abs(got[k] - arrs[f"{name}/{k}"][:400])) < 2e-10, (name, k)


def test_zvode_restatement_matches_golden_q_case(golden_cases):
    """Oracle Q (SciPy zvode-Adams driven like QuTiP) reproduces the golden produced through the reference's
    simulate_rare + fake_qutip at the sweep's tolerances."""
    from oracle import dense, solvers
    from quantumsimulations_b200 import compile_table

    meta, arrs = golden_cases
    name = "q_n7_d050k_center_on_short"
    p = params_from_meta(meta[name])
    table = compile_table(p)
    bm = table.bit_major()
    H = dense.csr_from_table(bm["h"], bm["J"])
    psi0 = np.zeros(table.dim, complex)
    psi0[table.psi0_index] = 1.0
    t = np.linspace(0.0, p.t_final, p.steps)
    # Iz_sea operator as a diagonal
    x = np.arange(table.dim)
    iz = sp.diags(0.5 * sum(1.0 - 2.0 * ((x >> b) & 1) for b in range(table.n) if table.sea_mask >> b & 1)).tocsr()
    traces, raw, _, st = solvers.evolve_csr(H, psi0, t, [iz], backend="zvode", atol=p.solver_atol, rtol=p.solver_rtol,
                                            nsteps=p.solver_nsteps, max_step=p.solver_max_step)
    assert np.max(np.abs(traces[0].real - arrs[f"{name}/Iz_sea"])) < 1e-8
    assert st["n_steps"] > 100


def test_metrics_restatement_reproduces_golden_sweep_rows(golden_cases):
    """oracle.metrics on the golden exact traces vs the rows the reference's sweep driver wrote (window 50)."""
    from oracle import metrics as om

    _, arrs = golden_cases
    with open(os.path.join(GOLDEN, "sweep_small.json")) as f:
        rows = {round(r["delta_Hz"]): r for r in json.load(f)["summary"]["sweep_results"]}
    for dk, delta in (("000k", 0), ("050k", 50000), ("100k", 100000), ("150k", 150000)):
        fit = {}
        for tag in ("center_off", "center_on", "shell_off"):
            name = f"n7_d{dk}_{tag}"
            fit[tag] = om.central_fit(*om.block_mean(arrs[f"{name}/t"], arrs[f"{name}/Iz_sea"], 50))
        row = rows[delta]
        for tag, sfx in (("center_off", "off_center"), ("center_on", "on_center"), ("shell_off", "off_sea_center")):
            assert fit[tag]["I_z_slope"] == pytest.approx(row[f"I_z_slope_{sfx}"], rel=1e-9, abs=1e-15)
            assert fit[tag]["t_value"] == pytest.approx(row[f"t_{sfx}"], rel=1e-8)
            assert fit[tag]["R_value"] == pytest.approx(row[f"R_{sfx}"], rel=1e-8)
        c = om.gated_contrast(fit["center_on"]["I_z_slope"], fit["center_off"]["I_z_slope"],
                              fit["center_on"]["t_value"], fit["center_off"]["t_value"])
        assert c == pytest.approx(row["contrast_rare_center"], rel=1e-9, abs=1e-15)


def test_survey_known_answers(golden_cases):
    """Known answers quoted in SURVEY.md 8(c) (regenerated here through the reference + exact propagation)."""
    _, arrs = golden_cases
    assert arrs["n7_d000k_center_on/Iz_sea"][:2] == pytest.approx([-3.0, -0.006917134766508], abs=1e-12)
    assert arrs["n7_d050k_center_on/Iz_sea"][1:4] == pytest.approx(
        [-2.890180450357937, -2.576824163560264, -2.105864137366082], abs=1e-12)
    assert arrs["n7_d100k_center_on/Iz_sea"][1] == pytest.approx(-2.137389662912731, abs=1e-12)


def test_c_port_matches_numpy_semantics():
    """oracle/c (the timed CPU baseline) against oracle.dense on a random dense table."""
    from conftest import random_table
    from oracle import dense
    from oracle.c import build as oc

    t = random_table(12, 5)
    bm = t.bit_major()
    rng = np.random.default_rng(2)
    v = rng.normal(size=t.dim) + 1j * rng.normal(size=t.dim)
    want = dense.apply_h(bm["h"], bm["J"], v)
    got = oc.apply_h(bm["h"], bm["J"], v)
    assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < 1e-13
    part = oc.apply_h(bm["h"], bm["J"], v, 1024, 2048)
    assert np.array_equal(part, got[1024:2048])


@needs_ref
def test_restatements_against_live_reference():
    """While the reference is mounted: model helpers bit-exact, K/table vs its builder, metrics vs its helpers."""
    from oracle import dense, kron, metrics as om, ref_loader
    from quantumsimulations_b200 import compile_table, model

    ref = ref_loader.load_reference_model("exact")
    for n in (1, 3, 4, 6, 7, 8, 12, 13, 20, 27):
        pr = ref.shell_positions_with_rare_center(n, 0.282393e-9)
        assert np.array_equal(pr, model.shell_positions_with_rare_center(n, 0.282393e-9))
        assert np.array_equal(ref.dipolar_couplings_from_positions(pr, 1.05e-41, 8.1812e7, 6.976e7),
                              model.dipolar_couplings_from_positions(pr, 1.05e-41, 8.1812e7, 6.976e7))
    kw = dict(n_sea=5, gamma_sea=2.0, gamma_rare=3.0, B0_sea=1.5, B0_rare=1.5, B1_sea=0.1, B1_rare=0.2,
              omega_rf_sea=2.9, omega_rf_rare=4.4, phi_sea=0.4, phi_rare=1.3, drive_sea=True, drive_rare=True,
              is_spin_three_half=False)
    for centre in (True, False):
        pr, pm = ref.DipolarRareParams(is_center_rare=centre, **kw), model.DipolarRareParams(is_center_rare=centre, **kw)
        assert ref.get_derived_frequencies(pr) == model.get_derived_frequencies(pm)
        Hr, _ = ref.build_hamiltonian_rare(pr)
        bm = compile_table(pm).bit_major()
        scale = np.abs(Hr.data.data).max()
        assert abs(Hr.data - dense.csr_from_table(bm["h"], bm["J"])).max() / scale < 1e-14
        Hk, _, _ = kron.build(pm, model.get_derived_frequencies, model.shell_positions_with_rare_center,
                              model.dipolar_couplings_from_positions)
        assert abs(Hr.data - Hk).max() / scale < 1e-14
    sw = ref_loader.load_reference_sweep()
    rng = np.random.default_rng(0)
    t = np.linspace(0, 0.03, 2000)
    y = -3 + 0.1 * t * 30 + 0.05 * rng.normal(size=t.size)
    for w in (25, 50, 100, 1, 3000):
        a, b = sw.coarse_grain(t, y, w), om.block_mean(t, y, w)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    tc, yc = sw.coarse_grain(t, y, 50)
    fa, fb = sw.iz_slope_from_coarse(tc, yc), om.central_fit(tc, yc)
    assert fa == fb
    assert sw.detuning_label(-1234.56) == om.label_for(-1234.56)
    for args in ((1.0, 2.0, 3.0, 0.5), (1.0, -2.0, 0.2, 0.3), (float("nan"), 1.0, 2.0, 2.0), (2.0, 1.0, 5.0, 5.0)):
        ra, rb = sw.contrast_michelson_with_t_gate(*args), om.gated_contrast(*args)
        assert (np.isnan(ra) and np.isnan(rb)) or ra == rb


def test_exact_oracle_matches_full_size_reference_sweep_samples():
    """tests/golden/sweep_full.json: the reference's unchanged driver at its own __main__ size (30 s, 20000 outputs,
    SW:1199-1252) on exact propagation.  The oracle's table + exact solver must reproduce the stored trace samples
    of all 39 trajectories (out to t = 30 s, i.e. phases of ~10^8 rad)."""
    import json

    from conftest import GOLDEN, exact_traces
    from quantumsimulations_b200 import compile_table
    from quantumsimulations_b200.metrics import detuning_label
    from quantumsimulations_b200.sweep import sweep_point_params

    with open(os.path.join(GOLDEN, "sweep_full.json")) as f:
        gold = json.load(f)
    gp = gold["summary"]["global_params"]
    t = np.linspace(0.0, gp["t_final_s"], gp["steps"])
    assert len(gold["samples"]) == 39
    for delta in gp["sea_detunings_Hz"]:
        variants = sweep_point_params(
            n_sea=gp["n_sea"], gamma_sea=gp["gamma_sea"], gamma_rare=gp["gamma_rare"], B0_common=gp["B0_common_T"],
            B1_sea=gp["B1_sea_T"], B1_rare=gp["B1_rare_T"], omega_rf_sea=2 * np.pi * (gp["f_Az_Hz"] - delta),
            omega_rf_rare=2 * np.pi * gp["f_Rz_Hz"], phi_sea=gp["phi_sea_rad"], phi_rare=gp["phi_rare_rad"],
            dipolar_scale=gp["dipolar_scale_SI"], shell_scale=gp["shell_scale_m"], t_final=gp["t_final_s"],
            steps=gp["steps"], is_spin_three_half=False, solver_atol=gp["solver_atol"], solver_rtol=gp["solver_rtol"],
            solver_nsteps=gp["solver_nsteps"], solver_max_step=gp["solver_max_step"])
        for tag, p in variants.items():
            s = gold["samples"][f"{detuning_label(delta)}/{tag}"]
            got = exact_traces(compile_table(p), t[s["index"]])
            assert np.max(np.abs(got["Iz_sea"] - np.array(s["Iz_sea"]))) < 1e-7, (delta, tag)
            assert np.max(np.abs(got["Iz_R"] - np.array(s["Iz_R"]))) < 1e-7, (delta, tag)


@pytest.mark.parametrize("name", ["n7_d050k_center_on", "n7_d100k_shell_off", "n9_cube"])
def test_product_state_closed_form_matches_reference_built_hamiltonians(name, golden_cases, golden_hams):
    """oracle.product (the O(n^2)-per-row closed form used at N = 28 .. 33) vs the reference-built CSR acting on the
    same product state, and vs the C row product (oracle/c) on a random table."""
    from oracle import product
    from quantumsimulations_b200 import compile_table

    meta, _ = golden_cases
    t = compile_table(params_from_meta(meta[name]))
    bm = t.bit_major()
    phi = product.random_phi(t.n, 11)
    psi = product.full_state(phi)
    G = _golden_csr(golden_hams, name)
    want = G @ psi
    got = product.apply_h_rows(bm["h"], bm["J"], phi, np.arange(t.dim))
    assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < 1e-13


def test_product_state_closed_form_matches_c_port_n20():
    from conftest import random_table
    from oracle import product
    from oracle.c import build as oc

    n = 20
    t = random_table(n, 91)
    bm = t.bit_major()
    phi = product.random_phi(n, 3)
    psi = product.full_state(phi)
    row0, row1 = 777 << 8, (777 << 8) + 4096
    want = oc.apply_h(bm["h"], bm["J"], psi, row0, row1)
    got = product.apply_h_rows(bm["h"], bm["J"], phi, np.arange(row0, row1))
    assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < 1e-13
