"""Local dimension 4 (the J = 3/2 rare spin of DE:22-25 / DE:499-501), host side and oracle.

The fixtures under tests/golden/spin32_cases.* come from the reference's OWN builder / simulate_rare with the one
function DE:486 misuses repaired (oracle/make_goldens_spin32.py).  Checked here, without a GPU: the product's two-bit
encoding + rotated frame + Pauli strings, and the literal Kronecker oracle, both reproduce them.
"""
import dataclasses
import os

import numpy as np
import pytest
import scipy.sparse as sp

from conftest import exact_traces, params_from_meta

KEYS = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")


def _kron_build(p):
    import quantumsimulations_b200 as q
    from oracle import kron

    return kron.build_spin32(p, q.get_derived_frequencies,
                             lambda n, r: q.shell_positions_with_rare_center(n_sea=n, radius=r),
                             q.dipolar_couplings_from_positions)


def test_two_bit_encoding_of_the_spin_three_half_operators():
    """J_z = 2 Iz_a + Iz_b, J_x = sqrt(3) Ix_b + 2 (Ix_a Ix_b + Iy_a Iy_b), J_y = sqrt(3) Iy_b + 2 (Iy_a Ix_b - Ix_a Iy_b)
    on level m = 2 a + b - against the jmat(3/2) matrices of the oracle (DE:22-24)."""
    from oracle import kron

    jx, jy, jz = (m.toarray() for m in kron._jmat32())
    sx, sy, sz = (m.toarray() for m in (kron._SX, kron._SY, kron._SZ))
    one = np.eye(2)
    a = lambda m: np.kron(m, one)   # noqa: E731  high bit
    b = lambda m: np.kron(one, m)   # noqa: E731  low bit
    assert np.allclose(jz, 2 * a(sz) + b(sz), atol=1e-15)
    assert np.allclose(jx, np.sqrt(3) * b(sx) + 2 * (a(sx) @ b(sx) + a(sy) @ b(sy)), atol=1e-15)
    assert np.allclose(jy, np.sqrt(3) * b(sy) + 2 * (a(sy) @ b(sx) - a(sx) @ b(sy)), atol=1e-15)
    assert np.allclose(jx @ jy - jy @ jx, 1j * jz, atol=1e-14)


def test_compiled_operator_equals_the_repaired_reference(golden_spin32):
    """H (lab frame, via SpinOperator.to_csr), psi0 index and - on the smallest case - all six observable matrices
    against what the reference's build_hamiltonian_rare / initial_state_rare returned."""
    import quantumsimulations_b200 as q

    meta, gold = golden_spin32
    seen = 0
    for name in meta:
        p = params_from_meta(meta[name])
        tb = q.compile_table(p)
        assert tb.n == p.n_sea + 2 and tb.psi0_index == int(gold[f"{name}/psi0_index"])
        if f"{name}/H_data" not in gold.files:
            continue
        dim = tb.dim
        want = sp.csr_matrix((gold[f"{name}/H_data"], gold[f"{name}/H_indices"], gold[f"{name}/H_indptr"]), shape=(dim, dim))
        H, obs = q.build_hamiltonian_rare(p)
        assert H.dims == [[2] * p.n_sea + [4]] * 2 and H.shape == (dim, dim)
        diff = (H.to_csr() - want)
        assert abs(diff).max() <= 1e-14 * abs(want).max(), name
        if f"{name}/op_Ix_R" in gold.files:
            for k in KEYS:
                assert np.abs(obs[k].full() - gold[f"{name}/op_{k}"]).max() < 1e-14, (name, k)
        seen += 1
    assert seen >= 4


@pytest.mark.parametrize("name", ["s32_n3_default_opts", "s32_n5_phases_detuned", "s32_n6_center_on", "s32_n8_cube"])
def test_host_logic_and_kron_oracle_reproduce_the_reference_traces(golden_spin32, name):
    """(i) the product's table in the rotated frame, evolved exactly from its basis state and read through its Pauli
    strings (oracle/pauli.py = the arithmetic qsb_propagate_expect defines), and (ii) the literal Kronecker oracle in
    the lab frame with the 4 x 4 J matrices: both equal the reference's traces."""
    import quantumsimulations_b200 as q
    from oracle import solvers

    meta, gold = golden_spin32
    p = params_from_meta(meta[name])
    t = gold[f"{name}/t"]
    assert np.array_equal(t, np.linspace(0.0, p.t_final, p.steps))
    got = exact_traces(q.compile_table(p), t)
    H, obs, psi0 = _kron_build(p)
    traces, _, _, _ = solvers.evolve_csr(H, psi0, t, [obs[k] for k in KEYS])
    for i, k in enumerate(KEYS):
        want = gold[f"{name}/{k}"]
        assert np.max(np.abs(got[k] - want)) < 1e-9, (name, k)
        assert np.max(np.abs(traces[i].real - want)) < 1e-9, (name, k)
    # DE:669 / QuTiP store_states: no norms without solver_* overrides
    assert len(gold[f"{name}/state_norm"]) == (0 if p.solver_atol is None else p.steps)


def test_rotated_frame_is_exact_for_a_random_state():
    """<psi_lab|O|psi_lab> with psi_lab = R psi equals the Pauli-string value on psi, for all six observables."""
    import quantumsimulations_b200 as q
    from oracle import pauli

    p = q.DipolarRareParams(n_sea=4, phi_rare=2.3, phi_sea=-0.4, drive_sea=True, drive_rare=True, B1_sea=1e-3, B1_rare=2e-3)
    tb = q.compile_table(p)
    rng = np.random.default_rng(5)
    psi = rng.normal(size=tb.dim) + 1j * rng.normal(size=tb.dim)
    cols = pauli.expect_terms(psi, tb.pauli_terms)[:, 0]
    lab = tb.frame_phases() * psi
    _, obs, _ = _kron_build(p)
    for i, k in enumerate(KEYS):
        want = np.vdot(lab, obs[k] @ lab).real / np.vdot(lab, lab).real
        assert abs(cols[i] - want) < 1e-13, k
    assert cols[6] == 1.0 and abs(cols[7] - np.linalg.norm(psi)) < 1e-13


@pytest.mark.reference
def test_live_repaired_reference_agrees(tmp_path):
    """The reference itself (dims repaired, fake QuTiP) against the product's operator on a case that is NOT a fixture."""
    from oracle import ref_loader

    if not ref_loader.reference_available():
        pytest.skip("needs /root/reference")
    import quantumsimulations_b200 as q
    from oracle.make_goldens_spin32 import load_repaired_reference, restore_reference

    ref = load_repaired_reference()
    try:
        rp = ref.DipolarRareParams(n_sea=4, drive_sea=True, drive_rare=True, B1_sea=2e-3, B1_rare=1e-3, phi_sea=0.9,
                                   phi_rare=-1.7, omega_rf_sea=8.1e7 * 2.999, omega_rf_rare=6.9e7 * 3.001, init_x_sign=+1)
        H, obs = ref.build_hamiltonian_rare(rp)
        psi0 = ref.initial_state_rare(rp).full().ravel()
    finally:
        restore_reference(ref)
    p = q.DipolarRareParams(**dataclasses.asdict(rp))
    Hq, oq = q.build_hamiltonian_rare(p)
    assert np.abs(Hq.full() - H.full()).max() <= 1e-14 * np.abs(H.full()).max()
    for k in KEYS:
        assert np.abs(oq[k].full() - obs[k].full()).max() < 1e-14
    assert np.array_equal(np.abs(q.initial_state_rare(p)), np.abs(psi0))


def lib_call_is_bad_arg(engine):
    """qsb_session_expect without a plan and qsb_propagate_hbm_expect without terms: QSB_ERR_BAD_ARG (1)."""
    lib = engine.load_library()
    return lib.qsb_session_expect(None, None, 0, None) == 1 and \
        lib.qsb_propagate_hbm_expect(None, 0, 1, 1.0, None, -1, None, 0, 0, None, None, 0) == 1


def test_pauli_entry_points_validate_before_touching_a_device():
    import quantumsimulations_b200 as q
    from quantumsimulations_b200 import engine

    tb = q.compile_table(q.DipolarRareParams(n_sea=3))
    o = engine.SolverOptions(method="taylor")
    with pytest.raises(ValueError, match="above n_qubits"):
        engine.propagate_expect(tb, 5, 1e-3, o, terms=[(1 << 9, 0, 0.5, 0)])
    with pytest.raises(ValueError, match="column"):
        engine.propagate_expect(tb, 5, 1e-3, o, terms=[(1, 0, 0.5, 6)])
    with pytest.raises(ValueError, match="Pauli terms"):
        engine.propagate_expect(tb, 5, 1e-3, o, terms=[])
    with pytest.raises(ValueError, match="dt must be"):
        engine.propagate_expect(tb, 5, 0.0, o)
    big = q.compile_table(q.DipolarRareParams(n_sea=13))         # 15 bits: beyond the HBM-resident propagator too
    with pytest.raises(engine.EngineError, match="n_qubits <= 14"):
        engine.propagate_expect(big, 5, 1e-3, o)
    assert lib_call_is_bad_arg(engine)
    with pytest.raises(ValueError, match="spin-3/2"):
        q.api.resolve_method(q.DipolarRareParams(n_sea=13), "propagator")
    assert q.api.resolve_method(q.DipolarRareParams(n_sea=9), "propagator") == ("taylor", True)   # HBM-resident propagator
    assert q.api.resolve_method(q.DipolarRareParams(n_sea=9), "taylor") == ("taylor", False)
    assert q.api.resolve_method(q.DipolarRareParams(n_sea=8), "dp5") == ("dp5", True)
    assert q.api.resolve_method(q.DipolarRareParams(n_sea=9), None) == ("taylor", False)
    other_phase = q.compile_table(q.DipolarRareParams(n_sea=3, phi_rare=0.5))
    with pytest.raises(ValueError, match="ONE list of Pauli strings"):
        engine.propagate_expect_batch([tb, other_phase], 5, 1e-3, [o, o])
    with pytest.raises(ValueError, match="same number of index bits"):
        engine.propagate_expect_batch([tb, q.compile_table(q.DipolarRareParams(n_sea=4))], 5, 1e-3, [o, o], terms=tb.pauli_terms)
    if engine.device_count() == 0:
        with pytest.raises(engine.EngineError, match="no CUDA device"):
            q.simulate_rare(q.DipolarRareParams(n_sea=3))
        with pytest.raises(engine.EngineError, match="no CUDA device"):
            q.simulate_rare_batch([q.DipolarRareParams(n_sea=3), q.DipolarRareParams(n_sea=3, drive_sea=True, B1_sea=1e-3)])
