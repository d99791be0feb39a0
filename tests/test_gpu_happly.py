"""GPU parity of one H-apply (through the C ABI) against the numpy oracle of the table semantics."""
import os

import numpy as np
import pytest

from conftest import params_from_meta, random_table

pytestmark = pytest.mark.gpu


def _rand_state(n, seed):
    rng = np.random.default_rng(seed)
    v = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    return v / np.linalg.norm(v)


def _check(table, seed, tol=2e-13):
    from oracle import dense
    from quantumsimulations_b200 import engine

    bm = table.bit_major()
    psi = _rand_state(table.n, seed)
    want = dense.apply_h(bm["h"], bm["J"], psi)
    with engine.Plan(table) as plan:
        got = plan.apply_h(psi)
    scale = np.max(np.abs(want))
    err = np.max(np.abs(got - want)) / scale
    assert err < tol, f"n={table.n}: rel err {err:.3e}"


@pytest.mark.parametrize("n", [1, 2, 3, 5, 7, 8, 10])
def test_small_path_random_table(n):
    _check(random_table(n, 100 + n), n)


@pytest.mark.parametrize("n", [11, 12, 13, 14, 16, 18])
def test_tiled_path_random_table(n):
    _check(random_table(n, 200 + n), n)


@pytest.mark.parametrize("n,gmax", [(16, 2), (17, 3), (18, 2), (19, 4), (21, 8), (21, 5)])
def test_tiled_path_multi_group(n, gmax, monkeypatch):
    """Force several bit groups (mid passes + a short centre group) at sizes the oracle can check."""
    monkeypatch.setenv("QSB200_MAX_GROUP", str(gmax))
    _check(random_table(n, 300 + n + gmax, dense_pairs=False), n)


@pytest.mark.parametrize("n,gmax", [(16, 2), (18, 2), (19, 4), (21, 8), (22, 5)])
def test_fused_strided_passes_are_bitwise_the_unfused_ones(n, gmax, monkeypatch):
    """The L2-resident persistent launch (mid .. centre .. mid) does the per-tile arithmetic of the separate passes."""
    from quantumsimulations_b200 import engine

    monkeypatch.setenv("QSB200_MAX_GROUP", str(gmax))
    monkeypatch.setenv("QSB200_FUSE_LB", "20")
    assert engine.strided_fusion(n) is not None
    table = random_table(n, 700 + n, dense_pairs=False)
    psi = _rand_state(n, n)
    with engine.Plan(table) as plan:
        fused = plan.apply_h(psi)
    monkeypatch.setenv("QSB200_FUSE_LB", "0")
    assert engine.strided_fusion(n) is None
    with engine.Plan(table) as plan:
        plain = plan.apply_h(psi)
    assert np.array_equal(fused, plain)


def test_reference_tables(golden_cases):
    from quantumsimulations_b200 import compile_table

    meta, _ = golden_cases
    for name in ("n7_d050k_center_on", "n7_d100k_shell_off", "n9_cube", "n13_dataclass_default"):
        _check(compile_table(params_from_meta(meta[name])), 7)


def test_linearity_and_hermiticity_n22():
    """Size-independent properties at a size the oracle does not reach: <a|H b> = <H a|b>, H(a+cb) = Ha + cHb."""
    from quantumsimulations_b200 import engine

    n = 22
    table = random_table(n, 4242, dense_pairs=True)
    a, b = _rand_state(n, 1), _rand_state(n, 2)
    c = 0.3 - 0.7j
    with engine.Plan(table) as plan:
        ha, hb, hab = plan.apply_h(a), plan.apply_h(b), plan.apply_h(a + c * b)
    scale = np.linalg.norm(ha)
    assert np.linalg.norm(hab - (ha + c * hb)) / scale < 1e-13
    assert abs(np.vdot(a, hb) - np.vdot(ha, b)) / scale < 1e-12
