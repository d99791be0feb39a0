"""GPU parity at the sizes the benchmark runs (BASELINE configs C4 = N 28, and the C5 shard size per GPU).

Three independent checks of the tiled path's H-apply at N = 28 on the C4 table (reference operator
dipolar_ensemble_with_rare.py:549-568 on the Fibonacci shell :232-247), all through the C ABI:
  * a random dense 2^28 state against the C restatement of the CSR row product (oracle/c) on row blocks chosen so that
    every layout group's bits (low 12, [12,20), [20,28)) take several values;
  * a product state against the O(n^2)-per-row closed form (oracle/product.py) on sampled rows;
  * <psi|H psi> real, and energy / norm conserved over 20 DP5 step attempts.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N_BIG = 28


def _c4_table(n=N_BIG):
    from bench import sweep_params
    from quantumsimulations_b200 import compile_table

    return compile_table(sweep_params(n - 1, 50e3))


def _sample_rows(dim, count, seed):
    rng = np.random.default_rng(seed)
    rows = rng.integers(0, dim, size=count, dtype=np.uint64)
    rows[:4] = (0, 1, dim - 1, dim - 2)
    return rows


@pytest.mark.parametrize("n", [14, 19, 22])
def test_product_state_rows_small_sizes(n):
    """The product-state hook against the closed form AND the numpy oracle at sizes both reach."""
    from conftest import random_table
    from oracle import dense, product
    from quantumsimulations_b200 import engine

    table = random_table(n, 900 + n)
    bm = table.bit_major()
    phi = product.random_phi(n, n)
    rows = _sample_rows(1 << n, 4096, n)
    want = product.apply_h_rows(bm["h"], bm["J"], phi, rows)
    with engine.Plan(table) as plan:
        got = plan.apply_h_product(phi, rows)
    scale = np.max(np.abs(want))
    assert np.max(np.abs(got - want)) / scale < 1e-12
    if n <= 19:
        full = dense.apply_h(bm["h"], bm["J"], product.full_state(phi))
        assert np.max(np.abs(got - full[rows.astype(np.int64)])) / scale < 1e-12


def test_n28_product_state_rows_vs_closed_form():
    from oracle import product
    from quantumsimulations_b200 import engine

    table = _c4_table()
    assert engine.tile_layouts(N_BIG) == [(12, 12, 12), (4, 12, 20), (4, 20, 28)]
    bm = table.bit_major()
    phi = product.random_phi(N_BIG, 28)
    rows = _sample_rows(1 << N_BIG, 1 << 16, 5)
    want = product.apply_h_rows(bm["h"], bm["J"], phi, rows)
    with engine.Plan(table) as plan:
        got = plan.apply_h_product(phi, rows)
    err = np.max(np.abs(got - want)) / np.max(np.abs(want))
    assert err < 1e-12, err


def test_n28_random_state_row_blocks_vs_c_port():
    from oracle.c import build as oc
    from quantumsimulations_b200 import engine

    table = _c4_table()
    bm = table.bit_major()
    n = N_BIG
    rng = np.random.default_rng(2828)
    psi = np.empty(1 << n, dtype=np.complex128)
    chunk = 1 << 24
    for c in range(0, 1 << n, chunk):   # standard normals in 268 MB pieces (keeps the transient memory small)
        psi[c:c + chunk] = rng.standard_normal(2 * chunk).view(np.complex128)
    psi *= 1.0 / np.sqrt(2.0 * (1 << n))
    with engine.Plan(table) as plan:
        got = plan.apply_h(psi)
    # <psi|H psi> is real for a Hermitian H
    e = np.vdot(psi, got)
    hnorm = np.linalg.norm(got)
    assert abs(e.imag) < 1e-12 * hnorm
    # 8 blocks of 2^18 rows: bits [0,18) run inside a block; the block offsets give bits [18,28) 8 different patterns
    blocks = [0, 0b1111111111, 0b1010101010, 0b0101010101, 0b1100110011, 0b0011001100, 0b1000000001, 0b0111111110]
    worst = 0.0
    for hi in blocks:
        row0 = hi << 18
        want = oc.apply_h(bm["h"], bm["J"], psi, row0, row0 + (1 << 18))
        err = np.max(np.abs(got[row0:row0 + (1 << 18)] - want)) / np.max(np.abs(want))
        worst = max(worst, err)
    assert worst < 1e-12, worst


def test_n28_dp5_attempts_conserve_energy_and_norm():
    """20 DP5 step attempts at the benchmark's own settings: the raw norm and <H> of the state stay where they
    started (basis state: <H> = d(x0), closed form), and <psi|H psi> is real."""
    from oracle import product
    from quantumsimulations_b200 import engine

    table = _c4_table()
    bm = table.bit_major()
    n = N_BIG
    basis = np.zeros((n, 2), dtype=np.complex128)   # |x0> as a product state: its diagonal element is <H>(0)
    for b in range(n):
        basis[b, (table.psi0_index >> b) & 1] = 1.0
    # closed-form diagonal element d(x0) (the flip terms of a basis state leave the row)
    s = 1.0 - 2.0 * np.array([(table.psi0_index >> b) & 1 for b in range(n)], dtype=np.float64)
    e0 = 0.5 * float(bm["h"][2] @ s) + 0.25 * float(s @ np.triu(bm["J"][2], 1) @ s)
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, max_step=1e-5, nsteps=0, method="dp5")
    with engine.Plan(table) as plan:
        plan.session_begin(opts)
        _, st, _ = plan.session_step(20)
        obs = plan.session_observe()
        psi = plan.session_state()
        plan.session_end()
        hpsi = plan.apply_h(psi)
    assert st.n_accepted + st.n_rejected == 20 and st.n_accepted >= 15
    nrm2 = float(np.vdot(psi, psi).real)
    assert abs(np.sqrt(nrm2) - obs[7]) < 1e-12
    # DP5 at atol 1e-10 / rtol 1e-9 in the RMS norm over 2^28 components: the raw norm drifts by ~2e-7 per attempt
    # (bench.py prints the same figure); a wrong operator or a lost tile would show up at O(1)
    assert abs(nrm2 - 1.0) < 2e-5
    e = np.vdot(psi, hpsi)
    scale = np.linalg.norm(hpsi)
    assert abs(e.imag) < 1e-12 * scale
    assert abs(e.real / nrm2 - e0) < 2e-5 * scale, (e.real / nrm2, e0, scale)
    del product


@pytest.mark.parametrize("world", [2, 8])
def test_sharded_product_state_rows(world):
    """The same closed-form check through the sharded plans (ranks = threads of this process), every rank sampling
    rows of its own slice: a peer mapping that permuted slabs would keep the norm but not these rows."""
    from conftest import random_table
    from oracle import product
    from quantumsimulations_b200 import shard

    n = 20
    table = random_table(n, 950 + world)
    bm = table.bit_major()
    phi = product.random_phi(n, world)
    with shard.LocalShardGroup(table, world) as grp:
        def work(plan, r):
            rows = _sample_rows(plan.local_dim, 2048, 10 + r)
            return rows, plan.apply_h_product(phi, rows)
        res = grp.run(work)
        ldim = grp.plans[0].local_dim
    for r, (rows, got) in enumerate(res):
        want = product.apply_h_rows(bm["h"], bm["J"], phi, rows + np.uint64(r * ldim))
        assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < 1e-12, r


@pytest.mark.parametrize("n,groups", [(21, "9"), (22, "10"), (23, "11"), (24, "12"), (24, "10,2"), (25, "11,2"),
                                      (26, "12,2"), (23, "3,8"), (26, "9,5"), (26, "5,9")])
def test_explicit_group_layouts(n, groups, monkeypatch):
    """QSB200_GROUPS: groups wider than 8 bits (tiles of 128- / 64- / 32- / 16-byte runs, the short-run passes timed
    for DESIGN.md section 10) go through the padded window-[0,4) code of pass_mid / pass_centre."""
    from conftest import random_table
    from oracle import product
    from quantumsimulations_b200 import engine

    monkeypatch.setenv("QSB200_GROUPS", groups)
    sizes = [int(x) for x in groups.split(",")]
    lays = engine.tile_layouts(n)
    assert [ge - gs for _, gs, ge in lays[1:]] == sizes and all(lk == 12 - (ge - gs) for lk, gs, ge in lays[1:])
    table = random_table(n, 1000 + n, dense_pairs=False)
    bm = table.bit_major()
    phi = product.random_phi(n, n)
    rows = _sample_rows(1 << n, 8192, n)
    want = product.apply_h_rows(bm["h"], bm["J"], phi, rows)
    with engine.Plan(table) as plan:
        got = plan.apply_h_product(phi, rows)
    assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < 1e-12


@pytest.mark.parametrize("n", [21, 29])
def test_nine_bit_group_layouts_product_rows(n):
    """Unsharded plans take a 9-bit group where that saves a layout (N = 21: one strided layout; N = 29: 12 | 9 | 8
    instead of four layouts).  Product-state rows against the closed form at those sizes (N = 29: 8.6 GB vectors)."""
    from oracle import product
    from quantumsimulations_b200 import engine

    lays = engine.tile_layouts(n)
    assert len(lays) == (2 if n == 21 else 3) and max(ge - gs for _, gs, ge in lays[1:]) == 9
    table = _c4_table(n)
    bm = table.bit_major()
    phi = product.random_phi(n, n)
    rows = _sample_rows(1 << n, 1 << 14, n)
    want = product.apply_h_rows(bm["h"], bm["J"], phi, rows)
    with engine.Plan(table) as plan:
        got = plan.apply_h_product(phi, rows)
    assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < 1e-12
