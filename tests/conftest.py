import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs /root/reference (build container only)")


@pytest.fixture(scope="session")
def golden_cases():
    with open(os.path.join(GOLDEN, "traj_cases.json")) as f:
        meta = json.load(f)
    arrs = np.load(os.path.join(GOLDEN, "traj_cases.npz"))
    return meta, arrs


@pytest.fixture(scope="session")
def golden_spin32():
    """Fixtures the reference's own code produced with its dims repaired (oracle/make_goldens_spin32.py)."""
    with open(os.path.join(GOLDEN, "spin32_cases.json")) as f:
        meta = json.load(f)
    return meta, np.load(os.path.join(GOLDEN, "spin32_cases.npz"))


@pytest.fixture(scope="session")
def golden_hams():
    return np.load(os.path.join(GOLDEN, "hamiltonians.npz"))


def params_from_meta(meta_entry):
    from quantumsimulations_b200 import DipolarRareParams

    return DipolarRareParams(**meta_entry["params"])


def random_table(n, seed, dense_pairs=True):
    """Random XYZ table (all three axes populated) as a SpinTable, plus bit-major arrays."""
    from quantumsimulations_b200.table import SpinTable

    rng = np.random.default_rng(seed)
    hx, hy, hz = rng.normal(size=(3, n))
    def tri():
        m = np.triu(rng.normal(size=(n, n)), 1)
        if not dense_pairs:
            m *= rng.random((n, n)) < 0.4
        return m
    sea = list(range(n - 1))
    return SpinTable(n, hx, hy, hz, tri(), tri(), tri(), sea, n - 1, int(rng.integers(0, 1 << n)))


def exact_traces(table, t):
    """Oracle X on a compiled table: dict of the six normalised traces (numpy index arithmetic + eigh)."""
    from oracle import dense, solvers

    bm = table.bit_major()
    H = dense.csr_from_table(bm["h"], bm["J"])
    psi0 = np.zeros(table.dim, complex)
    psi0[table.psi0_index] = 1.0
    states = solvers.exact_states(H, psi0, np.asarray(t))
    keys = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")
    if table.pauli_terms is not None:      # local dimension 4: the observables are weighted Pauli strings
        from oracle import pauli

        cols = pauli.expect_terms(states, table.pauli_terms)
        return {k: cols[i].copy() for i, k in enumerate(keys)}
    out = {k: np.empty(len(t)) for k in keys}
    for i in range(len(t)):
        o = dense.observable_sums(states[:, i], table.sea_mask, table.rare_mask)
        for k in keys:
            out[k][i] = o[k]
    return out
