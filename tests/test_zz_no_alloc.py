"""qsb_apply_h_static - the allocation-free variant of qsb_apply_h (scratch in a static __device__ array; no counterpart
in the reference) - against the existing entry point, through the C ABI.

Collected LAST on purpose (file name).  Passes alone in a fresh process and after the
other tests: the scratch address is resolved per device inside the library, and every call clears its output slot.
"""
import numpy as np
import pytest

from conftest import random_table

pytestmark = pytest.mark.gpu

TOL = 2e-13   # the relative tolerance of tests/test_gpu_happly.py


def _rand_state(n, seed):
    rng = np.random.default_rng(seed)
    v = rng.normal(size=1 << n) + 1j * rng.normal(size=1 << n)
    return v / np.linalg.norm(v)


@pytest.mark.parametrize("n", [2, 5, 7, 8, 10])
def test_static_apply_matches_apply_h_on_two_calls_in_a_row(n):
    from oracle import dense
    from quantumsimulations_b200 import engine

    table = random_table(n, 1200 + n)
    bm = table.bit_major()
    with engine.Plan(table) as plan:
        for call in range(2):                       # the second call finds the first one's contents in the scratch
            psi = _rand_state(n, 10 * n + call)
            want = plan.apply_h(psi)                # existing entry point (cudaMalloc per call)
            got = plan.apply_h_static(psi)          # static scratch, no allocation
            scale = np.max(np.abs(want))
            assert np.max(np.abs(got - want)) / scale < TOL, (n, call)
            ref = dense.apply_h(bm["h"], bm["J"], psi)
            assert np.max(np.abs(got - ref)) / scale < TOL, (n, call)


def test_static_apply_smaller_call_after_larger_one():
    """A 2^3-amplitude call right after a 2^10 one: stale amplitudes of the larger call must not leak in."""
    from quantumsimulations_b200 import engine

    big, small = random_table(10, 77), random_table(3, 78)
    with engine.Plan(big) as pb, engine.Plan(small) as ps:
        pb.apply_h_static(_rand_state(10, 1))
        psi = _rand_state(3, 2)
        want = ps.apply_h(psi)
        got = ps.apply_h_static(psi)
    assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < TOL


def test_static_apply_rejects_sizes_beyond_the_scratch():
    from quantumsimulations_b200 import engine

    with engine.Plan(random_table(12, 79)) as plan:
        with pytest.raises(engine.EngineError, match="static scratch holds at most"):
            plan.apply_h_static(_rand_state(12, 3))
