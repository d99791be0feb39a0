"""CPU tests of the host layer: model mirror, table compiler, metrics, sweep driver file contract, the C ABI's
exported symbols and loud failure without a GPU, and the N>1 sharding logic over gloo.  No compute kernels."""
import dataclasses
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, exact_traces, params_from_meta

KEYS = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R")


# ---- model mirror ------------------------------------------------------------------------------------------
def test_params_fields_and_defaults_match_golden_serialisation(golden_cases):
    from quantumsimulations_b200 import DipolarRareParams

    meta, _ = golden_cases
    ref_keys = list(meta["n7_d050k_center_on"]["params"].keys())       # asdict() of the REFERENCE dataclass
    assert [f.name for f in dataclasses.fields(DipolarRareParams)] == ref_keys
    d = meta["n13_dataclass_default"]["params"]                          # reference defaults (+2 overrides)
    mine = dataclasses.asdict(DipolarRareParams(is_spin_three_half=False, steps=201))
    assert mine == d


def test_derived_frequencies_bit_exact(golden_cases):
    from quantumsimulations_b200 import get_derived_frequencies

    meta, _ = golden_cases
    for name, entry in meta.items():
        got = get_derived_frequencies(params_from_meta(entry))
        assert list(got.keys()) == list(entry["freqs"].keys())
        for k, v in entry["freqs"].items():
            assert got[k] == v, (name, k)


def test_error_behaviour_matches_reference():
    import quantumsimulations_b200 as q

    with pytest.raises(ValueError, match="n_sea must be at least 1"):
        q.shell_positions_with_rare_center(0)
    with pytest.raises(ValueError, match="identical positions"):
        q.dipolar_couplings_from_positions(np.zeros((3, 3)), 1.0, 1.0, 1.0)
    with pytest.raises(ValueError, match="Bad time grid"):
        q.simulate_rare(q.DipolarRareParams(n_sea=6, is_spin_three_half=False, steps=1))
    with pytest.raises(ValueError, match="Bad time grid"):
        q.simulate_rare(q.DipolarRareParams(n_sea=6, is_spin_three_half=False, t_final=0.0))
    # spin-3/2 flag on the sea-centre control: there is no rare spin, the record compiles as spin 1/2 (table.local_dim4)
    ctl = q.compile_table(q.DipolarRareParams(n_sea=4, is_center_rare=False))
    assert ctl.n == 5 and ctl.pauli_terms is None and ctl.frame_phi is None


def test_geometry_shapes_and_symmetry():
    import quantumsimulations_b200 as q

    for n in (1, 4, 5, 6, 8, 12, 20, 27, 32):
        pos = q.shell_positions_with_rare_center(n, 2.0)
        assert pos.shape == (n + 1, 3)
        assert np.allclose(np.linalg.norm(pos[:-1], axis=1), 2.0)
        assert np.all(pos[-1] == 0.0)
        b = q.dipolar_couplings_from_positions(pos, 1.0, 2.0, 3.0)
        assert np.array_equal(b, b.T) and np.all(np.diag(b) == 0.0)


# ---- table compiler --------------------------------------------------------------------------------------------
def test_spin_operator_views(golden_cases, golden_hams):
    import scipy.sparse as sp
    from quantumsimulations_b200 import build_hamiltonian_rare, initial_state_rare

    meta, _ = golden_cases
    name = "n5_tetra"
    p = params_from_meta(meta[name])
    H, obs = build_hamiltonian_rare(p)
    dim = 32
    G = sp.csr_matrix((golden_hams[f"{name}/data"], golden_hams[f"{name}/indices"], golden_hams[f"{name}/indptr"]),
                      shape=(dim, dim))
    assert abs(H.to_csr() - G).max() / np.abs(G.data).max() < 1e-14
    assert set(obs) == set(KEYS) and H.shape == (dim, dim)
    iz = obs["Iz_sea"].full()
    assert np.allclose(iz, np.diag(np.diag(iz))) and np.isclose(np.trace(iz), 0.0)
    psi0 = initial_state_rare(p)
    assert psi0[int(golden_hams[f"{name}/psi0_index"])] == 1.0 and np.count_nonzero(psi0) == 1
    assert np.allclose(obs["Ix_R"].full(), obs["Ix_R"].full().conj().T)


def test_bit_major_layout_is_site_reversed():
    from quantumsimulations_b200 import DipolarRareParams, compile_table

    t = compile_table(DipolarRareParams(n_sea=4, is_spin_three_half=False, drive_sea=True, B1_sea=0.3, omega_rf_sea=0.5))
    bm = t.bit_major()
    n = t.n
    assert np.array_equal(bm["h"][0], t.hx[::-1])
    assert bm["J"][2][0, n - 1] == t.jzz[0, n - 1]        # bit 0 = rare site, bit n-1 = site 0
    assert np.array_equal(bm["J"], bm["J"].transpose(0, 2, 1))
    assert t.rare_mask == 1 and t.sea_mask == (1 << n) - 2


# ---- product metrics vs oracle metrics ---------------------------------------------------------------------------
def test_product_metrics_agree_with_oracle():
    from oracle import metrics as om
    from quantumsimulations_b200 import metrics as pm

    rng = np.random.default_rng(3)
    t = np.linspace(0.0, 30.0, 20000)
    y = -3.0 + 1e-3 * t + 0.02 * rng.normal(size=t.size)
    for w in (1, 25, 100, 30000):
        a, b = pm.coarse_grain(t, y, w), om.block_mean(t, y, w)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    tc, yc = pm.coarse_grain(t, y, 100)
    assert pm.iz_slope_from_coarse(tc, yc) == om.central_fit(tc, yc)
    assert all(np.isnan(v) for v in pm.iz_slope_from_coarse(tc[:3], yc[:3]).values())
    for args in ((1.0, 2.0, 3.0, 0.5), (1.0, -2.0, 0.2, 0.3), (2.0, 1.0, 5.0, 5.0), (0.0, 0.0, 2.0, 2.0)):
        assert pm.contrast_michelson_with_t_gate(*args) == om.gated_contrast(*args)
    assert np.isnan(pm.contrast_michelson_with_t_gate(np.nan, 1.0, 2.0, 2.0))
    assert pm.detuning_label(50000.0) == "delta_p50000.0Hz" and pm.detuning_label(-12.34) == "delta_m12.3Hz"
    assert pm.f1R_for_resonance(50e3, 50e3) == (50e3**2 + 50e3**2) ** 0.5


# ---- sweep driver: file contract and numbers against the reference's own sweep output ---------------------------
def _oracle_batch(plist):
    from quantumsimulations_b200 import compile_table

    out = []
    for p in plist:
        t = np.linspace(0.0, p.t_final, p.steps)
        obs = exact_traces(compile_table(p), t)
        obs["state_norm"] = np.ones_like(t)
        out.append((t, obs))
    return out


def test_sweep_driver_reproduces_reference_sweep_files(tmp_path):
    """The engine's batched sweep driver, fed by the exact oracle instead of the GPU, must write the same file
    set and the same numbers as the reference's unchanged run_sweep_sea_detuning did (tests/golden/sweep_small.json)."""
    from quantumsimulations_b200.sweep import run_sweep_sea_detuning

    with open(os.path.join(GOLDEN, "sweep_small.json")) as f:
        gold = json.load(f)
    gp = gold["summary"]["global_params"]
    dets = gp["sea_detunings_Hz"][:3]
    base = run_sweep_sea_detuning(
        f_Az=gp["f_Az_Hz"], f1A=gp["f1A_Hz"], target_sea_detuning=gp["target_sea_detuning"],
        gamma_sea=gp["gamma_sea"], gamma_rare=gp["gamma_rare"], sea_detunings_Hz=dets, n_sea=gp["n_sea"],
        t_final=gp["t_final_s"], steps=gp["steps"], phi_sea=gp["phi_sea_rad"], phi_rare=gp["phi_rare_rad"],
        out_root=str(tmp_path), is_spin_three_half=False, solver_atol=gp["solver_atol"], solver_rtol=gp["solver_rtol"],
        solver_nsteps=gp["solver_nsteps"], solver_max_step=gp["solver_max_step"], coarse_window=gp["coarse_window"],
        simulate_batch=_oracle_batch, timestamp="test", verbose=False)
    with open(os.path.join(base, "summary.json")) as f:
        mine = json.load(f)
    assert list(mine["global_params"].keys()) == list(gp.keys())
    for k, v in gp.items():
        if k != "sea_detunings_Hz":
            assert mine["global_params"][k] == v, k
    for got, want in zip(mine["sweep_results"], gold["summary"]["sweep_results"]):
        assert list(got.keys()) == list(want.keys())
        for k in want:
            if k.startswith("t_") or k.startswith("R_"):
                assert got[k] == pytest.approx(want[k], rel=1e-6), k
            else:
                assert got[k] == pytest.approx(want[k], rel=1e-7, abs=1e-12), k
    written = sorted(os.path.relpath(os.path.join(d, fn), base) for d, _, fns in os.walk(base) for fn in fns)
    expected = [f for f in gold["files"] if not f.endswith((".png", ".pdf"))
                and (os.sep not in f or f.split(os.sep)[0] in {"delta_p0.0Hz", "delta_p12500.0Hz", "delta_p25000.0Hz"})]
    assert written == sorted(expected)
    z = np.load(os.path.join(base, "delta_p12500.0Hz", "time_and_obs_center_on.npz"))
    assert set(z.files) == {"t", "Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R", "state_norm"}


def test_replay_shim_batches_an_unchanged_script(tmp_path):
    """record -> batch -> replay around a stand-in 'reference script' module that calls simulate_rare serially."""
    import types
    from quantumsimulations_b200 import DipolarRareParams
    from quantumsimulations_b200.sweep import replay_reference_script

    calls = []
    mod = types.ModuleType("fake_script")
    mod.simulate_rare = lambda p: (_ for _ in ()).throw(AssertionError("serial path must not be used"))

    def script(out_root, n):
        tot = 0.0
        for k in range(n):
            t, obs = mod.simulate_rare(DipolarRareParams(n_sea=3, is_spin_three_half=False, t_final=1e-3 * (k + 1), steps=5))
            tot += float(obs["Iz_sea"][-1])
        with open(os.path.join(out_root, "out.txt"), "w") as f:
            f.write(repr(tot))
        return out_root

    mod.script = script

    def batch(plist):
        calls.append(len(plist))
        return [(np.linspace(0, p.t_final, p.steps), {k: np.full(p.steps, p.t_final) for k in KEYS + ("state_norm",)})
                for p in plist]

    out = replay_reference_script(mod, "script", simulate_batch=batch, out_root=str(tmp_path), n=4)
    assert calls == [4]                                   # one batched solve for the four serial calls
    assert float(open(os.path.join(out, "out.txt")).read()) == pytest.approx(1e-3 * 10)


def _check_sweep_dir_against_golden(base, rel_slope=1e-7, golden="sweep_small.json"):
    with open(os.path.join(GOLDEN, golden)) as f:
        gold = json.load(f)
    with open(os.path.join(base, "summary.json")) as f:
        mine = json.load(f)
    gp = gold["summary"]["global_params"]
    assert list(mine["global_params"].keys()) == list(gp.keys())
    for k, v in gp.items():
        assert mine["global_params"][k] == v, k
    assert len(mine["sweep_results"]) == len(gold["summary"]["sweep_results"]) == 13
    for got, want in zip(mine["sweep_results"], gold["summary"]["sweep_results"]):
        assert list(got.keys()) == list(want.keys())
        for k in want:
            if k.startswith("t_") or k.startswith("R_"):
                assert got[k] == pytest.approx(want[k], rel=max(1e-6, 10 * rel_slope)), k
            elif k.startswith("contrast"):
                # discontinuous at |t| = 1 (SW:308-315): only compared when both t-values are away from the gate
                continue
            else:
                assert got[k] == pytest.approx(want[k], rel=rel_slope, abs=1e-12), k
    written = sorted(os.path.relpath(os.path.join(d, fn), base) for d, _, fns in os.walk(base) for fn in fns)
    # the reference run that made the golden had the same inert plotting stubs: identical listing, PDFs/PNGs included
    assert written == sorted(gold["files"])
    return mine, gold


def test_unchanged_reference_sweep_script_runs_on_the_dropin(tmp_path):
    """Row (b) of SURVEY 8: the reference's sweep_sea_detuning.py, UNMODIFIED, imported with the drop-in directory
    first on sys.path, binds the engine's names (SW:103-109) and, run through the record -> batch -> replay shim at
    its function-default grid (SW:356-376, 13 detunings x 3 variants), makes exactly ONE batch call of 39 trajectories
    and writes the summary / file set the reference itself wrote (tests/golden/sweep_small.json).  The batch solver
    here is the exact CPU oracle (no GPU in this suite); tests/test_gpu_evolve.py repeats it on the engine."""
    import subprocess

    from oracle import ref_loader

    if not ref_loader.reference_available():
        pytest.skip("reference checkout not present")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "run_reference_sweep_on_dropin.py"), str(tmp_path), "oracle"],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    info = json.loads([ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1])
    assert info["calls"] == [39]
    assert info["bound"] == {"simulate_rare": True, "params": True, "freqs": True, "model_module_is_dropin": True,
                             "qutip_imported": False}
    mine, gold = _check_sweep_dir_against_golden(info["base"])
    for got, want in zip(mine["sweep_results"], gold["summary"]["sweep_results"]):
        for k in ("contrast_rare_center", "contrast_sea_center"):
            assert got[k] == pytest.approx(want[k], rel=1e-6, abs=1e-12), k


def test_unchanged_reference_sweep_script_runs_with_a_spin_three_half_rare_spin(tmp_path):
    """The same unmodified script with ``is_spin_three_half=True`` (SW:370, SW:651 hand the flag to every parameter
    record): one batch of 39 through the drop-in, whose rare-centre variants compile to the two-bit four-level site
    (table.py) and whose sea-centre control has no rare spin; summary and file set equal what the reference's own
    driver wrote on top of its model with the dims repaired (tests/golden/sweep_small_spin32.json,
    oracle/make_goldens_spin32.py)."""
    import subprocess

    from oracle import ref_loader

    if not ref_loader.reference_available():
        pytest.skip("reference checkout not present")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "run_reference_sweep_on_dropin.py"), str(tmp_path), "oracle",
                          "spin32"], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    info = json.loads([ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1])
    assert info["calls"] == [39] and info["bound"]["simulate_rare"] and not info["bound"]["qutip_imported"]
    mine, gold = _check_sweep_dir_against_golden(info["base"], golden="sweep_small_spin32.json")
    assert mine["global_params"]["rare_spin_type"] == "3/2"
    for got, want in zip(mine["sweep_results"], gold["summary"]["sweep_results"]):
        for k in ("contrast_rare_center", "contrast_sea_center"):
            assert got[k] == pytest.approx(want[k], rel=1e-6, abs=1e-12), k


def test_auto_method_picks_the_cheaper_path_by_size():
    """"auto": on-chip propagator for N <= 10 when there are more output intervals than basis states; for 11 <= N <= 13
    the HBM-resident propagator only where the ||H|| t_final bound says stepping is the expensive way (the sweep's
    parameters), not for the cheap dataclass-default trajectory; stepping beyond."""
    import dataclasses

    from bench import sweep_params
    from quantumsimulations_b200 import DipolarRareParams, api

    assert api.resolve_method(sweep_params(6, 50e3, t_final=30.0, steps=20000), None) == ("taylor", True)
    assert api.resolve_method(sweep_params(6, 50e3, t_final=1e-3, steps=50), None) == ("taylor", False)
    assert api.resolve_method(DipolarRareParams(is_spin_three_half=False), None) == ("taylor", False)          # N = 13, cheap
    assert api.resolve_method(sweep_params(12, 50e3, t_final=30.0, steps=20000), None) == ("taylor", True)     # N = 13, sweep
    assert api.resolve_method(sweep_params(10, 50e3, t_final=30.0, steps=20000), None) == ("taylor", True)     # N = 11
    assert api.resolve_method(sweep_params(13, 50e3, t_final=30.0, steps=20000), None) == ("taylor", True)     # N = 14
    assert api.resolve_method(sweep_params(14, 50e3, t_final=30.0, steps=20000), None) == ("taylor", False)    # N = 15
    assert api.resolve_method(sweep_params(12, 50e3, t_final=30.0, steps=20000), "dp5") == ("dp5", False)
    assert api.resolve_method(dataclasses.replace(sweep_params(12, 50e3), t_final=30.0, steps=20000), "propagator-dp5") == ("dp5", True)


# ---- C ABI -------------------------------------------------------------------------------------------------------
def test_library_exports_every_declared_symbol():
    from quantumsimulations_b200 import engine

    header = open(os.path.join(ROOT, "include", "qsb200.h")).read()
    declared = set(re.findall(r"^(?:int|const char\*|void\*)\s+(qsb_\w+)\s*\(", header, flags=re.M))
    assert len(declared) >= 15
    lib = engine.load_library()
    for sym in declared:
        assert hasattr(lib, sym), sym
    assert declared == set(engine.EXPORTS)
    assert lib.qsb_version() == 200


def test_no_silent_cpu_fallback():
    """Without a CUDA device every compute entry point must fail loudly (never route through the oracle)."""
    from quantumsimulations_b200 import DipolarRareParams, engine, simulate_rare

    if engine.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(engine.EngineError, match="no CUDA device"):
        simulate_rare(DipolarRareParams(n_sea=6, is_spin_three_half=False))
    src = ""
    pkg = os.path.join(ROOT, "quantumsimulations_b200")
    for d, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".inl", ".h")):
                src += open(os.path.join(d, fn)).read()
    assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M)


# ---- N > 1: sharding + gathering over gloo --------------------------------------------------------------------------
def test_partition_is_balanced_and_complete():
    from quantumsimulations_b200.dist import partition

    costs = [1.0, 4.0, 2.0, 8.0, 3.0, 3.0, 1.0, 6.0]
    bins = partition(costs, 3)
    assert sorted(i for b in bins for i in b) == list(range(len(costs)))
    loads = [sum(costs[i] for i in b) for b in bins]
    assert max(loads) - min(loads) <= max(costs)
    assert partition(costs, 1) == [list(range(len(costs)))]
    assert sum(len(b) for b in partition(costs[:2], 4)) == 2


_GLOO_WORKER = r"""
import os, sys
import numpy as np
import torch.distributed as dist
sys.path.insert(0, {root!r})
from quantumsimulations_b200 import DipolarRareParams
from quantumsimulations_b200.dist import simulate_distributed

dist.init_process_group("gloo")
rank = dist.get_rank()
plist = [DipolarRareParams(n_sea=3, is_spin_three_half=False, t_final=1e-3 * (k + 1), steps=4, drive_sea=True,
                           B1_sea=0.1 * k) for k in range(7)]
seen = []
def compute(sub, device):
    seen.extend(p.t_final for p in sub)
    return [(np.linspace(0, p.t_final, p.steps), {{"Iz_sea": np.full(p.steps, p.t_final)}}) for p in sub]
out = simulate_distributed(plist, device=0, compute=compute)
assert len(out) == 7 and all(o is not None for o in out)
for p, (t, obs) in zip(plist, out):
    assert t[-1] == p.t_final and obs["Iz_sea"][0] == p.t_final
assert 0 < len(seen) < 7                 # each rank integrated only its share
counts = [None, None]
dist.all_gather_object(counts, len(seen))
assert sum(counts) == 7
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_distributed_sweep_sharding_world_size_2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER.format(root=ROOT))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29531", str(script)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert res.stdout.count("ok") == 2


# ---- L2-resident fusion of the strided passes: the ticket schedule, checked on the host ------------------------
def _deposit(x, mask):
    out, pos = 0, 0
    while mask:
        low = mask & -mask
        if x & 1:
            out |= low
        x >>= 1
        mask ^= low
    return out


def _tile_indices(lay, v, n):
    """Index set of tile v of layout (lk, gs, ge): outer bits ascending = bits of v."""
    lk, gs, ge = lay
    tile_bits = list(range(lk)) + list(range(gs, ge))
    outer = [b for b in range(n) if b not in tile_bits]
    base = sum(((v >> i) & 1) << b for i, b in enumerate(outer))
    idx = np.zeros(1, dtype=np.int64) + base
    for b in tile_bits:
        idx = np.concatenate([idx, idx | (1 << b)])
    return idx


@pytest.mark.parametrize("n,gmax", [(16, 2), (18, 2), (17, 3), (19, 4), (21, 8)])
def test_fused_strided_schedule_covers_tiles_and_respects_dependencies(n, gmax, monkeypatch):
    """Replays pass_strided_fused's ticket decoding: every tile of every fused layout is visited once per phase,
    the phases of a unit touch one and the same index set, and a phase's dependency always has smaller tickets."""
    from quantumsimulations_b200 import engine

    monkeypatch.setenv("QSB200_MAX_GROUP", str(gmax))
    monkeypatch.setenv("QSB200_FUSE_LB", "20")
    lays = engine.tile_layouts(n)
    fu = engine.strided_fusion(n, 148)
    assert fu is not None and fu["f"] + fu["m"] == len(lays) - 1
    f, m, lag, T = fu["f"], fu["m"], fu["lag"], 1 << fu["tiles_lb"]
    P, U = 2 * m + 1, 1 << fu["fbits"]  # one branch is enough: the branch is the low bit of the unit number
    n_tiles = 1 << (n - 12)
    assert U * T == n_tiles and fu["vmask"] == n_tiles - 1
    first_ticket, last_ticket, seen, sets = {}, {}, {}, {}
    ticket = 0
    for slot in range(U + (P - 1) * lag):
        for p in range(P - 1, -1, -1):
            u = slot - p * lag
            for q in range(T):
                if 0 <= u < U:
                    li = f + (p if p <= m else 2 * m - p)
                    fm = fu["fixed_vmask"][li]
                    v = _deposit(u, fm) | _deposit(q, fu["vmask"] & ~fm)
                    seen.setdefault(p, []).append(v)
                    first_ticket.setdefault((u, p), ticket)
                    last_ticket[(u, p)] = ticket
                    if u in (0, U - 1, U // 2):
                        sets.setdefault((u, p), []).append(_tile_indices(lays[li], v, n))
                ticket += 1
    for p in range(P):
        assert sorted(seen[p]) == list(range(n_tiles))
    for u in range(U):
        for p in range(1, P):
            assert last_ticket[(u, p - 1)] < first_ticket[(u, p)]
    for u in (0, U - 1, U // 2):
        ref = np.sort(np.concatenate(sets[(u, 0)]))
        assert len(np.unique(ref)) == len(ref) == (1 << n) // U
        for p in range(1, P):
            assert np.array_equal(np.sort(np.concatenate(sets[(u, p)])), ref)


def test_fusion_plan_at_bench_sizes(monkeypatch):
    from quantumsimulations_b200 import engine

    assert engine.strided_fusion(28, 148) is None  # opt-in: off by default
    monkeypatch.setenv("QSB200_FUSE_LB", "20")
    fu = engine.strided_fusion(28, 148)
    assert fu and (fu["f"], fu["m"], fu["fbits"], fu["tiles_lb"]) == (1, 1, 8, 8)  # 16 MB units
    assert engine.strided_fusion(20, 148) is None  # two layouts: nothing to fuse
    fu30 = engine.strided_fusion(30, 148)
    assert fu30 and fu30["f"] == 2  # the first mid layout stays a full-grid launch


def test_multi_sweep_grid_is_one_batch_and_one_directory_per_f1a(tmp_path):
    """C3: the 2D (f1A x detuning) grid goes through ONE batch call and lands as one reference-layout sweep
    directory per f1A (what 2D_sweep_report.py walks for summary.json); each sweep equals the stand-alone one."""
    from quantumsimulations_b200.sweep import run_multi_sweep, run_sweep_sea_detuning

    with open(os.path.join(GOLDEN, "sweep_small.json")) as f:
        gp = json.load(f)["summary"]["global_params"]
    calls = []

    def batch(plist):
        calls.append(len(plist))
        return _oracle_batch(plist)

    common = dict(f_Az=gp["f_Az_Hz"], gamma_sea=gp["gamma_sea"], gamma_rare=gp["gamma_rare"], n_sea=gp["n_sea"],
                  t_final=2.0e-3, steps=101, phi_sea=gp["phi_sea_rad"], phi_rare=gp["phi_rare_rad"],
                  is_spin_three_half=False, solver_atol=gp["solver_atol"], solver_rtol=gp["solver_rtol"],
                  solver_nsteps=gp["solver_nsteps"], solver_max_step=gp["solver_max_step"], coarse_window=5)
    f1as = [20e3, 35e3]
    rule = lambda f1a: np.linspace(0.0, 3.0 * f1a, 3)
    dirs = run_multi_sweep(f1A_values=f1as, detunings_for=rule, out_root=str(tmp_path / "grid"), simulate_batch=batch,
                           timestamp="20250101_000000", verbose=False, **common)
    assert calls == [2 * 3 * 3]
    assert [os.path.basename(d) for d in dirs] == ["sea_detuning_sweep_20250101_000000", "sea_detuning_sweep_20250101_000001"]
    for f1a, d in zip(f1as, dirs):
        with open(os.path.join(d, "summary.json")) as f:
            got = json.load(f)
        assert got["global_params"]["f1A_Hz"] == f1a and got["global_params"]["target_sea_detuning"] == f1a
        assert [r["delta_Hz"] for r in got["sweep_results"]] == list(rule(f1a))
        alone = run_sweep_sea_detuning(f1A=f1a, target_sea_detuning=f1a, sea_detunings_Hz=rule(f1a),
                                       out_root=str(tmp_path / "alone"), simulate_batch=_oracle_batch,
                                       timestamp=f"{int(f1a)}", verbose=False, **common)
        with open(os.path.join(alone, "summary.json")) as f:
            want = json.load(f)
        assert json.dumps(got, sort_keys=True) == json.dumps(want, sort_keys=True)


def test_trajectory_cache_serves_repeated_and_interrupted_sweeps(tmp_path):
    """8(f) rank 3 for the on-chip path: finished trajectories are cached by their parameters, so the second sweep
    integrates nothing and a sweep that shares points with an earlier one integrates only the new ones."""
    from quantumsimulations_b200.checkpoint import trajectory_key
    from quantumsimulations_b200.sweep import run_sweep_sea_detuning

    with open(os.path.join(GOLDEN, "sweep_small.json")) as f:
        gp = json.load(f)["summary"]["global_params"]
    calls = []

    def batch(plist):
        calls.append(len(plist))
        return _oracle_batch(plist)

    kw = dict(f_Az=gp["f_Az_Hz"], f1A=gp["f1A_Hz"], target_sea_detuning=gp["target_sea_detuning"], gamma_sea=gp["gamma_sea"],
              gamma_rare=gp["gamma_rare"], n_sea=gp["n_sea"], t_final=2.0e-3, steps=101, phi_sea=gp["phi_sea_rad"],
              phi_rare=gp["phi_rare_rad"], solver_atol=1e-10, solver_rtol=1e-9, solver_nsteps=10**7, solver_max_step=1e-5,
              coarse_window=5, simulate_batch=batch, cache_dir=str(tmp_path / "cache"), verbose=False)
    a = run_sweep_sea_detuning(sea_detunings_Hz=[0.0, 25e3], out_root=str(tmp_path / "a"), timestamp="a", **kw)
    b = run_sweep_sea_detuning(sea_detunings_Hz=[0.0, 25e3, 50e3], out_root=str(tmp_path / "b"), timestamp="b", **kw)
    c = run_sweep_sea_detuning(sea_detunings_Hz=[0.0, 25e3, 50e3], out_root=str(tmp_path / "c"), timestamp="c", **kw)
    assert calls == [6, 3]          # b integrates only delta = 50 kHz, c nothing
    sb, sc = (json.load(open(os.path.join(d, "summary.json")))["sweep_results"] for d in (b, c))
    assert sb == sc
    sa = json.load(open(os.path.join(a, "summary.json")))["sweep_results"]
    assert sa == sb[:2]
    za = np.load(os.path.join(a, "delta_p25000.0Hz", "time_and_obs_center_on.npz"))
    zc = np.load(os.path.join(c, "delta_p25000.0Hz", "time_and_obs_center_on.npz"))
    assert all(np.array_equal(za[k], zc[k]) for k in za.files)
    # the key sees every field of the record and the solver choice
    from conftest import params_from_meta  # noqa: F401
    from quantumsimulations_b200 import DipolarRareParams
    p = DipolarRareParams(n_sea=6, is_spin_three_half=False)
    assert trajectory_key(p) != trajectory_key(dataclasses.replace(p, gamma_sea=np.nextafter(p.gamma_sea, np.inf)))
    assert trajectory_key(p, method="dp5") != trajectory_key(p, method="taylor")
    # ... and the RESOLVED numerics: environment overrides, explicit tol_scale, numpy scalars in the record
    base_key = trajectory_key(p)
    os.environ["QSB200_TOL_SCALE"] = "1e-4"
    try:
        assert trajectory_key(p) != base_key
    finally:
        del os.environ["QSB200_TOL_SCALE"]
    assert trajectory_key(p, tol_scale=1e-2) != base_key
    assert trajectory_key(p, method="auto") == base_key                                   # auto is the default
    assert trajectory_key(dataclasses.replace(p, steps=5)) != trajectory_key(dataclasses.replace(p, steps=5), method="propagator")
    assert trajectory_key(dataclasses.replace(p, n_sea=np.int64(6))) == base_key


# ---- the reference's consumers, unchanged, on the engine's output files -------------------------------------------
def _load_reference_script(name):
    """Import one of the reference's report scripts (file names may start with a digit) on the inert plotting stubs."""
    import importlib.util

    from oracle import ref_loader

    ref_loader.load_reference_sweep()           # fake qutip + inert matplotlib / tkinter on sys.path
    alias = "ref_" + name.replace("2D", "twoD")
    spec = importlib.util.spec_from_file_location(alias, os.path.join(ref_loader.REFERENCE_DIR, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[alias] = mod
    spec.loader.exec_module(mod)
    return mod


def test_reference_report_scripts_consume_the_engine_sweep_files(tmp_path):
    """reprocess_sweep_results.py, reprocess_detuning_sweep_exponential.py and the two 2D_sweep_report*.py scripts
    (pure file consumers) run UNCHANGED on what run_multi_sweep writes, and the numbers they
    re-derive from the npz files equal the metrics the engine's driver stored."""
    from oracle import ref_loader

    if not ref_loader.reference_available():
        pytest.skip("reference checkout not present")
    from quantumsimulations_b200.sweep import run_multi_sweep

    with open(os.path.join(GOLDEN, "sweep_small.json")) as f:
        gp = json.load(f)["summary"]["global_params"]
    common = dict(f_Az=gp["f_Az_Hz"], gamma_sea=gp["gamma_sea"], gamma_rare=gp["gamma_rare"], n_sea=gp["n_sea"],
                  t_final=3.0e-3, steps=151, phi_sea=gp["phi_sea_rad"], phi_rare=gp["phi_rare_rad"],
                  is_spin_three_half=False, solver_atol=gp["solver_atol"], solver_rtol=gp["solver_rtol"],
                  solver_nsteps=gp["solver_nsteps"], solver_max_step=gp["solver_max_step"], coarse_window=5)
    f1as = [20e3, 35e3]
    root = str(tmp_path / "grid")
    dirs = run_multi_sweep(f1A_values=f1as, detunings_for=lambda f: np.linspace(0.0, 3.0 * f, 3), out_root=root,
                           simulate_batch=_oracle_batch, timestamp="20250101_000000", verbose=False, **common)
    mine = {d: json.load(open(os.path.join(d, "summary.json"))) for d in dirs}

    # 2D_sweep_report.py / 2D_sweep_report_stable_region.py: walk the root for summary.json (R2:199-303)
    for script in ("2D_sweep_report", "2D_sweep_report_stable_region"):
        r2 = _load_reference_script(script)
        pts = r2.aggregate_points(root)
        finite = [(s["global_params"]["f1A_Hz"], r) for s in mine.values() for r in s["sweep_results"]
                  if np.isfinite(r["DeltaOmega_over_geff"]) and np.isfinite(r["contrast_rare_center"])]
        assert len(pts) == len(finite) > 0
        assert sorted({p["f1A_Hz"] for p in pts}) == f1as
        assert sorted(p["contrast"] for p in pts) == sorted(r["contrast_rare_center"] for _, r in finite)
    _load_reference_script("2D_sweep_report").make_plots(root, os.path.join(root, "report2d.pdf"))

    # reprocess_sweep_results.py: re-reads summary.json + the three npz per detuning and recomputes the slopes (RS:95-)
    rs = _load_reference_script("reprocess_sweep_results")
    rs.reprocess_sweep(dirs[0], window=5)
    with open(os.path.join(dirs[0], "summary_reprocessed_win5.json")) as f:
        rep = json.load(f)
    rows = rep["sweep_results"] if isinstance(rep, dict) else rep
    assert len(rows) == 3
    for got, want in zip(rows, mine[dirs[0]]["sweep_results"]):
        assert got["delta_Hz"] == want["delta_Hz"]
        for k in ("I_z_slope_off_center", "I_z_slope_on_center", "I_z_slope_off_sea_center"):
            assert k in got
            if np.isfinite(want[k]):
                assert got[k] == pytest.approx(want[k], rel=1e-9, abs=1e-15), k

    # reprocess_detuning_sweep_exponential.py: needs metrics.json + the two centre npz per detuning directory
    rex = _load_reference_script("reprocess_detuning_sweep_exponential")
    det_dirs = rex.discover_detuning_dirs(dirs[0])
    assert len(det_dirs) == 3
    t, obs = rex.load_npz(det_dirs[0], "center_on")
    assert t.shape == (151,) and "Iz_sea" in obs


# ---- propagator path: host-side logic (the kernels themselves are covered by tests/test_gpu_propagator.py) -----------
def test_method_strings_and_uniform_grid_detection(monkeypatch):
    from quantumsimulations_b200 import api, engine

    monkeypatch.delenv("QSB200_METHOD", raising=False)
    assert api.split_method(None) == ("taylor", False)            # "auto" without a size: stepping
    assert api.split_method(None, 7, 2000) == ("taylor", True)    # more output intervals than basis states
    assert api.split_method("auto", 7, 129) == ("taylor", False)
    assert api.split_method("auto", 7, 130) == ("taylor", True)
    assert api.split_method("auto", 13, 20000) == ("taylor", False)  # tiled path: no dense propagator
    assert api.split_method("taylor", 7, 2000) == ("taylor", False)
    assert api.split_method("dp5") == ("dp5", False)
    assert api.split_method("propagator") == ("taylor", True)
    assert api.split_method("propagator-dp5") == ("dp5", True)
    monkeypatch.setenv("QSB200_METHOD", "propagator")
    assert api.split_method(None) == ("taylor", True)
    with pytest.raises(ValueError, match="unknown method"):
        api.split_method("zvode")
    # np.linspace grids (DE:626) are uniform; anything else is refused by the propagator path
    assert engine.uniform_step(np.linspace(0.0, 30.0, 20000)) == pytest.approx(30.0 / 19999, rel=1e-14)
    assert engine.uniform_step(np.linspace(1.0, 1.5, 7)) == pytest.approx(0.5 / 6)
    assert engine.uniform_step(np.array([0.0, 1.0, 3.0])) is None
    assert engine.uniform_step(np.array([0.0])) is None
    # the integrator options of the column trajectories are those of the stepping path
    p = api.DipolarRareParams(n_sea=6, is_spin_three_half=False, solver_atol=1e-10, solver_rtol=1e-9)
    o, override = api._options_from(p, "propagator")
    assert (o.method, o.order, override) == ("taylor", 20, True) and o.atol == pytest.approx(1e-15)
    o, _ = api._options_from(p, "propagator-dp5")
    assert o.method == "dp5" and o.atol == 1e-10
    monkeypatch.delenv("QSB200_METHOD")
    o, _ = api._options_from(dataclasses.replace(p, steps=2000), None)      # auto -> propagator tolerances
    assert o.method == "taylor" and o.atol == pytest.approx(1e-15)
    o, _ = api._options_from(dataclasses.replace(p, steps=100), None)       # auto -> stepping tolerances
    assert o.method == "taylor" and o.atol == pytest.approx(1e-13)


def test_propagate_kernel_launch_shapes():
    """Host-only query: threads = 2^n x split (>= 32), split a power of two, cached columns fit the opt-in shared
    memory, and the whole propagator is cached whenever it fits (n <= 6)."""
    from quantumsimulations_b200 import engine

    for n in range(1, engine.SMALL_MAX_QUBITS + 1):
        s = engine.propagate_shape(n)
        dim = 1 << n
        assert s["threads"] == max(32, dim * s["split"]) and s["threads"] <= 1024
        assert s["split"] in (1, 2, 4) and s["split"] <= dim     # the partial sums of a row meet inside one warp
        assert 2 <= s["ring"] <= max(2, s["threads"] // 32)
        assert 1 <= s["cached_columns"] <= dim
        assert s["smem_bytes"] <= 227 * 1024
        if n <= 6:
            assert s["cached_columns"] == dim
    with pytest.raises(ValueError):
        engine.propagate_shape(11)


def test_propagator_needs_small_n_and_no_device_fails_loudly():
    from quantumsimulations_b200 import DipolarRareParams, engine, simulate_rare, simulate_rare_batch

    big = DipolarRareParams(n_sea=14, is_spin_three_half=False)      # N = 15: beyond the HBM-resident propagator too
    with pytest.raises(ValueError, match="N <= 14"):
        simulate_rare(big, method="propagator")
    with pytest.raises(ValueError, match="N <= 14"):
        simulate_rare_batch([big], method="propagator")
    if engine.device_count() == 0:
        with pytest.raises(engine.EngineError, match="no CUDA device"):
            simulate_rare(DipolarRareParams(n_sea=6, is_spin_three_half=False), method="propagator")


def test_sweep_driver_forwards_method_to_the_batch_solver(tmp_path, monkeypatch):
    """run_sweep_sea_detuning(method=...) reaches simulate_rare_batch and becomes part of the cache key."""
    from quantumsimulations_b200 import dist, sweep

    seen = []

    def fake_batch(plist, device=0, **kw):
        seen.append(kw)
        return _oracle_batch(plist), [None] * len(plist)

    monkeypatch.setattr("quantumsimulations_b200.api.simulate_rare_batch", fake_batch)
    with open(os.path.join(GOLDEN, "sweep_small.json")) as f:
        gp = json.load(f)["summary"]["global_params"]
    kw = dict(f_Az=gp["f_Az_Hz"], f1A=gp["f1A_Hz"], target_sea_detuning=gp["target_sea_detuning"],
              gamma_sea=gp["gamma_sea"], gamma_rare=gp["gamma_rare"], sea_detunings_Hz=[0.0, 50e3], n_sea=gp["n_sea"],
              t_final=1.0e-3, steps=51, phi_sea=gp["phi_sea_rad"], phi_rare=gp["phi_rare_rad"],
              is_spin_three_half=False, solver_atol=gp["solver_atol"], solver_rtol=gp["solver_rtol"],
              solver_nsteps=gp["solver_nsteps"], solver_max_step=gp["solver_max_step"], coarse_window=5, verbose=False)
    cache = str(tmp_path / "cache")
    sweep.run_sweep_sea_detuning(out_root=str(tmp_path / "a"), method="propagator", cache_dir=cache, **kw)
    assert seen == [{"method": "propagator"}]
    n_prop = len(os.listdir(cache))
    assert n_prop == 6
    sweep.run_sweep_sea_detuning(out_root=str(tmp_path / "b"), cache_dir=cache, **kw)   # default method: new entries
    assert seen[1] == {} and len(os.listdir(cache)) == 2 * n_prop
    sweep.run_sweep_sea_detuning(out_root=str(tmp_path / "c"), method="propagator", cache_dir=cache, **kw)
    assert len(seen) == 2                                                                 # served from the cache
    assert dist is not None
