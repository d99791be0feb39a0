"""Generate tests/golden/spin32_cases.* by executing the REFERENCE's own code with its dims repaired (build container only).

    python -m oracle.make_goldens_spin32       # from the repo root; needs /root/reference

TEST INFRASTRUCTURE ONLY.  ``is_spin_three_half=True`` (the J = 3/2 rare spin, DE:22-25 / DE:499-501; the dataclass
default, DE:377) cannot run in the reference as written: DE:486 calls ``dims_with_rare(n_sea, False)`` whenever the
centre IS the rare spin, so every local dimension is 2 while the rare operators are 4 x 4.  The one-line
repair - ``dims_with_rare`` returns ``[2] * n_sea + [4]``, which is what its own docstring (DE:28-34) and
``initial_state_rare`` (DE:602-603, a 4-level rare ket) describe - is applied here by rebinding that ONE module-level
function; everything else (``build_hamiltonian_rare``, ``initial_state_rare``, ``simulate_rare``) is the reference's
unmodified code on top of ``oracle/fake_qutip`` (exact propagation).  The fixtures pin the operator, the observables
and the traces of the local-dimension-4 row; like every other fixture they leave the integrator half unpinned.

  spin32_cases.json / spin32_cases.npz   params + (t, 7 traces) per case, CSR of H + psi0 index for the small ones
  sweep_small_spin32.json                global_params / metrics rows / file listing of one sweep through the
                                         reference's unchanged driver with is_spin_three_half=True (see ``sweep_golden``)
"""
from __future__ import annotations

import dataclasses
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402
from oracle.make_goldens import B0_T, F1A, G_RARE, G_SEA, GOLD, TRACE_KEYS, sweep_point  # noqa: E402


def load_repaired_reference():
    """The reference module with ``dims_with_rare`` giving the rare site local dimension 4 (see module docstring)."""
    ref = ref_loader.load_reference_model("exact")
    if not getattr(ref, "_qsb_dims_repaired", False):
        original = ref.dims_with_rare
        ref.dims_with_rare = lambda n_sea, is_spin_three_half=False: original(n_sea, True)
        ref._qsb_dims_repaired = True
        ref._qsb_dims_original = original
    return ref


def restore_reference(ref) -> None:
    if getattr(ref, "_qsb_dims_repaired", False):
        ref.dims_with_rare = ref._qsb_dims_original
        ref._qsb_dims_repaired = False


def case_list(ref):
    def point(n_sea, delta, variant, t_final, steps, **kw):
        solver = kw.pop("solver", True)
        p = sweep_point(ref, n_sea, delta, variant, t_final, steps, solver=solver)
        return dataclasses.replace(p, is_spin_three_half=True, **kw)

    cases = [
        ("s32_n3_default_opts", point(3, 25e3, "center_on", 1e-3, 101, solver=False)),
        # every knob away from its symmetric value: both drive phases, spin-up start (rare level m = 3), rare detuning
        ("s32_n5_phases_detuned", point(5, 75e3, "center_on", 2e-3, 301, init_x_sign=+1, phi_sea=0.3, phi_rare=1.1,
                                        B0_rare=3.0 * (1.0 + 2.0e-4))),
        ("s32_n6_center_on", point(6, 50e3, "center_on", 3.0e-2, 2000)),
        ("s32_n6_center_off", point(6, 100e3, "center_off", 3.0e-2, 2000)),
        ("s32_n8_cube", point(8, 50e3, "center_on", 1e-3, 101)),
        # beyond the on-chip path (n_sea + 2 > 10 bits): stepping sessions on the tiled path
        ("s32_n9_tiled", point(9, 50e3, "center_on", 2e-4, 21, phi_rare=0.7)),
        ("s32_n10_tiled", point(10, 25e3, "center_on", 1e-4, 11)),
        # the reference's dataclass defaults as they stand (DE:332-384: is_spin_three_half=True, n_sea = 12 -> 14 bits,
        # drives off, QuTiP-default options), 201 of the 2000 outputs
        ("s32_dataclass_default", ref.DipolarRareParams(steps=201)),
    ]
    return cases


def sweep_golden(ref):
    """One sweep through the reference's UNMODIFIED ``run_sweep_sea_detuning`` (SW:356-1165) at its function-default grid
    with ``is_spin_three_half=True``.  Its two rare-centre variants run ``ref.simulate_rare`` with the dims repaired.  Its
    sea-centre control (``is_center_rare=False``, SW:640-668) has no rare spin: the centre is one more sea spin 1/2
    (DE:488-489), the reference's own rare observables for it mix 4 x 4 operators into all-2 dims (DE:499-501 /
    DE:576-578) under any repair, so that variant runs the unrepaired reference with the meaningless flag off - the
    semantics ``quantumsimulations_b200.table.local_dim4`` states.  Only ``sw.simulate_rare`` is rebound."""
    sw = ref_loader.load_reference_sweep()
    original = sw.simulate_rare

    def simulate(params):
        if params.is_center_rare:
            return ref.simulate_rare(params)
        restore_reference(ref)
        try:
            return ref.simulate_rare(dataclasses.replace(params, is_spin_three_half=False))
        finally:
            load_repaired_reference()

    sw.simulate_rare = simulate
    try:
        with tempfile.TemporaryDirectory() as tmp:
            base = sw.run_sweep_sea_detuning(
                f_Az=G_SEA * B0_T / (2 * np.pi), f1A=F1A, target_sea_detuning=F1A, gamma_sea=G_SEA, gamma_rare=G_RARE,
                sea_detunings_Hz=np.linspace(0.0, 3.0 * F1A, 13), n_sea=6, phi_sea=np.pi / 2.0, phi_rare=np.pi / 2.0,
                out_root=tmp, is_spin_three_half=True, solver_atol=1e-10, solver_rtol=1e-9, solver_nsteps=10_000_000,
                solver_max_step=1e-5)
            with open(os.path.join(base, "summary.json")) as f:
                summary = json.load(f)
            listing = sorted(os.path.relpath(os.path.join(d, fn), base) for d, _, fns in os.walk(base) for fn in fns)
    finally:
        sw.simulate_rare = original
    with open(os.path.join(GOLD, "sweep_small_spin32.json"), "w") as f:
        json.dump({"summary": summary, "files": listing}, f, indent=1)


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref = load_repaired_reference()
    meta, arrays = {}, {}
    for name, p in case_list(ref):
        t, out = ref.simulate_rare(p)
        meta[name] = {"params": dataclasses.asdict(p), "freqs": ref.get_derived_frequencies(p)}
        arrays[f"{name}/t"] = t
        for k in TRACE_KEYS:
            arrays[f"{name}/{k}"] = np.asarray(out[k], dtype=float)
        psi0 = ref.initial_state_rare(p).full().ravel()
        arrays[f"{name}/psi0_index"] = np.int64(np.argmax(np.abs(psi0)))
        if p.n_sea + 2 <= 8:
            H, obs = ref.build_hamiltonian_rare(p)
            csr = H.data.tocsr()
            csr.sort_indices()
            arrays[f"{name}/H_data"] = csr.data
            arrays[f"{name}/H_indices"] = csr.indices.astype(np.int32)
            arrays[f"{name}/H_indptr"] = csr.indptr.astype(np.int32)
            if p.n_sea + 2 <= 5:            # the six observables as dense matrices, smallest case only
                for k, op in obs.items():
                    arrays[f"{name}/op_{k}"] = op.full()
        print(f"{name:28s} Iz_R[1]={out['Iz_R'][1]:+.15f} Ix_R[1]={out['Ix_R'][1]:+.15f}", flush=True)
    if "--cases-only" not in sys.argv:
        sweep_golden(ref)
    restore_reference(ref)
    with open(os.path.join(GOLD, "spin32_cases.json"), "w") as f:
        json.dump(meta, f, indent=1, default=float)
    np.savez_compressed(os.path.join(GOLD, "spin32_cases.npz"), **arrays)
    print("wrote spin-3/2 goldens to", GOLD)


if __name__ == "__main__":
    main()
