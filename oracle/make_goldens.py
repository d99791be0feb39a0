"""Generate tests/golden/* by executing the REFERENCE's own code (build container only).

    python -m oracle.make_goldens            # from the repo root; needs /root/reference

TEST INFRASTRUCTURE ONLY.  The reference's ``dipolar_ensemble_with_rare.py`` and
``sweep_sea_detuning.py`` run unmodified; QuTiP is replaced by ``oracle/fake_qutip`` (exact
propagation, or SciPy zvode-Adams for the "Q" cases) and matplotlib by inert stubs.  Outputs:

  traj_cases.json / traj_cases.npz   params + (t, 7 traces, raw norm) per case  <- ref simulate_rare
  hamiltonians.npz                   CSR of H, psi0 index per small case        <- ref build_hamiltonian_rare
  sweep_small.json                   global_params / metrics rows of one sweep  <- ref run_sweep_sea_detuning
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

GOLD = os.path.join(ROOT, "tests", "golden")

# Ga/Al constants of the sweep's __main__ (SW:1205-1228)
G_SEA, G_RARE, B0_T = 8.1812e7, 6.976e7, 3.0
F1A = 50_000
TRACE_KEYS = ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R", "state_norm")


def sweep_point(ref, n_sea, delta_hz, variant, t_final, steps, solver=True, f1a=F1A):
    """DipolarRareParams exactly as SW:414-668 derives them for one (detuning, variant)."""
    f_az = G_SEA * B0_T / (2 * np.pi)
    f1r = (float(f1a) ** 2 + float(f1a) ** 2) ** 0.5
    b0 = 2 * np.pi * f_az / G_SEA
    f_rz = G_RARE * b0 / (2 * np.pi)
    kw = dict(
        n_sea=n_sea, gamma_sea=G_SEA, gamma_rare=G_RARE, B0_sea=b0, B0_rare=b0,
        B1_sea=2 * np.pi * f1a / G_SEA, B1_rare=2 * np.pi * f1r / G_RARE,
        omega_rf_sea=2 * np.pi * (f_az - delta_hz), omega_rf_rare=2 * np.pi * f_rz,
        phi_sea=np.pi / 2.0, phi_rare=np.pi / 2.0,
        dipolar_scale=1.0e-7 * 1.054571817e-34, shell_scale=0.282393e-9,
        t_final=t_final, steps=steps, drive_sea=True,
        drive_rare=(variant == "center_on"), init_x_sign=-1, init_rare_level=3,
        is_spin_three_half=False, is_center_rare=(variant != "shell_off"),
    )
    if solver:
        kw.update(solver_atol=1e-10, solver_rtol=1e-9, solver_nsteps=10_000_000, solver_max_step=1e-5)
    return ref.DipolarRareParams(**kw)


def case_list(ref):
    cases = []
    for delta in (0.0, 50e3, 100e3, 150e3):
        for variant in ("center_off", "center_on", "shell_off"):
            cases.append((f"n7_d{int(delta/1e3):03d}k_{variant}", "exact",
                          sweep_point(ref, 6, delta, variant, 3.0e-2, 2000)))
    # small / odd geometries, QuTiP-default options (state_norm comes back empty)
    cases.append(("n4_fib3_default_opts", "exact", sweep_point(ref, 3, 25e3, "center_on", 1e-3, 101, solver=False)))
    cases.append(("n5_tetra", "exact", sweep_point(ref, 4, 50e3, "center_on", 2e-3, 201)))
    cases.append(("n6_fib5_shell_off", "exact", sweep_point(ref, 5, 75e3, "shell_off", 2e-3, 201)))
    cases.append(("n9_cube", "exact", sweep_point(ref, 8, 50e3, "center_on", 1e-3, 101)))
    up = dataclasses.replace(sweep_point(ref, 6, 50e3, "center_on", 1e-3, 101), init_x_sign=+1, phi_sea=0.3, phi_rare=1.1)
    cases.append(("n7_spin_up_phases", "exact", up))
    # dataclass defaults with the only runnable spin setting (C1 (ii)): N=13 icosahedron, drives off
    cases.append(("n13_dataclass_default", "exact", ref.DipolarRareParams(is_spin_three_half=False, steps=201)))
    # reference-equivalent integrator (zvode-Adams) at the sweep's tolerances, short window
    cases.append(("q_n7_d050k_center_on_short", "zvode", sweep_point(ref, 6, 50e3, "center_on", 1e-4, 201)))
    cases.append(("q_n7_d050k_center_on_30ms", "zvode", sweep_point(ref, 6, 50e3, "center_on", 3.0e-2, 2000)))
    cases.append(("x_n7_d050k_center_on_short", "exact", sweep_point(ref, 6, 50e3, "center_on", 1e-4, 201)))
    # same integrator with tolerances tightened until it is converged (1e-8 at 0.03 s)
    tight = dataclasses.replace(sweep_point(ref, 6, 50e3, "center_on", 3.0e-2, 2000), solver_atol=1e-14, solver_rtol=1e-13)
    cases.append(("q_n7_d050k_center_on_30ms_tight", "zvode", tight))
    return cases


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref = ref_loader.load_reference_model("exact")
    import qutip  # the fake

    meta, arrays, hams = {}, {}, {}
    for name, backend, p in case_list(ref):
        qutip.SESOLVE_BACKEND = backend
        t, out = ref.simulate_rare(p)
        meta[name] = {"backend": backend, "params": dataclasses.asdict(p),
                      "freqs": ref.get_derived_frequencies(p)}
        arrays[f"{name}/t"] = t
        for k in TRACE_KEYS:
            arrays[f"{name}/{k}"] = np.asarray(out[k], dtype=float)
        if p.n_sea + 1 <= 9:
            H, _ = ref.build_hamiltonian_rare(p)
            csr = H.data.tocsr()
            csr.sort_indices()
            hams[f"{name}/data"] = csr.data
            hams[f"{name}/indices"] = csr.indices.astype(np.int32)
            hams[f"{name}/indptr"] = csr.indptr.astype(np.int32)
            psi0 = ref.initial_state_rare(p).full().ravel()
            hams[f"{name}/psi0_index"] = np.int64(np.argmax(np.abs(psi0)))
        print(f"{name:36s} {backend:6s} Iz_sea[1]={out['Iz_sea'][1]:+.15f}", flush=True)
    qutip.SESOLVE_BACKEND = "exact"
    with open(os.path.join(GOLD, "traj_cases.json"), "w") as f:
        json.dump(meta, f, indent=1, default=float)
    np.savez_compressed(os.path.join(GOLD, "traj_cases.npz"), **arrays)
    np.savez_compressed(os.path.join(GOLD, "hamiltonians.npz"), **hams)

    # one whole sweep through the reference's unchanged driver (function-default grid, SW:365-375)
    sw = ref_loader.load_reference_sweep()
    with tempfile.TemporaryDirectory() as tmp:
        base = sw.run_sweep_sea_detuning(
            f_Az=G_SEA * B0_T / (2 * np.pi), f1A=F1A, target_sea_detuning=F1A,
            gamma_sea=G_SEA, gamma_rare=G_RARE,
            sea_detunings_Hz=np.linspace(0.0, 3.0 * F1A, 13), n_sea=6,
            phi_sea=np.pi / 2.0, phi_rare=np.pi / 2.0, out_root=tmp, is_spin_three_half=False,
            solver_atol=1e-10, solver_rtol=1e-9, solver_nsteps=10_000_000, solver_max_step=1e-5,
        )
        with open(os.path.join(base, "summary.json")) as f:
            summary = json.load(f)
        listing = sorted(
            os.path.relpath(os.path.join(d, fn), base) for d, _, fns in os.walk(base) for fn in fns
        )
    with open(os.path.join(GOLD, "sweep_small.json"), "w") as f:
        json.dump({"summary": summary, "files": listing}, f, indent=1)
    print("wrote goldens to", GOLD)


if __name__ == "__main__":
    main()
