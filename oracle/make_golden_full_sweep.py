"""Generate tests/golden/sweep_full.json: the reference's UNCHANGED sweep driver at its own __main__ settings
(sweep_sea_detuning.py:1199-1252: N = 7, t_final = 30 s, 20000 outputs, coarse_window = 100) with QuTiP replaced by
oracle/fake_qutip on its exact-propagation backend.  Build container only (needs /root/reference).

    python -m oracle.make_golden_full_sweep

TEST INFRASTRUCTURE ONLY.  The real QuTiP run would take ~10^9 zvode steps per trajectory; exact propagation gives
the traces that run converges to, and the reference's own coarse-grain / slope / contrast code turns them into the
metrics rows stored here.
"""
from __future__ import annotations

import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    ref_loader.load_reference_model("exact")
    import qutip  # the stand-in

    qutip.SESOLVE_BACKEND = "exact"
    sw = ref_loader.load_reference_sweep()
    gamma_sea, gamma_rare, b0, f1a = 8.1812e7, 6.976e7, 3.0, 50_000
    with tempfile.TemporaryDirectory() as tmp:
        base = sw.run_sweep_sea_detuning(
            f_Az=gamma_sea * b0 / (2 * np.pi), f1A=f1a, target_sea_detuning=f1a, gamma_sea=gamma_sea,
            gamma_rare=gamma_rare, sea_detunings_Hz=np.linspace(0.0, 3.0 * f1a, 13), n_sea=6, t_final=30, steps=20000,
            phi_sea=np.pi / 2.0, phi_rare=np.pi / 2.0, out_root=tmp, is_spin_three_half=False, solver_atol=1e-10,
            solver_rtol=1e-9, solver_nsteps=10_000_000, solver_max_step=1e-5, coarse_window=100)
        with open(os.path.join(base, "summary.json")) as f:
            summary = json.load(f)
        # a few raw samples per trajectory so that the traces themselves are pinned, not only the fitted metrics
        samples = {}
        for d in sorted(os.listdir(base)):
            for tag in ("center_off", "center_on", "shell_off"):
                p = os.path.join(base, d, f"time_and_obs_{tag}.npz")
                if os.path.exists(p):
                    with np.load(p) as z:
                        idx = [0, 1, 2, 999, 5000, 12345, 19999]
                        samples[f"{d}/{tag}"] = {"index": idx, "Iz_sea": [float(z["Iz_sea"][i]) for i in idx],
                                                 "Iz_R": [float(z["Iz_R"][i]) for i in idx]}
    with open(os.path.join(GOLD, "sweep_full.json"), "w") as f:
        json.dump({"summary": summary, "samples": samples}, f, indent=1)
    print("wrote", os.path.join(GOLD, "sweep_full.json"))


if __name__ == "__main__":
    main()
