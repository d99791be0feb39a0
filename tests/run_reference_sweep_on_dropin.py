"""Helper of test_cpu_host.py::test_unchanged_reference_sweep_script_runs_on_the_dropin (run as a subprocess so that
no earlier import of the reference's own model module is in sys.modules).

Puts the drop-in directory FIRST on sys.path, imports the reference's UNMODIFIED sweep_sea_detuning.py (its
`from dipolar_ensemble_with_rare import ...`, SW:103-109, must bind to the engine), and runs its
run_sweep_sea_detuning (SW:356-376) at the function-default grid through the record -> batch -> replay shim.
argv[1]: output directory; argv[2]: "oracle" (exact CPU oracle as the batch solver: no GPU) or "engine";
argv[3] (optional) "spin32": the sweep runs with is_spin_three_half=True (the J = 3/2 rare spin, DE:22-25).
Prints one JSON line: batch sizes, sweep directory, bound-module check.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

import quantumsimulations_b200 as q  # noqa: E402
from oracle import ref_loader  # noqa: E402  (only for REFERENCE_DIR; the reference MODEL module is never loaded here)

out_root, solver = sys.argv[1], sys.argv[2]
spin32 = len(sys.argv) > 3 and sys.argv[3] == "spin32"
sys.path.insert(0, os.path.join(ROOT, "oracle", "fake_mpl"))      # inert matplotlib (absent on this image)
sys.path.insert(0, q.dropin_path())                               # the drop-in wins the import of the model module
sys.path.append(ref_loader.REFERENCE_DIR)                         # ... and only sweep_sea_detuning.py resolves here
import sweep_sea_detuning as sw  # noqa: E402  the reference's file, unmodified

assert os.path.realpath(sw.__file__).startswith(os.path.realpath(ref_loader.REFERENCE_DIR)), sw.__file__
import dipolar_ensemble_with_rare as de  # noqa: E402

bound = {
    "simulate_rare": sw.simulate_rare is q.api.simulate_rare,
    "params": sw.DipolarRareParams is q.DipolarRareParams,
    "freqs": sw.get_derived_frequencies is q.get_derived_frequencies,
    "model_module_is_dropin": os.path.realpath(de.__file__).startswith(os.path.realpath(q.dropin_path())),
    "qutip_imported": "qutip" in sys.modules,
}
calls = []


def oracle_batch(plist):
    from conftest import exact_traces

    calls.append(len(plist))
    out = []
    for p in plist:
        t = np.linspace(0.0, p.t_final, p.steps)
        obs = exact_traces(q.compile_table(p), t)
        obs["state_norm"] = np.ones_like(t)
        out.append((t, obs))
    return out


def engine_batch(plist):
    calls.append(len(plist))
    res, _ = q.simulate_rare_batch(plist)
    return res


G_SEA, G_RARE, B0_T, F1A = 8.1812e7, 6.976e7, 3.0, 50_000.0
from quantumsimulations_b200.sweep import replay_reference_script  # noqa: E402

base = replay_reference_script(
    sw, "run_sweep_sea_detuning", simulate_batch=oracle_batch if solver == "oracle" else engine_batch,
    f_Az=G_SEA * B0_T / (2 * np.pi), f1A=F1A, target_sea_detuning=F1A, gamma_sea=G_SEA, gamma_rare=G_RARE,
    sea_detunings_Hz=np.linspace(0.0, 3.0 * F1A, 13), n_sea=6, phi_sea=np.pi / 2.0, phi_rare=np.pi / 2.0,
    out_root=out_root, is_spin_three_half=spin32, solver_atol=1e-10, solver_rtol=1e-9, solver_nsteps=10_000_000,
    solver_max_step=1e-5)
print(json.dumps({"calls": calls, "base": base, "bound": bound}), flush=True)
