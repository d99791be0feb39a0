"""Batched detuning sweep: same inputs and on-disk outputs as the reference's ``run_sweep_sea_detuning``
(sweep_sea_detuning.py:356-1165), with all 3 x n_det trajectories integrated concurrently on the GPU(s).

Files written under ``{out_root}/sea_detuning_sweep_{timestamp}/``:
geometry_and_couplings.npz, global_params.json, summary.json and, per detuning, ``delta_{p|m}X.XHz/`` with
time_and_obs_{center_off|center_on|shell_off}.npz, params_{tag}.json, freqs_{tag}.json, metrics.json.
The reference also draws a PDF/PNG report with matplotlib; that plotting layer is outside the hot path and
matplotlib is not installed here, so no figures are produced (the reprocess / 2D-report scripts only read
the npz/json files).

``replay_reference_script`` is the other way to get batching: it runs the reference's *unchanged* sweep
function twice (record -> batch on the GPU -> replay from cache), see SURVEY.md 7.4.
"""
from __future__ import annotations

import datetime as _dt
import json
import os
import time
from dataclasses import asdict, replace
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np

from . import metrics as M
from .model import (
    DipolarRareParams,
    dipolar_couplings_from_positions,
    get_derived_frequencies,
    shell_positions_with_rare_center,
)

VARIANTS = ("center_off", "center_on", "shell_off")
BatchFn = Callable[[Sequence[DipolarRareParams]], List]


def _dump(path: str, obj: Any) -> None:
    with open(path, "w", encoding="utf-8") as f:
        json.dump(obj, f, indent=2, default=float)


def sweep_point_params(*, n_sea, gamma_sea, gamma_rare, B0_common, B1_sea, B1_rare, omega_rf_sea, omega_rf_rare,
                       phi_sea, phi_rare, dipolar_scale, shell_scale, t_final, steps, is_spin_three_half,
                       solver_atol, solver_rtol, solver_nsteps, solver_max_step) -> Dict[str, DipolarRareParams]:
    """The three variants the reference runs per detuning (SW:631-668)."""
    base = DipolarRareParams(
        n_sea=n_sea, gamma_sea=gamma_sea, gamma_rare=gamma_rare, B0_sea=B0_common, B0_rare=B0_common,
        B1_sea=B1_sea, B1_rare=B1_rare, omega_rf_sea=omega_rf_sea, omega_rf_rare=omega_rf_rare,
        phi_sea=phi_sea, phi_rare=phi_rare, dipolar_scale=dipolar_scale, shell_scale=shell_scale,
        t_final=t_final, steps=steps, drive_sea=True, drive_rare=False, init_x_sign=-1, init_rare_level=3,
        is_spin_three_half=is_spin_three_half, is_center_rare=True,
        solver_atol=solver_atol, solver_rtol=solver_rtol, solver_nsteps=solver_nsteps, solver_max_step=solver_max_step,
    )
    return {
        "center_off": replace(base, drive_rare=False, is_center_rare=True),
        "center_on": replace(base, drive_rare=True, is_center_rare=True),
        "shell_off": replace(base, drive_rare=False, is_center_rare=False),
    }


def _with_cache(batch: BatchFn, cache_dir: Optional[str], method: Optional[str] = None) -> BatchFn:
    if not cache_dir:
        return batch
    from .checkpoint import TrajectoryCache, cached_batch

    solver = {"method": method} if method else {}  # part of the cache key: different numerics, different entry
    return cached_batch(batch, TrajectoryCache(cache_dir), **solver)


def _default_batch(devices: Optional[Sequence[int]], method: Optional[str] = None) -> BatchFn:
    from .dist import simulate_on_devices

    kw = {"method": method} if method else {}
    return lambda plist: simulate_on_devices(plist, devices, **kw)


def run_sweep_sea_detuning(
    *,
    f_Az: float,
    f1A: float,
    target_sea_detuning: float,
    gamma_sea: float,
    gamma_rare: float,
    sea_detunings_Hz: Sequence[float],
    n_sea: int = 12,
    t_final: float = 3.0e-2,
    steps: int = 2000,
    phi_sea: float = 0.0,
    phi_rare: float = 0.0,
    out_root: str = "results",
    is_spin_three_half: bool = False,
    solver_atol: Optional[float] = None,
    solver_rtol: Optional[float] = None,
    solver_nsteps: Optional[int] = None,
    solver_max_step: Optional[float] = None,
    coarse_window: int = 50,
    devices: Optional[Sequence[int]] = None,
    simulate_batch: Optional[BatchFn] = None,
    timestamp: Optional[str] = None,
    verbose: bool = True,
    cache_dir: Optional[str] = None,
    method: Optional[str] = None,
    _defer: bool = False,
):
    """Reference keyword arguments (SW:356-376) + ``devices`` (GPUs to spread trajectories over),
    ``simulate_batch`` (override of the batch solver: list[params] -> list[(t, obs)]), ``timestamp`` and
    ``cache_dir`` (finished trajectories are kept there, keyed by their parameters, and an interrupted or repeated
    sweep integrates only what is missing; the reference starts every run from zero, SW:483-485) and ``method``
    (integrator of the default batch solver: "taylor", "dp5", "propagator" ...; see ``api.simulate_rare``).
    Returns the sweep directory.  (``_defer``: return ``(plist, finish)`` instead of integrating, so that
    ``run_multi_sweep`` can put the trajectories of many sweeps into one batch.)"""
    say = print if verbose else (lambda *a, **k: None)
    f1R = M.f1R_for_resonance(f1A, target_sea_detuning, 0.0)
    dets = np.asarray(sea_detunings_Hz, dtype=float)

    # fields from frequencies (SW:421-429) and the physical dipolar scale (SW:434-437)
    B0_common = 2 * np.pi * f_Az / gamma_sea
    f_Rz = gamma_rare * B0_common / (2 * np.pi)
    B1_sea = 2 * np.pi * f1A / gamma_sea
    B1_rare = 2 * np.pi * f1R / gamma_rare if gamma_rare != 0.0 else 0.0
    dipolar_scale_SI = 1.0e-7 * 1.054571817e-34
    shell_scale = 0.282393e-9

    positions = shell_positions_with_rare_center(n_sea=n_sea, radius=shell_scale)
    idx_rare = n_sea
    b = dipolar_couplings_from_positions(positions=positions, scale=dipolar_scale_SI, gamma_sea=gamma_sea,
                                         gamma_rare=gamma_rare)
    sea_idx = list(range(n_sea))
    sea_rare_vals = np.array([b[i, idx_rare] for i in sea_idx], dtype=float)
    sea_rare_abs_Hz = np.abs(sea_rare_vals) / (2 * np.pi)
    sea_rare_rms_Hz = np.sqrt(np.mean(np.abs(sea_rare_vals) ** 2)) / (2 * np.pi)
    sea_sea_vals = np.array([b[i, j] for i in sea_idx for j in sea_idx if j > i], dtype=float)
    sea_sea_abs_Hz = np.abs(sea_sea_vals) / (2 * np.pi)
    sea_sea_rms_Hz = np.sqrt(np.mean(np.abs(sea_sea_vals) ** 2)) / (2 * np.pi)

    stamp = timestamp or _dt.datetime.now().strftime("%Y%m%d_%H%M%S")
    base_dir = os.path.join(out_root, f"sea_detuning_sweep_{stamp}")
    os.makedirs(base_dir, exist_ok=True)
    np.savez(os.path.join(base_dir, "geometry_and_couplings.npz"), positions=positions, b=b,
             sea_indices=np.array(sea_idx, dtype=int), idx_rare=int(idx_rare), sea_rare_vals=sea_rare_vals,
             sea_sea_vals=sea_sea_vals)

    summary: Dict[str, Any] = {"global_params": {}, "sweep_results": []}
    summary["global_params"] = {
        "f_Az_Hz": float(f_Az), "f_Rz_Hz": float(f_Rz), "f1A_Hz": float(f1A), "f1R_Hz": float(f1R),
        "gamma_sea": float(gamma_sea), "gamma_rare": float(gamma_rare), "B0_common_T": float(B0_common),
        "B1_sea_T": float(B1_sea), "B1_rare_T": float(B1_rare), "dipolar_scale_SI": float(dipolar_scale_SI),
        "shell_scale_m": float(shell_scale), "t_final_s": float(t_final), "steps": int(steps), "n_sea": int(n_sea),
        "phi_sea_rad": float(phi_sea), "phi_rare_rad": float(phi_rare),
        "sea_detunings_Hz": [float(x) for x in dets], "sea_spin_type": "1/2",
        "rare_spin_type": "3/2" if is_spin_three_half else "1/2",
        "solver_atol": solver_atol, "solver_rtol": solver_rtol, "solver_nsteps": solver_nsteps,
        "solver_max_step": solver_max_step, "target_sea_detuning": target_sea_detuning,
        "coarse_window": int(coarse_window),
        "avg_b_AR_Hz": float(sea_rare_abs_Hz.mean()), "rms_b_AR_Hz": float(sea_rare_rms_Hz),
        "avg_b_AA_Hz": float(sea_sea_abs_Hz.mean()), "rms_b_AA_Hz": float(sea_sea_rms_Hz),
    }

    # ---- all trajectories of the sweep in one batch ------------------------------------------------------
    plist: List[DipolarRareParams] = []
    per_det: List[Dict[str, DipolarRareParams]] = []
    for delta_Hz in dets:
        f_rf_sea = f_Az - delta_Hz
        variants = sweep_point_params(
            n_sea=n_sea, gamma_sea=gamma_sea, gamma_rare=gamma_rare, B0_common=B0_common, B1_sea=B1_sea,
            B1_rare=B1_rare, omega_rf_sea=2 * np.pi * f_rf_sea, omega_rf_rare=2 * np.pi * f_Rz, phi_sea=phi_sea,
            phi_rare=phi_rare, dipolar_scale=dipolar_scale_SI, shell_scale=shell_scale, t_final=t_final, steps=steps,
            is_spin_three_half=is_spin_three_half, solver_atol=solver_atol, solver_rtol=solver_rtol,
            solver_nsteps=solver_nsteps, solver_max_step=solver_max_step)
        per_det.append(variants)
        plist.extend(variants[tag] for tag in VARIANTS)
    say(f"Sea detuning sweep: {len(dets)} detunings x {len(VARIANTS)} variants = {len(plist)} trajectories "
        f"(N = {n_sea + 1} spins) -> {base_dir}", flush=True)

    def finish(results) -> str:
        return _write_sweep(results, base_dir=base_dir, dets=dets, per_det=per_det, summary=summary, f_Az=f_Az, f1A=f1A,
                            f1R=f1R, sea_rare_rms_Hz=sea_rare_rms_Hz, coarse_window=coarse_window, say=say)

    if _defer:
        return plist, finish
    t0 = time.perf_counter()
    results = _with_cache(simulate_batch or _default_batch(devices, method), cache_dir, method)(plist)
    say(f"  integrated in {time.perf_counter() - t0:.2f} s", flush=True)
    return finish(results)


def _write_sweep(results, *, base_dir, dets, per_det, summary, f_Az, f1A, f1R, sea_rare_rms_Hz, coarse_window, say) -> str:
    """Per-detuning files and metrics (SW:671-792), global_params.json and summary.json (SW:1153-1157)."""
    for i, delta_Hz in enumerate(dets):
        det_dir = os.path.join(base_dir, M.detuning_label(delta_Hz))
        os.makedirs(det_dir, exist_ok=True)
        fits = {}
        for j, tag in enumerate(VARIANTS):
            t_arr, obs = results[3 * i + j]
            params = per_det[i][tag]
            np.savez(os.path.join(det_dir, f"time_and_obs_{tag}.npz"), t=t_arr, **obs)
            _dump(os.path.join(det_dir, f"params_{tag}.json"), asdict(params))
            _dump(os.path.join(det_dir, f"freqs_{tag}.json"), get_derived_frequencies(params))
            fits[tag] = M.iz_slope_from_coarse(*M.coarse_grain(t_arr, obs["Iz_sea"], window=coarse_window))
        off, on, ctl = fits["center_off"], fits["center_on"], fits["shell_off"]
        Omega_A = np.sqrt(delta_Hz**2 + f1A**2)
        Omega_R = np.sqrt(0.0**2 + f1R**2)
        DeltaOmega = Omega_A - Omega_R
        sin_A = f1A / Omega_A if Omega_A != 0.0 else 0.0
        sin_R = f1R / Omega_R if Omega_R != 0.0 else 0.0
        g_eff_Hz = (sea_rare_rms_Hz / 4.0) * sin_A * sin_R
        ratio = float("nan") if (g_eff_Hz == 0.0 or np.isnan(g_eff_Hz)) else float(DeltaOmega / abs(g_eff_Hz))
        row = {
            "delta_Hz": float(delta_Hz), "f_rf_sea_Hz": float(f_Az - delta_Hz),
            "I_z_slope_off_center": float(off["I_z_slope"]), "R_off_center": float(off["R_value"]),
            "t_off_center": float(off["t_value"]),
            "I_z_slope_on_center": float(on["I_z_slope"]), "R_on_center": float(on["R_value"]),
            "t_on_center": float(on["t_value"]),
            "contrast_rare_center": float(M.contrast_michelson_with_t_gate(
                on["I_z_slope"], off["I_z_slope"], on["t_value"], off["t_value"])),
            "I_z_slope_off_sea_center": float(ctl["I_z_slope"]), "R_off_sea_center": float(ctl["R_value"]),
            "t_off_sea_center": float(ctl["t_value"]),
            "contrast_sea_center": float(M.contrast_michelson_with_t_gate(
                on["I_z_slope"], ctl["I_z_slope"], on["t_value"], ctl["t_value"])),
            "DeltaOmega_Hz": float(DeltaOmega), "g_eff_Hz": float(g_eff_Hz), "DeltaOmega_over_geff": float(ratio),
        }
        _dump(os.path.join(det_dir, "metrics.json"), row)
        summary["sweep_results"].append(row)

    _dump(os.path.join(base_dir, "global_params.json"), summary["global_params"])
    _dump(os.path.join(base_dir, "summary.json"), summary)
    say(f"Sweep complete: {base_dir}", flush=True)
    return base_dir


def run_multi_sweep(
    *,
    f1A_values: Sequence[float],
    detunings_for: Optional[Callable[[float], Sequence[float]]] = None,
    out_root: str = "results",
    devices: Optional[Sequence[int]] = None,
    simulate_batch: Optional[BatchFn] = None,
    timestamp: Optional[str] = None,
    verbose: bool = True,
    cache_dir: Optional[str] = None,
    method: Optional[str] = None,
    **sweep_kwargs,
) -> List[str]:
    """The 2D grid that ``2D_sweep_report.py`` consumes: one detuning sweep per sea drive amplitude f1A, each
    with ``target_sea_detuning = f1A`` and detunings ``linspace(0, 3 f1A, 13)`` (the rule of SW:1231, override with
    ``detunings_for``), written as one ``sea_detuning_sweep_*`` directory per f1A under ``out_root``.

    The reference would run the sweeps (and every trajectory inside them) one after the other; here ALL
    trajectories of ALL sweeps go through one batch call, so the grid fills the GPU(s): 19 x 13 x 3 = 741
    one-CTA trajectories are five waves on one B200 and less than one on eight.
    ``sweep_kwargs`` are the remaining keyword arguments of ``run_sweep_sea_detuning`` (f_Az, gamma_sea, ...).
    """
    say = print if verbose else (lambda *a, **k: None)
    rule = detunings_for or (lambda f1a: np.linspace(0.0, 3.0 * f1a, 13))
    t_base = _dt.datetime.strptime(timestamp, "%Y%m%d_%H%M%S") if timestamp else _dt.datetime.now()
    plist: List[DipolarRareParams] = []
    parts = []
    for k, f1a in enumerate(f1A_values):
        stamp = (t_base + _dt.timedelta(seconds=k)).strftime("%Y%m%d_%H%M%S")  # one directory per sweep
        sub, finish = run_sweep_sea_detuning(
            f1A=float(f1a), target_sea_detuning=float(f1a), sea_detunings_Hz=rule(float(f1a)), out_root=out_root,
            timestamp=stamp, verbose=False, _defer=True, **sweep_kwargs)
        parts.append((len(plist), len(sub), finish))
        plist.extend(sub)
    say(f"2D grid: {len(parts)} sweeps, {len(plist)} trajectories in one batch", flush=True)
    t0 = time.perf_counter()
    results = _with_cache(simulate_batch or _default_batch(devices, method), cache_dir, method)(plist)
    say(f"  integrated in {time.perf_counter() - t0:.2f} s", flush=True)
    return [finish(results[a:a + n]) for a, n, finish in parts]


def replay_reference_script(module, func_name: str = "run_sweep_sea_detuning", *, devices=None,
                            simulate_batch: Optional[BatchFn] = None, **kwargs) -> str:
    """Run an UNCHANGED reference sweep function with its serial simulate_rare calls batched.

    ``module`` is the imported reference script (it must have bound ``simulate_rare`` at import, SW:103-109).
    Pass 1 records every DipolarRareParams against placeholder traces in a scratch directory, the engine then
    integrates all of them at once, pass 2 re-runs the script with simulate_rare served from that cache.
    """
    import shutil
    import tempfile

    func = getattr(module, func_name)
    original = module.simulate_rare
    recorded: List[DipolarRareParams] = []

    def recorder(p):
        recorded.append(replace(p))
        t = np.linspace(0.0, p.t_final, p.steps)
        z = np.zeros_like(t)
        return t, {k: z.copy() for k in ("Ix_sea", "Iy_sea", "Iz_sea", "Iz_R", "Ix_R", "Iy_R", "state_norm")}

    scratch = tempfile.mkdtemp(prefix="qsb200_record_")
    try:
        module.simulate_rare = recorder
        func(**{**kwargs, "out_root": scratch})
        results = (simulate_batch or _default_batch(devices))(recorded)
        cache = {json.dumps(asdict(p), sort_keys=True, default=float): r for p, r in zip(recorded, results)}

        def cached(p):
            return cache[json.dumps(asdict(p), sort_keys=True, default=float)]

        module.simulate_rare = cached
        return func(**kwargs)
    finally:
        module.simulate_rare = original
        shutil.rmtree(scratch, ignore_errors=True)
