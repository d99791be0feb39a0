"""Envelope metrics of the detuning sweep, computed by the engine's own sweep driver.

Same definitions as the reference's helpers so that ``metrics.json`` / ``summary.json`` mean the same thing:
block average (SW:116-141), ordinary-least-squares line on the central 60 % of the coarse points with slope
standard error, t statistic and Pearson R (SW:148-268), Michelson contrast with |t| >= 1 gating (SW:276-317),
and the per-detuning directory label (SW:342-349).
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import numpy as np

SLOPE_T_MIN = 1.0
_FIT_KEYS = ("I_z_slope", "t_start", "t_end", "I_z_start", "I_z_end", "slope", "slope_std", "t_value", "R_value",
             "R2_value")


def coarse_grain(t: np.ndarray, y: np.ndarray, window: int = 25) -> Tuple[np.ndarray, np.ndarray]:
    """Means over consecutive blocks of ``window`` samples; the ragged tail is dropped."""
    n_used = (len(t) // window) * window
    if n_used == 0 or window <= 1:
        return t, y
    shape = (-1, window)
    return t[:n_used].reshape(shape).mean(axis=1), y[:n_used].reshape(shape).mean(axis=1)


def _all_nan() -> Dict[str, float]:
    return {k: float("nan") for k in _FIT_KEYS}


def iz_slope_from_coarse(t_coarse: np.ndarray, iz_coarse: np.ndarray) -> Dict[str, float]:
    """Drift of <Iz_sea> across the fitted (central 60 %) segment, with fit diagnostics."""
    n = t_coarse.size
    if n < 4 or iz_coarse.size < 4:
        return _all_nan()
    first = max(0, min(int(0.2 * n), n - 2))
    last = max(first + 2, min(int((1.0 - 0.2) * n), n))
    ts, ys = t_coarse[first:last], iz_coarse[first:last]
    if ts.size < 2:
        return _all_nan()
    slope, intercept = np.polyfit(ts, ys, 1)
    t_a, t_b = float(ts[0]), float(ts[-1])
    y_a, y_b = float(intercept + slope * t_a), float(intercept + slope * t_b)

    tc = ts - np.mean(ts)
    yc = ys - np.mean(ys)
    s_tt = float(np.sum(tc * tc))
    s_yy = float(np.sum(yc * yc))
    if s_tt > 0.0 and s_yy > 0.0:
        r_val = float(np.dot(tc, yc) / np.sqrt(s_tt * s_yy))
        r2_val = float(r_val * r_val)
    else:
        r_val = r2_val = float("nan")

    if ts.size > 2 and s_tt > 0.0:
        resid = ys - (intercept + slope * ts)
        var_b = (float(np.sum(resid**2)) / (ts.size - 2)) / s_tt
        se_b = float(np.sqrt(var_b)) if var_b > 0.0 else float("nan")
        t_val = float(slope / se_b) if (se_b > 0.0 and np.isfinite(se_b)) else float("nan")
    else:
        se_b = t_val = float("nan")

    return {
        "I_z_slope": float(y_b - y_a), "t_start": t_a, "t_end": t_b, "I_z_start": y_a, "I_z_end": y_b,
        "slope": float(slope), "slope_std": se_b, "t_value": t_val, "R_value": r_val, "R2_value": r2_val,
    }


def contrast_michelson_with_t_gate(slope_on: float, slope_off: float, t_on: float, t_off: float,
                                   t_min: float = SLOPE_T_MIN) -> float:
    """(|on| - |off|) / (|on| + |off|) with slopes whose |t| < t_min treated as zero; 0 when both vanish."""
    if not all(np.isfinite(v) for v in (slope_on, slope_off, t_on, t_off)):
        return float("nan")
    on = slope_on if abs(t_on) >= t_min else 0.0
    off = slope_off if abs(t_off) >= t_min else 0.0
    total = abs(on) + abs(off)
    if not math.isfinite(total) or total <= 1e-16:
        return 0.0
    return (abs(on) - abs(off)) / total


def detuning_label(delta_hz: float) -> str:
    """'delta_p1000.0Hz' / 'delta_m1000.0Hz'."""
    return f"delta_{delta_hz:+.1f}Hz".replace("+", "p").replace("-", "m")


def f1R_for_resonance(f1a_hz: float, delta_a_hz: float, delta_r_hz: float = 0.0) -> float:
    """Rare Rabi frequency matching the sea's effective field: sqrt(dA^2 + f1A^2 - dR^2)  (SW:1168-1194)."""
    return (delta_a_hz**2 + f1a_hz**2 - delta_r_hz**2) ** 0.5
