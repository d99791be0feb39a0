"""Restatement of the sweep's envelope metrics.  TEST INFRASTRUCTURE ONLY.

coarse_grain SW:116-141, iz_slope_from_coarse SW:148-268, SLOPE_T_MIN SW:276,
contrast_michelson_with_t_gate SW:279-317, detuning_label SW:342-349.  Written from the formulas
(block mean; OLS line on the central 60 %; t = b / se(b); Pearson R; t-gated Michelson contrast) and
checked against the reference functions themselves in tests/test_oracle_vs_reference.py.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import numpy as np

T_GATE = 1.0
_KEYS = ("I_z_slope", "t_start", "t_end", "I_z_start", "I_z_end", "slope", "slope_std", "t_value", "R_value", "R2_value")


def block_mean(t: np.ndarray, y: np.ndarray, window: int) -> Tuple[np.ndarray, np.ndarray]:
    usable = (len(t) // window) * window
    if usable == 0 or window <= 1:
        return t, y
    return t[:usable].reshape(-1, window).mean(axis=1), y[:usable].reshape(-1, window).mean(axis=1)


def central_fit(tc: np.ndarray, yc: np.ndarray) -> Dict[str, float]:
    nan = {k: float("nan") for k in _KEYS}
    n = tc.size
    if n < 4 or yc.size < 4:
        return nan
    lo = int(0.2 * n)
    hi = int((1.0 - 0.2) * n)
    lo = max(0, min(lo, n - 2))
    hi = max(lo + 2, min(hi, n))
    ts, ys = tc[lo:hi], yc[lo:hi]
    if ts.size < 2:
        return nan
    slope, icpt = np.polyfit(ts, ys, 1)
    t0, t1 = float(ts[0]), float(ts[-1])
    y0, y1 = float(icpt + slope * t0), float(icpt + slope * t1)
    dt = ts - np.mean(ts)
    dy = ys - np.mean(ys)
    stt = float(np.sum(dt * dt))
    syy = float(np.sum(dy * dy))
    if stt > 0.0 and syy > 0.0:
        r = float(np.dot(dt, dy) / np.sqrt(stt * syy))
        r2 = float(r * r)
    else:
        r = r2 = float("nan")
    if ts.size > 2 and stt > 0.0:
        res = ys - (icpt + slope * ts)
        var = (float(np.sum(res**2)) / (ts.size - 2)) / stt
        se = float(np.sqrt(var)) if var > 0.0 else float("nan")
        tval = float(slope / se) if (se > 0.0 and np.isfinite(se)) else float("nan")
    else:
        se = tval = float("nan")
    return {
        "I_z_slope": float(y1 - y0), "t_start": t0, "t_end": t1, "I_z_start": y0, "I_z_end": y1,
        "slope": float(slope), "slope_std": se, "t_value": tval, "R_value": r, "R2_value": r2,
    }


def gated_contrast(s_on: float, s_off: float, t_on: float, t_off: float, gate: float = T_GATE) -> float:
    if not (np.isfinite(s_on) and np.isfinite(s_off) and np.isfinite(t_on) and np.isfinite(t_off)):
        return float("nan")
    a = 0.0 if abs(t_on) < gate else s_on
    b = 0.0 if abs(t_off) < gate else s_off
    den = abs(a) + abs(b)
    if not math.isfinite(den) or den <= 1e-16:
        return 0.0
    return (abs(a) - abs(b)) / den


def label_for(delta_hz: float) -> str:
    return f"delta_{delta_hz:+.1f}Hz".replace("+", "p").replace("-", "m")
