"""Host-side model layer: parameter record, derived frequencies, shell geometry, dipolar couplings.

Mirrors the reference's model surface (names, argument meaning, error behaviour) so callers written
against ``dipolar_ensemble_with_rare.py`` keep working:

* ``DipolarRareParams``                    -> reference DE:307-384
* ``get_derived_frequencies``              -> reference DE:387-450
* ``shell_positions_with_rare_center``     -> reference DE:205-251 (+ vertex tables DE:107-202)
* ``dipolar_couplings_from_positions``     -> reference DE:255-299

Everything here is pure float64 numpy.  The expressions are evaluated in the same order as the
reference evaluates them, because downstream coefficients are differences of ~2.5e8 rad/s numbers
 and the sweep's json files store these floats verbatim.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np

_TWO_PI = 2 * np.pi


@dataclass
class DipolarRareParams:
    """Physical + solver knobs of one trajectory (field order and defaults follow DE:331-384).

    ``asdict(params)`` is what the sweep serialises to ``params_{tag}.json`` (SW:683), so the field
    names and their order are part of the on-disk contract.
    """

    n_sea: int = 12
    gamma_sea: float = 1.0
    gamma_rare: float = 1.0
    B0_sea: float = 1.0
    B0_rare: float = 1.0
    B1_sea: float = 0.0
    B1_rare: float = 0.0
    omega_rf_sea: Optional[float] = None
    omega_rf_rare: Optional[float] = None
    phi_sea: float = 0.0
    phi_rare: float = 0.0
    dipolar_scale: float = 2 * np.pi
    shell_scale: float = 0.282393
    t_final: float = 0.02
    steps: int = 2_000
    drive_sea: bool = False
    drive_rare: bool = False
    init_x_sign: int = -1
    init_rare_level: int = 3          # never read by the reference either
    is_spin_three_half: bool = True   # J = 3/2 rare spin: runs here (table.py, two index bits); not in the reference
    is_center_rare: bool = True
    solver_atol: Optional[float] = None
    solver_rtol: Optional[float] = None
    solver_nsteps: Optional[int] = None
    solver_max_step: Optional[float] = None


def get_derived_frequencies(params: DipolarRareParams) -> Dict[str, float]:
    """Larmor / Rabi / carrier frequencies and detunings (rad/s and Hz); 16 keys, order as DE:431-450.

    A species' detuning is non-zero only while its drive flag is set (DE:417-425).
    """
    w_az = params.gamma_sea * params.B0_sea
    w_rz = params.gamma_rare * params.B0_rare
    w1_a = params.gamma_sea * params.B1_sea
    w1_r = params.gamma_rare * params.B1_rare
    w_rf_a = w_az if params.omega_rf_sea is None else params.omega_rf_sea
    w_rf_r = w_rz if params.omega_rf_rare is None else params.omega_rf_rare
    d_a = (w_az - w_rf_a) if params.drive_sea else 0.0
    d_r = (w_rz - w_rf_r) if params.drive_rare else 0.0

    ang = {
        "omega_Az": w_az,
        "omega_Rz": w_rz,
        "omega1_sea": w1_a,
        "omega1_rare": w1_r,
        "omega_rf_sea": w_rf_a,
        "omega_rf_rare": w_rf_r,
        "delta_sea": d_a,
        "delta_rare": d_r,
    }
    hz_names = ("f_Az", "f_Rz", "f1_sea", "f1_rare", "f_rf_sea", "f_rf_rare", "delta_sea_Hz", "delta_rare_Hz")
    out = dict(ang)
    for name, w in zip(hz_names, ang.values()):
        out[name] = w / _TWO_PI
    return out


def _unit_polyhedron(n_sea: int) -> np.ndarray:
    """Unit-sphere vertices of the Platonic solid with ``n_sea`` vertices, in the reference's row order.

    n_sea in {4, 6, 8, 12, 20}; anything else raises ValueError (the caller falls back to a Fibonacci
    sphere).  Row order matters: site k maps to qubit bit N-1-k, so it is part of the parity contract.
    """
    g = (1.0 + np.sqrt(5.0)) / 2.0
    ig = 1.0 / g
    if n_sea == 4:
        rows = [(1.0, 1.0, 1.0), (-1.0, -1.0, 1.0), (-1.0, 1.0, -1.0), (1.0, -1.0, -1.0)]
    elif n_sea == 6:
        rows = []
        for axis in range(3):
            for sgn in (1.0, -1.0):
                v = [0.0, 0.0, 0.0]
                v[axis] = sgn
                rows.append(tuple(v))
    elif n_sea == 8:
        rows = [(x, y, z) for x in (1.0, -1.0) for y in (1.0, -1.0) for z in (1.0, -1.0)]
    elif n_sea == 12:
        rows = []
        # three golden rectangles: (0, ±1, ±g) with z-sign outer, then cyclic permutations
        rows += [(0.0, y, z) for z in (g, -g) for y in (1.0, -1.0)]
        rows += [(x, y, 0.0) for y in (g, -g) for x in (1.0, -1.0)]
        rows += [(x, 0.0, z) for x in (g, -g) for z in (1.0, -1.0)]
    elif n_sea == 20:
        rows = [(x, y, z) for x in (-1.0, 1.0) for y in (-1.0, 1.0) for z in (-1.0, 1.0)]
        rows += [(0.0, y, z) for y in (-ig, ig) for z in (-g, g)]
        rows += [(x, y, 0.0) for x in (-ig, ig) for y in (-g, g)]
        rows += [(x, 0.0, z) for x in (-g, g) for z in (-ig, ig)]
    else:
        raise ValueError(f"No Platonic solid with {n_sea} vertices.")
    pts = np.array(rows, dtype=float)
    return pts / np.linalg.norm(pts, axis=1, keepdims=True)


def shell_positions_with_rare_center(n_sea: int, radius: float = 0.282393) -> np.ndarray:
    """(n_sea+1, 3) positions: sea sites on a shell of ``radius``, centre site (0,0,0) appended last.

    Platonic vertices when they exist, otherwise the Fibonacci-sphere fallback of DE:232-247.
    """
    if n_sea < 1:
        raise ValueError("n_sea must be at least 1.")
    try:
        shell = radius * _unit_polyhedron(n_sea)
    except ValueError:
        shell = np.zeros((n_sea, 3), dtype=float)
        g = (1.0 + np.sqrt(5.0)) / 2.0
        for i in range(n_sea):
            y = 1.0 - 2.0 * (i + 0.5) / n_sea
            rho = np.sqrt(max(0.0, 1.0 - y * y))
            ang = 2.0 * np.pi * i / g
            shell[i] = radius * np.array([rho * np.cos(ang), y, rho * np.sin(ang)], dtype=float)
    return np.vstack([shell, np.zeros((1, 3), dtype=float)])


def dipolar_couplings_from_positions(
    positions: np.ndarray,
    scale: float,
    gamma_sea: float,
    gamma_rare: float,
) -> np.ndarray:
    """Secular dipolar couplings b_ij = g_i g_j * scale * (1 - 3 cos^2(theta_ij)) / r_ij^3  (DE:255-299).

    theta_ij is the angle of r_i - r_j to z.  The LAST site carries ``gamma_rare``.  Symmetric, zero
    diagonal.  Raises ValueError when two sites coincide.
    """
    n = positions.shape[0]
    b = np.zeros((n, n), dtype=float)
    last = n - 1
    for i in range(n):
        for j in range(i + 1, n):
            d = positions[i] - positions[j]
            dist = np.linalg.norm(d)
            if dist == 0:
                raise ValueError("Two sites have identical positions.")
            c = d[2] / dist
            shape = (1.0 - 3.0 * c**2) / (dist**3)
            g_i = gamma_rare if i == last else gamma_sea
            g_j = gamma_rare if j == last else gamma_sea
            b[i, j] = b[j, i] = g_i * g_j * scale * shape
    return b
