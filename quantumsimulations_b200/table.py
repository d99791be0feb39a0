"""Coupling-table compiler: DipolarRareParams -> compact "XYZ" spin table the CUDA engine consumes.

The reference assembles H as a 2^N x 2^N sparse matrix out of Kronecker products (DE:453-570).  The
engine never materialises H.  Every term the reference can produce is of the form

    H = sum_k ( hx_k Ix_k + hy_k Iy_k + hz_k Iz_k )
      + sum_{i<j} ( Jxx_ij Ix_i Ix_j + Jyy_ij Iy_i Iy_j + Jzz_ij Iz_i Iz_j ),      I = sigma/2,

so the table is three length-N field vectors and three strictly-upper-triangular N x N matrices
(coefficients in rad/s).  Reference mapping:

    sea detuning   DE:505-508   hz_k += delta_A                       k in sea set
    rare detuning  DE:510-512   hz_R += delta_R
    sea drive      DE:515-519   hx_k += w1A cos(phi_A), hy_k += w1A sin(phi_A)
    rare drive     DE:523-528   hx_R, hy_R likewise with w1R, phi_R
    sea-sea pair   DE:559-561   Jzz = b_ij, Jxx = -b_ij/4, Jyy = +b_ij/4     (double-quantum form)
    sea-rare pair  DE:563-568   Jzz = b_iR

Index convention (DE:43-45, appendix A.1): site 0 is the most significant tensor factor, i.e. site k
is bit N-1-k of the amplitude index; bit value 0 <-> Iz = +1/2.
"""
from __future__ import annotations

import functools

from dataclasses import dataclass, field
from typing import Dict, List

import numpy as np

from .model import (
    DipolarRareParams,
    dipolar_couplings_from_positions,
    get_derived_frequencies,
    shell_positions_with_rare_center,
)


@dataclass
class SpinTable:
    """Compiled Hamiltonian + observable + initial-state description for N spin-1/2 sites."""

    n: int
    hx: np.ndarray
    hy: np.ndarray
    hz: np.ndarray
    jxx: np.ndarray
    jyy: np.ndarray
    jzz: np.ndarray
    sea_sites: List[int] = field(default_factory=list)   # sites summed into the *_sea observables
    rare_site: int = 0                                   # site of the *_R observables
    psi0_index: int = 0                                  # computational-basis index of psi(0)

    # ---- helpers shared by host code and tests ------------------------------------------------
    def bit_of_site(self, k: int) -> int:
        return self.n - 1 - k

    @property
    def dim(self) -> int:
        return 1 << self.n

    @property
    def sea_mask(self) -> int:
        m = 0
        for k in self.sea_sites:
            m |= 1 << self.bit_of_site(k)
        return m

    @property
    def rare_mask(self) -> int:
        return 1 << self.bit_of_site(self.rare_site)

    def pairs(self):
        """Non-zero pair rows (i, j, jxx, jyy, jzz), i < j, in (i, j) lexicographic order."""
        rows = []
        for i in range(self.n):
            for j in range(i + 1, self.n):
                c = (self.jxx[i, j], self.jyy[i, j], self.jzz[i, j])
                if any(v != 0.0 for v in c):
                    rows.append((i, j) + c)
        return rows

    def bit_major(self) -> Dict[str, np.ndarray]:
        """Arrays re-indexed by BIT position (what the C ABI takes): h[3][n], J[3][n][n] symmetric."""
        n = self.n
        perm = [n - 1 - b for b in range(n)]          # bit b <- site n-1-b
        h = np.stack([self.hx[perm], self.hy[perm], self.hz[perm]]).astype(np.float64)
        J = np.zeros((3, n, n), dtype=np.float64)
        for a, m in enumerate((self.jxx, self.jyy, self.jzz)):
            full = m + m.T
            J[a] = full[np.ix_(perm, perm)]
        return {"h": np.ascontiguousarray(h), "J": np.ascontiguousarray(J)}

    def spectral_bound(self) -> float:
        """Cheap upper bound on ||H||_2 (sum of term norms); used for step-size sanity checks only."""
        s = 0.5 * float(np.sum(np.sqrt(self.hx**2 + self.hy**2)) + np.sum(np.abs(self.hz)))
        s += 0.25 * float(np.sum(np.abs(self.jxx)) + np.sum(np.abs(self.jyy)) + np.sum(np.abs(self.jzz)))
        return s


def initial_basis_index(params: DipolarRareParams) -> int:
    """Computational-basis index of the reference's product state (DE:591-606).

    Sea sites take |Iz=+1/2> (bit 0) when ``init_x_sign >= 0`` else |Iz=-1/2> (bit 1) (DE:71-72); a
    centre-rare site takes the *opposite* sign, via sign=-init_x_sign and ``sign >= 0`` -> largest
    eigenvalue (DE:99, DE:602).  The sea-centre control prepares all n_sea+1 sites like sea sites.
    """
    n_total = params.n_sea + 1
    sea_bit = 0 if params.init_x_sign >= 0 else 1
    bits = [sea_bit] * n_total
    if params.is_center_rare:
        bits[-1] = 0 if (-params.init_x_sign) >= 0 else 1
    idx = 0
    for q in bits:
        idx = (idx << 1) | q
    return idx


@functools.lru_cache(maxsize=256)
def _couplings(n_shell: int, shell_scale: float, dipolar_scale: float, gamma_sea: float, gamma_centre: float) -> np.ndarray:
    """b_ij of DE:533-545 for one geometry.  A sweep compiles hundreds of tables that differ only in drive fields and
    detunings (SW:631-668), so the geometry + couplings (the reference's own functions, bit-exact) are computed once
    per distinct argument tuple; the returned array is read-only."""
    n = n_shell + 1
    pos = shell_positions_with_rare_center(n_sea=n_shell, radius=shell_scale)
    if pos.shape != (n, 3):
        raise RuntimeError("Shell geometry returned unexpected number of sites.")
    b = dipolar_couplings_from_positions(pos, dipolar_scale, gamma_sea, gamma_centre)
    b.setflags(write=False)
    return b


def compile_table(params: DipolarRareParams) -> SpinTable:
    """Compile the reference's Hamiltonian/observable/initial-state recipe (DE:483-606) into a SpinTable."""
    if params.is_spin_three_half:
        # DE:486/DE:499-501 build operators of inconsistent dimension for spin-3/2: the
        # reference itself cannot run it.  Refuse loudly instead of silently substituting spin-1/2.
        raise NotImplementedError(
            "is_spin_three_half=True is not runnable in the reference either (operator dimensions "
            "disagree, DE:486 vs DE:499-501); pass is_spin_three_half=False."
        )
    n_shell = params.n_sea
    n = n_shell + 1
    rare = n_shell
    centre_is_rare = bool(params.is_center_rare)
    n_like_sea = n_shell if centre_is_rare else n          # DE:488-489

    fr = get_derived_frequencies(params)
    w1a, w1r = fr["omega1_sea"], fr["omega1_rare"]
    da, dr = fr["delta_sea"], fr["delta_rare"]

    hx = np.zeros(n)
    hy = np.zeros(n)
    hz = np.zeros(n)
    if params.drive_sea and da != 0.0:
        hz[:n_like_sea] += da
    if centre_is_rare and params.drive_rare and dr != 0.0:
        hz[rare] += dr
    if params.drive_sea and w1a != 0.0:
        hx[:n_like_sea] += w1a * np.cos(params.phi_sea)
        hy[:n_like_sea] += w1a * np.sin(params.phi_sea)
    if centre_is_rare and params.drive_rare and w1r != 0.0:
        hx[rare] += w1r * np.cos(params.phi_rare)
        hy[rare] += w1r * np.sin(params.phi_rare)

    b = _couplings(n_shell, float(params.shell_scale), float(params.dipolar_scale), float(params.gamma_sea),
                   float(params.gamma_rare if centre_is_rare else params.gamma_sea))

    jzz = np.triu(b, 1)                      # every pair has the Ising term (sea-rare pairs ONLY that one)
    jyy = np.zeros((n, n))
    jyy[:n_like_sea, :n_like_sea] = 0.25 * jzz[:n_like_sea, :n_like_sea]
    jxx = -jyy

    return SpinTable(
        n=n,
        hx=hx,
        hy=hy,
        hz=hz,
        jxx=jxx,
        jyy=jyy,
        jzz=jzz,
        sea_sites=list(range(n_like_sea)),
        rare_site=rare,
        psi0_index=initial_basis_index(params),
    )
