"""Host-side grid construction for the real-gas path: the wavenumber grid, the spectral bands
and the fixed-size pressure grid that ``RealGas.__init__`` derives before any flux is computed
(real_gas.py:300-334 get_wavenumber_array, :336-402 get_wavenumber_bands, :493-503 the integer-nz
branch of get_p_grid).  One-off O(n_nu) set-up work, NumPy on the host (SURVEY.md section 2,
row 3).  The Planck values behind the band / grid edges are evaluated in host NumPy with exactly
the reference's expression (``B_freq`` / ``B_wavenumber``, real_gas.py:14-42: ``freq ** 3``, the
same operation order), because the edges are arg-min decisions: a last-bit difference of another
Planck routine (the CUDA one cubes with two multiplications) could flip a near-tie and move an edge
by one bin.  The CUDA Planck function stays on the hot path only.

Results must match the reference *exactly* (band edges are argmin decisions), which the tests
check against the golden cases.
"""
from math import ceil

import numpy as np


def planck_wavenumber_host(nu, T):
    """B(nu, T) per wavenumber (cm^-1), W m^-2 / cm^-1 -- the reference's B_wavenumber (real_gas.py:30-42)
    on B_freq (:14-27) written out in NumPy, operation for operation."""
    from ..constants import h_planck, speed_of_light, k_boltzmann
    freq = 100 * np.asarray(nu, dtype=float) * speed_of_light
    B = 2 * h_planck * freq ** 3 / speed_of_light ** 2 / (np.exp(h_planck * freq / (k_boltzmann * T)) - 1)
    return 100 * speed_of_light * B


def _default_planck():
    return planck_wavenumber_host


def fixed_pressure_grid(nz, p_surface, p_toa):
    """Interfaces for an integer ``nz`` (real_gas.py:434, :493-503): log-pressure spacing that
    shrinks towards the surface.  Returned ascending (index 0 = top of atmosphere), which is the
    order RealGas stores after its np.sort (real_gas.py:281)."""
    n_fine = int(nz * 1000)
    log_p = np.log10(np.logspace(np.log10(p_surface), np.log10(p_toa), n_fine))
    top, bottom = log_p[0], log_p[-1]
    alpha = np.log10(top - bottom + 1) / (nz - 1)
    log_levels = top + 1 - 10 ** (alpha * np.arange(nz))
    if log_levels[-1] != bottom:
        raise ValueError('Too few grid points to cover pressure grid')
    return np.sort(10 ** log_levels)


def _cdf_crossing(nu, B, level):
    """Wavenumber at which the normalised running sum of B is closest to ``level``."""
    frac = np.cumsum(B) / sum(B)
    return nu[np.abs(frac - level).argmin()]


def wavenumber_array(d_nu, T_star, T_g, fract_to_ignore=0.001, fract_to_ignore_overlap=0.001, planck=None):
    """(nu, nu_lw, nu_overlap, nu_sw) as real_gas.py:300-334: the range holding all but
    ``fract_to_ignore`` of the planetary (low end) and stellar (high end) Planck energy, split
    where the planetary spectrum stops / the stellar spectrum starts to matter."""
    planck = planck or _default_planck()
    trial = np.arange(10.0, 100000.0 + d_nu, d_nu)
    B_star = planck(trial, T_star)
    B_planet = planck(trial, T_g)
    nu_max = _cdf_crossing(trial, B_star, 1 - fract_to_ignore)
    nu_min = _cdf_crossing(trial, B_planet, fract_to_ignore)
    sw_start = _cdf_crossing(trial, B_star, fract_to_ignore_overlap)
    lw_end = _cdf_crossing(trial, B_planet, 1 - fract_to_ignore_overlap)
    nu = np.arange(nu_min, nu_max + d_nu, d_nu)
    overlap = nu[np.logical_and(nu <= lw_end, nu >= sw_start)]
    return nu, nu[nu <= lw_end], overlap, nu[nu >= lw_end]


def _equal_energy_bands(nu, B, n_bands):
    """Split ``nu`` into ``n_bands`` runs over which a monotone remap of B rises by equal steps
    (get_equal_bands, real_gas.py:346-363).  Adjacent bands share their end-point bin."""
    shape = B / max(B)
    peak = shape.argmax()
    shape[peak:] = 1 + (1 - shape[peak:])          # fold the falling side upwards
    shape = shape - min(shape)
    shape = shape / max(shape)
    targets = np.linspace(0, 1, n_bands + 1)[1:]
    ranges, centre, delta = [], np.zeros(len(targets)), np.zeros(len(targets))
    first = 0
    for i, target in enumerate(targets):
        last = max(np.abs(shape - target).argmin(), first + 1)
        run = nu[first:last + 1]
        ranges.append(run)
        centre[i] = run[round((len(run) + 1) / 2) - 1]
        delta[i] = run[-1] - run[0]
        first = last
    return ranges, centre, delta


def wavenumber_bands(nu_lw, nu_overlap, nu_sw, n_nu_bands, T_star, T_g, planck=None):
    """The ``nu_bands`` dict of real_gas.py:336-388: long-wave-only, overlap and short-wave band
    groups, each of equal (remapped) Planck energy; 'sw' marks bands whose emission integrals are
    skipped (centre above the start of nu_sw)."""
    planck = planck or _default_planck()
    B_star = planck(nu_sw, T_star)
    lw_only = np.setdiff1d(nu_lw, nu_overlap)
    B_planet = planck(lw_only, T_g)
    B_ov_planet = planck(nu_overlap, T_g)
    B_ov_star = planck(nu_overlap, T_star)
    share_planet = (1 - np.sum(B_planet) / (np.sum(B_planet) + np.sum(B_ov_planet))) * n_nu_bands / 2
    share_star = (1 - (np.sum(B_star) / (np.sum(B_star) + np.sum(B_ov_star)))) * n_nu_bands / 2
    n_overlap = ceil(share_planet + share_star)
    n_lw = ceil(n_nu_bands / 2 - share_planet)
    n_sw = n_nu_bands - n_lw - n_overlap
    lw = _equal_energy_bands(lw_only, B_planet, n_lw)
    sw = _equal_energy_bands(nu_sw, B_star, n_sw)
    ov_planet = B_ov_planet / max(B_planet)
    ov_star = B_ov_star / max(B_star)
    if max(ov_star) == 1 or max(ov_planet) == 1:
        raise ValueError('Peak of planet or star spectrum is in overlap region')
    combined = ov_planet + ov_star[0] - (ov_star - ov_star[0])
    ov = _equal_energy_bands(nu_overlap, combined, n_overlap)
    bands = {'range': lw[0] + ov[0] + sw[0],
             'centre': np.concatenate((lw[1], ov[1], sw[1])),
             'delta': np.concatenate((lw[2], ov[2], sw[2])),
             'sw': np.ones(n_nu_bands, dtype=bool)}
    bands['sw'][bands['centre'] <= nu_sw.min()] = False
    return bands
