"""Synthetic inputs for benchmarks and smoke tests: the reference ships only
a 10 cm^-1 CO2 table (the four headline tables are stripped), so larger spectral grids need
generated tables in the same dict format, and "1M columns" needs generated atmospheres.

Everything here is host-side NumPy input generation with fixed seeds -- no flux arithmetic.
"""
import numpy as np

from . import constants as k

T_REFERENCE = 296.0          # hitran.py:31
P_REFERENCE = k.p_one_atmosphere


def lorentz_table(nu, p, T, lines, n_line_widths=1000):
    """k[p, T, nu] (m^2/kg) as a sum of pressure-broadened Lorentz lines, the model of
    hitran.py:138-247 (gamma ~ p (T0/T)^n, S ~ (T/T0)^n exp(-h c nu0/k (1/T - 1/T0)), each line
    kept within ``n_line_widths`` widths of its centre).  ``lines``: dict with 'nu', 'sw',
    'gamma_air', 'n_air' arrays."""
    nu = np.asarray(nu, dtype=float)
    p = np.asarray(p, dtype=float)
    T = np.asarray(T, dtype=float)
    d_nu = nu[1] - nu[0]
    out = np.zeros((p.size, T.size, nu.size))
    pp, TT = np.meshgrid(p, T, indexing='ij')
    for c, s, ga, n in zip(lines['nu'], lines['sw'], lines['gamma_air'], lines['n_air']):
        gamma = ga * (pp / P_REFERENCE) * (T_REFERENCE / TT) ** n
        freq = 100 * c * k.speed_of_light
        strength = s * (TT / T_REFERENCE) ** n * np.exp(-(k.h_planck * freq / k.k_boltzmann) * (1 / TT - 1 / T_REFERENCE))
        centre = np.abs(nu - c).argmin()
        half = int(n_line_widths * gamma.max() / d_nu)
        i1, i2 = max(0, centre - half), min(nu.size - 1, centre + half) + 1
        x = nu[i1:i2] - c
        g3 = gamma[:, :, None]
        out[:, :, i1:i2] += strength[:, :, None] * (1 / np.pi) * g3 / (g3 ** 2 + x ** 2)
    return out


def synthetic_lines(kind, seed=0, n_lines=240):
    rng = np.random.default_rng(seed)
    if kind == 'co2':
        centres = np.concatenate((rng.normal(667, 60, n_lines // 2), rng.normal(2350, 60, n_lines // 4),
                                  rng.uniform(900, 1100, n_lines - n_lines // 2 - n_lines // 4)))
    elif kind == 'h2o':
        third = n_lines // 3
        centres = np.concatenate((rng.uniform(20, 600, third), rng.uniform(1300, 2000, third),
                                  rng.uniform(3000, 4200, n_lines - 2 * third)))
    else:
        raise ValueError(kind)
    return {'nu': np.sort(np.abs(centres)), 'sw': 10 ** rng.uniform(-3, 3, n_lines),
            'gamma_air': rng.uniform(0.05, 0.1, n_lines), 'n_air': rng.uniform(0.5, 0.8, n_lines)}


def synthetic_table(kind, d_nu, nu_max, n_p=48, seed=0):
    """Table dict {'p','T','nu','absorption_coef'} in the reference's format (hitran.py:315-356)
    with descending p (like CO2_no_sw.npy) and T = 250..350 step 20."""
    nu = np.arange(0.0, nu_max + d_nu, d_nu)
    p = np.logspace(np.log10(k.p_surface_earth), np.log10(k.p_toa_earth), n_p)
    T = np.arange(250.0, 351.0, 20.0)
    return {'p': p, 'T': T, 'nu': nu, 'absorption_coef': lorentz_table(nu, p, T, synthetic_lines(kind, seed))}


def co2_mass_ratio(ppmv):
    return ppmv * 1e-6 * 44.01 / 28.97


def columns(n_col, P, seed=0):
    """Synthetic lat-lon atmosphere: per column a tropospheric lapse up to a latitude-dependent
    tropopause, isothermal above, plus noise.  Returns T [n_col, L-1] on cell centres,
    T_g [n_col], T_tau [n_col, L] on interfaces, q [2, n_col, L] (CO2-like well mixed, H2O-like
    decaying with height and scaled per column), albedo [n_col]."""
    rng = np.random.default_rng(seed)
    P = np.asarray(P, dtype=float)
    p_c = 0.5 * (P[:-1] + P[1:])
    lat = np.arcsin(rng.uniform(-1, 1, n_col))
    T_ground = 300.0 - 50.0 * np.sin(lat) ** 2 + rng.normal(0, 2.0, n_col)
    T_trop = 195.0 + 25.0 * np.sin(lat) ** 2
    kappa = k.R_specific * 6.5e-3 / k.g

    def profile(p):
        t = T_ground[:, None] * (p[None, :] / P[-1]) ** kappa
        return np.maximum(t, T_trop[:, None])
    T = profile(p_c) + rng.normal(0, 0.3, (n_col, p_c.size))
    T_tau = profile(P)
    q = np.empty((2, n_col, P.size))
    q[0] = co2_mass_ratio(370.0)
    q[1] = (rng.uniform(0.2, 1.0, n_col) * 0.01)[:, None] * (P[None, :] / P[-1]) ** 3
    albedo = rng.uniform(0.2, 0.4, n_col)
    return T, T_ground, T_tau, q, albedo


def workload(d_nu=5.0, nz=100, n_nu_bands=40, T_g=288.0, planck=None):
    """Grids of the BASELINE configs: nz interfaces between p_surface_earth and p_toa_earth, the
    wavenumber grid / bands the reference would build for (T_g, T_sun) at spacing d_nu, and two
    synthetic tables covering it."""
    from .radiation import grids
    from .plan import bands_from_nu_bands
    P = grids.fixed_pressure_grid(nz, k.p_surface_earth, k.p_toa_earth)
    nu, nu_lw, nu_ov, nu_sw = grids.wavenumber_array(d_nu, k.T_sun, T_g, planck=planck)
    nu_bands = grids.wavenumber_bands(nu_lw, nu_ov, nu_sw, n_nu_bands, k.T_sun, T_g, planck=planck)
    bands = bands_from_nu_bands(nu, nu_bands)
    tables = [synthetic_table('co2', d_nu, 4500.0, seed=1), synthetic_table('h2o', d_nu, 4500.0, seed=2)]
    return {'P': P, 'nu': nu, 'bands': bands, 'tables': tables}
