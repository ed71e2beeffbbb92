"""Host-side input generators: specific humidity q(p) = rho_molecule / rho_air per molecule, the
pressure <-> altitude map and the ``molecules`` registry the reference's ``RealGas`` uses for its
default ``q_funcs`` (real_gas_data/specific_humidity.py:21-266).  These only *feed* the hot path
(they are called through ``q_funcs[mol](p, *args)``, real_gas.py:118) and stay NumPy.

The profiles are piecewise-linear curves in altitude read off Solomon et al. (whole-atmosphere
climate change) figure 1; the knot data below are that paper's numbers as used by the reference.
Function names, arguments, defaults and units match the reference so that ``q_funcs`` /
``q_funcs_args`` written for it keep working.
"""
import numpy as np
from scipy.interpolate import interp1d

from ...constants import p_surface_earth

M_air = 28.97                       # g / mol
temp_kelvin_to_celsius = 273.15

# log10(p) is linear in altitude in two pieces: surface -> 90 km (0.1 Pa) -> 130 km (1e-3 Pa)
_H_BREAK, _LOGP_BREAK, _H_TOP, _LOGP_TOP = 90000.0, -1.0, 130000.0, -3.0


def _slopes():
    logp0 = np.log10(p_surface_earth)
    return logp0, (_LOGP_BREAK - logp0) / _H_BREAK, (_LOGP_TOP - _LOGP_BREAK) / (_H_TOP - _H_BREAK)


def p_altitude_convert(altitude=None, p=None):
    """altitude (m) -> pressure (Pa) when ``altitude`` is given, pressure -> altitude when ``p`` is."""
    logp0, lower, upper = _slopes()
    if p is None:
        altitude = np.asarray(altitude)
        log_p = np.where(altitude <= _H_BREAK, logp0 + lower * altitude,
                         _LOGP_BREAK + upper * (altitude - _H_BREAK)).astype(float)
        return 10 ** log_p
    if altitude is None:
        log_p = np.log10(p)
        return np.where(log_p >= _LOGP_BREAK, (log_p - logp0) / lower,
                        (log_p - _LOGP_BREAK) / upper + _H_BREAK).astype(float)
    raise ValueError('give either altitude or p')


def humidity_from_ppmv(conc_ppmv, molecule_name):
    """ppmv -> kg/kg."""
    return conc_ppmv / 10 ** 6 * molecules[molecule_name]['M'] / M_air


def ppmv_from_humidity(humidity, molecule_name):
    """kg/kg -> ppmv."""
    return humidity * 10 ** 6 * M_air / molecules[molecule_name]['M']


def _curve(p, h_km, values, log_values=False):
    """Piecewise-linear profile in altitude through (h_km, values); zero at and above the last
    knot; negative values clipped.  ``log_values`` interpolates log10(values)."""
    h_knots = np.asarray(h_km, dtype=float) * 1000
    y = np.log10(values) if log_values else np.asarray(values, dtype=float)
    h = p_altitude_convert(p=p)
    out = np.zeros_like(p, dtype=float)
    below = h < h_knots.max()
    inside = interp1d(h_knots, y)(h[below])       # same interpolator as the reference (raises below 0 m)
    out[below] = 10 ** inside if log_values else inside
    out[out < 0] = 0
    return out


def co2(p, q_surface=370, h_change=80000):
    """Well mixed (q_surface ppmv) below h_change, then falling linearly to 60 ppmv at 120 km."""
    if q_surface == 0:
        return np.zeros_like(p)
    h = p_altitude_convert(p=p)
    slope = (q_surface - 60) / (h_change - 120000)
    q = np.ones_like(p) * q_surface
    above = h > h_change
    q[above] = (q_surface - slope * h_change) + slope * h[above]
    q[q < 0] = 0
    return humidity_from_ppmv(q, 'CO2')


def ch4(p, scale_factor=1):
    if scale_factor == 0:
        return np.zeros_like(p)
    values = np.array([1.75, 1.75, 1.68, 1.32, 1.19, 0.4, 0.19, 0.04, 0]) * scale_factor
    values[1] = values[0]
    values[values > values[0]] = values[0]
    return humidity_from_ppmv(_curve(p, [0, 10, 17, 22, 28, 50, 68, 80, 90], values), 'CH4')


def h2o(p, scale_factor=1):
    if scale_factor == 0:
        return np.zeros_like(p)
    values = np.array([20000, 2500, 250, 12, 4, 4.3, 4.9, 5.1, 5.7, 5.9, 6, 6.1, 6, 5.8, 5, 4, 2.5, 1])
    return humidity_from_ppmv(_curve(p, np.arange(0, 90, 5), values, log_values=True), 'H2O') * scale_factor


def o3(p, scale_factor=1):
    if scale_factor == 0:
        return np.zeros_like(p)
    h_km = np.sort(np.concatenate((np.arange(0, 125, 5), np.array([32, 78, 92]))))
    values = np.array([0.05, 0.07, 0.09, 0.25, 1.8, 5.25, 7.8, 7.9, 7.85, 6, 3.8, 2.4, 1.6, 1, 0.75, 0.3, 0.15,
                       0.1, 0.15, 0.8, 1.75, 1.8, 1.7, 1, 0.3, 0.07, 0.05, 0]) * scale_factor
    return humidity_from_ppmv(_curve(p, h_km, values), 'O3')


def constant_q(p, q_surface, molecule_name):
    return humidity_from_ppmv(np.ones_like(p) * q_surface, molecule_name.upper())


def gradient_q(p, q_sfc, q_upper, h_upper, molecule_name='CO2'):
    """q_sfc at the surface changing linearly in height to q_upper at h_upper, constant above."""
    h = p_altitude_convert(p=p)
    q = np.where(h >= h_upper, q_upper, q_sfc + h * (q_upper - q_sfc) / h_upper).astype(float)
    return humidity_from_ppmv(q, molecule_name.upper())


def saturation_vapor_pressure(temp):
    """Bolton (1980) eq. 10; temp in K, result in Pa."""
    celsius = temp - temp_kelvin_to_celsius
    return 611.2 * np.exp(17.67 * celsius / (celsius + 243.5))


def constant_rh(p, temp_func, rh=0.7, h_upper=None, molecule_name='H2O'):
    """Constant relative humidity for the temperature profile temp_func(p); zero above h_upper."""
    conc_ppmv = rh * saturation_vapor_pressure(temp_func(p)) / p * 10 ** 6
    if h_upper is not None:
        conc_ppmv[p_altitude_convert(p=p) >= h_upper] = 0
    return humidity_from_ppmv(conc_ppmv, molecule_name.upper())


molecules = {'H2O': {'hitran_id': 1, 'M': 18, 'q': h2o, 'q_args': (1,)},
             'CO2': {'hitran_id': 2, 'M': 44, 'q': co2, 'q_args': (370, 80000)},
             'O3': {'hitran_id': 3, 'M': 48, 'q': o3, 'q_args': (1,)},
             'CH4': {'hitran_id': 6, 'M': 16, 'q': ch4, 'q_args': (1,)},
             'CFC12': {'hitrad_id': 10, 'M': 120.91, 'q': o3, 'q_args': (1,)}}
