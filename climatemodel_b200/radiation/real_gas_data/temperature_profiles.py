"""Host-side input generators: temperature profiles T(p) in K
(real_gas_data/temperature_profiles.py:16-59).  Piecewise-linear in altitude through a few
(height, temperature) knots; above the last knot the last temperature holds.  ``earth_temp``
uses the knots of Solomon et al., figure 3a, as the reference does."""
import numpy as np
from scipy.interpolate import interp1d

from .specific_humidity import p_altitude_convert


def _through_knots(p, h_km, T_knots):
    h_knots = np.asarray(h_km, dtype=float) * 1000
    T_knots = np.asarray(T_knots, dtype=float)
    h = p_altitude_convert(p=p)
    T = np.zeros_like(p, dtype=float)
    inside = h <= h_knots[-1]
    T[inside] = interp1d(h_knots, T_knots)(h[inside])
    T[~inside] = T_knots[-1]
    return T


def earth_temp(p):
    return _through_knots(p, [0, 12, 19, 21, 30, 40, 46, 50, 70, 79, 81, 88, 99, 140],
                          [288, 210, 205, 215, 226, 250, 260, 260, 210, 199, 199, 202, 195, 610])


def fixed_tropopause_temp(p, h_tropopause=19, T_tropopause=205, T_ground=288):
    """Linear troposphere up to h_tropopause (km), isothermal at T_tropopause above."""
    return _through_knots(p, [0, h_tropopause, 140], [T_ground, T_tropopause, T_tropopause])


def two_lapse_temp(p, h_tropopause=10, h_top=20, T_ground=288, lapse_trop=9, lapse_strat=0):
    """Two constant lapse rates (K/km): lapse_trop up to h_tropopause, lapse_strat up to h_top."""
    T_tropopause = T_ground - h_tropopause * lapse_trop
    T_top = T_tropopause - (h_top - h_tropopause) * lapse_strat
    return _through_knots(p, [0, h_tropopause, h_top], [T_ground, T_tropopause, T_top])
