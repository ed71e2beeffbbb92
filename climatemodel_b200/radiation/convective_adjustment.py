"""Dry convective adjustment of a temperature profile, conserving enthalpy (integral of T dp)
-- the step right after the flux path in a radiative-convective time step
(convective_adjustment.py:7-135 in the reference).  Host-side NumPy:
data-dependent iteration over groups of unstable levels, O(nz) per column.

Same call signature and in-place semantics as the reference: ``T`` is modified (callers pass a
copy, base.py:188) and returned.
"""
import numpy as np

from ..constants import g, c_p_dry, p_surface_earth, R_specific


def get_theta(T, p, p_reference, alpha):
    return T / (p / p_reference) ** alpha


def get_T(theta, p, p_reference, alpha):
    return theta * (p / p_reference) ** alpha


def get_enthalpy(T, p):
    """Proportional to the column enthalpy: -trapz(T, p) for descending p."""
    return -(np.trapezoid if hasattr(np, 'trapezoid') else np.trapz)(T, p)


def nearest_value_in_array(array, value):
    return array[np.abs(array - value).argmin()]


def _runs(levels):
    """Split a sorted index array into lists of consecutive indices."""
    if len(levels) == 0:
        return []
    cuts = np.where(np.diff(levels) != 1)[0] + 1
    return [list(map(int, r)) for r in np.split(np.asarray(levels), cuts)]


def _unstable(theta, small=1e-10):
    d = np.ediff1d(theta)
    d = np.concatenate((d, [d[-1]]))
    return np.where(d < -small)[0]          # theta must not decrease with height (p descending)


def convective_adjustment_single(p, T, lapserate=g / c_p_dry, delta_T_thresh='auto'):
    """One column, p[0] = surface > p[-1] = top.  Wherever potential temperature decreases with
    height, two candidate profiles are built -- the unstable region (grown downwards / upwards
    until stable) set to the theta just above it, or to the theta at its base -- and blended so
    that the column enthalpy is unchanged.  A group whose adjustment would change T by more than
    ``delta_T_thresh`` (default median(T)/4) is left alone.  Repeats until no unstable level
    remains."""
    if isinstance(delta_T_thresh, str) and delta_T_thresh == 'auto':
        delta_T_thresh = np.median(T) / 4.0
    p_ref = p_surface_earth
    alpha = R_specific * lapserate / g
    theta = get_theta(T, p, p_ref, alpha)
    unstable = _unstable(theta)
    ignored = []
    while len(unstable) > 0:
        enthalpy = get_enthalpy(T, p)
        for group in _runs(unstable):
            candidates = {}
            # (a) pull the region down to the theta of the level just above it
            top = group[-1] + 1
            theta_low = theta[top]
            colder = np.where(theta < theta_low)[0]
            colder = colder[colder < top]
            first = 0 if len(colder) == 0 else nearest_value_in_array(colder, top) + 1
            candidates['lower'] = (theta_low, np.arange(first, top + 1))
            # (b) lift the region up to the theta at its base
            base = group[0]
            theta_up = theta[base]
            warmer = np.where(theta > theta_up)[0]
            warmer = warmer[warmer > base]
            last = len(p) - 1 if len(warmer) == 0 else nearest_value_in_array(warmer, base)
            candidates['upper'] = (theta_up, np.arange(base, last + 1))
            profile, h = {}, {}
            for key, (value, levels) in candidates.items():
                th = theta.copy()
                th[levels] = value
                profile[key] = get_T(th, p, p_ref, alpha)
                h[key] = get_enthalpy(profile[key], p)
            beta = (enthalpy - h['lower']) / (h['upper'] - h['lower'])
            new_T = beta * profile['upper'] + (1 - beta) * profile['lower']
            if abs(new_T - T).max() < delta_T_thresh:
                T = new_T
            else:
                ignored = ignored + group
            theta = get_theta(T, p, p_ref, alpha)
        unstable = np.setdiff1d(_unstable(theta), ignored)
    return T


def convective_adjustment(p, T, lapserate=g / c_p_dry, delta_T_thresh='auto'):
    """All columns of T [nz x ny]; p [nz] in either order."""
    if len(np.shape(p)) > 1:
        raise ValueError('Pressure has wrong dimension, must be only in the z dimension')
    flip = p[1] > p[0]
    if flip:
        p, T = p[::-1], T[::-1, :]
    for j in range(np.shape(T)[1]):
        T[:, j] = convective_adjustment_single(p, T[:, j], lapserate)
    return T[::-1, :] if flip else T
