"""``Atmosphere``: the adaptive time-stepper that drives the flux path towards equilibrium
(base.py:76-335 in the reference; SURVEY.md 8f row f1 -- the immediate caller of the hot path).
Host-side control logic: it consumes the net flux the CUDA path returns and decides the step.

Behaviour contract kept from the reference, branch for branch (the heuristics are stateful and a
different branch order diverges after a few steps):
* heating rate  dT/dt = g/c_p d(net)/dp on cell centres                         (base.py:166-168)
* dt such that the fastest-changing updated level moves by ``DeltaT`` K          (base.py:197-246)
* oscillation -> smaller DeltaT; steady sign -> larger; persistent offenders are frozen
* ``delta_net_flux`` = percentile of |net - previous net|                         (base.py:173)
"""
import inspect
from math import ceil, floor

import numpy as np

from ..constants import g, c_p_dry, sigma, p_surface_earth, p_toa_earth
from .convective_adjustment import convective_adjustment

_trapz = np.trapezoid if hasattr(np, 'trapezoid') else np.trapz
SECONDS_PER_DAY = 24 * 60 ** 2


def round_any(x, base, round_type='round'):
    """x rounded to a multiple of base ('round', 'ceil' or 'floor')."""
    op = {'round': round, 'ceil': ceil, 'floor': floor}.get(round_type)
    return None if op is None else base * op(x / base)


def t_years_days(t):
    """seconds -> (years, days)."""
    return divmod(t / SECONDS_PER_DAY, 365)


def latitudinal_solar_distribution(latitude, c=0.477):
    """Annual-mean insolation factor versus latitude (degrees), normalised so that
    0.5 * integral(factor * cos(lat)) dlat = 1; 1 for a single latitude."""
    if np.size(latitude) <= 1:
        return 1
    lat = np.radians(latitude)
    shape = 1 - 0.5 * c * (3 * np.sin(lat) ** 2 - 1)
    return shape / _trapz(0.5 * shape * np.cos(lat), lat)


def get_isothermal_temp(albedo, F_stellar=None, latitude=None, T_star=None, R_star=None, star_planet_dist=None):
    """Equilibrium temperature of an airless planet."""
    if F_stellar is None:
        F_stellar = sigma * T_star ** 4 * R_star ** 2 / star_planet_dist ** 2
    if latitude is not None:
        F_stellar = F_stellar * latitudinal_solar_distribution(latitude)
    return np.power(F_stellar / sigma * (1 - albedo) / 4, 1 / 4)


class Atmosphere:
    def __init__(self, nz, ny, F_stellar_constant, albedo=0.3, p_surface=p_surface_earth, p_toa=p_toa_earth,
                 temp_change=1, delta_temp_change=0.01):
        self.nz, self.ny = nz, ny
        self.p_surface, self.p_toa = p_surface, p_toa
        self.latitude = np.linspace(-90, 90, self.ny)
        if inspect.isfunction(albedo):
            self.albedo = albedo(self.latitude)
        elif np.size(albedo) == 1 and np.size(albedo) < self.ny:
            self.albedo = np.repeat(albedo, self.ny)
        else:
            self.albedo = albedo
        self.F_stellar_constant = F_stellar_constant
        self.solar_latitude_factor = latitudinal_solar_distribution(self.latitude)
        self.T0 = get_isothermal_temp(self.albedo, self.F_stellar_constant, self.latitude)
        self.time_step_info = {'DeltaT': temp_change, 'MaxTend': 0, 'MaxTendInd': -1,
                               'DeltaT_step': delta_temp_change, 'dt': 0, 'MaxDeltaT': temp_change,
                               'nSameMaxInd': 0, 'nSameMaxInd2': 0, 'RemoveInd': []}

    # -- one step ------------------------------------------------------------------------------
    def update_temp(self, t, net_flux, changing_tau=False, convective_adjust=False, net_flux_thresh=1e-7,
                    net_flux_percentile=95, conv_thresh=1e-5, conv_t_multiplier=5):
        """Advance self.T by one adaptive step given net_flux [nz x ny]; returns (t, delta_net_flux)."""
        info = self.time_step_info
        tendency = g / c_p_dry * (net_flux[1:, :] - net_flux[:-1, :]) / (self.p_interface[1:, :] - self.p_interface[:-1, :])
        if t > 0 and changing_tau is False:
            active = np.where(abs(net_flux[:-1].flatten()) > net_flux_thresh)[0]
            active = np.setdiff1d(active, info['RemoveInd'])
            delta_net_flux = np.percentile(abs(net_flux - self.net_flux), net_flux_percentile)
        else:
            active = np.arange((self.nz - 1) * self.ny)
            delta_net_flux = 1e6
        if 'convective_levels' not in info:
            info['convective_levels'] = np.array([])
        if len(active) > 0:
            self.update_time_step(tendency, active)
            if convective_adjust and info['MaxTendInd'] in info['convective_levels']:
                info['dt'] = info['dt'] * conv_t_multiplier
        self.net_flux = net_flux
        where = np.unravel_index(active, self.T.shape)
        self.T[where] = self.T[where] + info['dt'] * tendency[where]
        if convective_adjust:
            adjusted = convective_adjustment(self.p[:, 0], self.T.copy())
            moved = abs(adjusted.flatten()[active] - self.T.flatten()[active]) > conv_thresh
            info['convective_levels'] = active[np.where(moved)[0]]
            self.T = adjusted
        return t + info['dt'], delta_net_flux

    def update_time_step(self, T_tendency, levels_to_update):
        """dt = DeltaT / |largest tendency among updated levels|, with DeltaT adapted to
        oscillation / stagnation of the level that sets it (state in self.time_step_info)."""
        info = self.time_step_info
        flat = T_tendency.flatten()
        ind = levels_to_update[np.argmax(abs(flat[levels_to_update]))]
        tend = flat[ind]
        same_level = ind == info['MaxTendInd']
        same_sign = np.sign(tend) == np.sign(info['MaxTend'])
        if same_level and not same_sign:
            info['DeltaT'] = info['DeltaT'] - info['DeltaT_step']
            info['nSameMaxInd'] += 1
            if info['DeltaT'] < info['DeltaT_step']:
                info['DeltaT'] = info['DeltaT_step']
        elif same_level and same_sign and info['DeltaT'] < info['MaxDeltaT']:
            info['DeltaT'] = info['DeltaT'] + info['DeltaT_step']
            info['nSameMaxInd2'] += 1
            if info['DeltaT'] > info['MaxDeltaT']:
                info['DeltaT'] = info['MaxDeltaT']
        else:
            info['nSameMaxInd'] = 0
            info['nSameMaxInd2'] = 0
        n1, n2 = info['nSameMaxInd'], info['nSameMaxInd2']
        if (n1 > 1 and n2 > 10) or (n1 > 20 and n2 == 0) or (len(info['RemoveInd']) > 3 and n1 + n2 > 0):
            info['RemoveInd'].append(ind)
            info['nSameMaxInd'] = 0
            info['nSameMaxInd2'] = 0
        info['MaxTendInd'] = ind
        info['MaxTend'] = tend
        with np.errstate(divide='ignore'):
            info['dt'] = float(info['DeltaT'] / abs(tend).max())
        if np.isinf(info['dt']):
            info['dt'] = SECONDS_PER_DAY

    def check_equilibrium(self, delta_net_flux, flux_thresh=1e-3):
        return bool(max(abs(self.net_flux.flatten())) < flux_thresh or delta_net_flux < flux_thresh)

    # -- driver ----------------------------------------------------------------------------------
    def evolve_to_equilibrium(self, data_dict=None, flux_thresh=1e-3, T_initial=None, convective_adjust=False,
                              save=True, t_end=4.0, conv_thresh=1e-5, conv_t_multiplier=5):
        """Step until equilibrium (or t_end years); returns data_dict with 't' and 'T' histories."""
        if data_dict is None:
            if T_initial is None:
                T_initial = self.T.copy()
            data_dict = {'t': [0], 'T': [T_initial.copy()]}
        t = data_dict['t'][-1]
        years0 = t_years_days(t)[0] + t_years_days(t)[1] / 365
        step = 0
        done = False
        while not done:
            t, delta_net_flux = self.take_time_step(t, T_initial, changing_tau=False,
                                                    convective_adjust=convective_adjust, conv_thresh=conv_thresh,
                                                    conv_t_multiplier=conv_t_multiplier)
            if save:
                data_dict = self.save_data(data_dict, t)
            if step == 1:
                flux_thresh = min(flux_thresh, 0.99 * delta_net_flux)
                print("Trying to reach equilibrium (flux_thresh = {:.4f})...".format(flux_thresh))
            done = self.check_equilibrium(delta_net_flux, flux_thresh)
            if min(self.T.flatten()) < 0:
                raise ValueError('Temperature is below zero')
            years, days = t_years_days(t)
            if years + days / 365 - years0 > t_end:
                done = True
            print("{:.0f} Years, {:.0f} Days: delta_net_flux = {:.4f}".format(years, days, delta_net_flux), end="\r")
            step += 1
        print("{:.0f} Years, {:.0f} Days: delta_net_flux = {:.4f}".format(years, days, delta_net_flux))
        print("Done!")
        self.time_step_info.update({'RemoveInd': [], 'nSameMaxInd': 0, 'nSameMaxInd2': 0, 'MaxTendInd': -1})
        return data_dict
