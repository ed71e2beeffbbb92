"""``RealGas`` -- drop-in for the reference class (real_gas.py:187-840): same constructor
signature, public attributes and methods, with ``get_flux`` (and the net flux / heating rate
behind ``take_time_step``) evaluated by the CUDA path through a ``FluxPlan``.

What stays host-side NumPy here is the one-off set-up the reference does in ``__init__`` --
wavenumber grid, bands (``grids.py``), pressure grid, q(p) callables, Newton solves for T_g --
exactly the split of SURVEY.md section 2, row 3.  Plot methods are out of scope.
"""
import contextlib
import os

import numpy as np
from scipy import optimize
from scipy.interpolate import interp1d

from ..constants import T_sun, R_sun, AU, p_surface_earth, p_toa_earth, sigma
from ..plan import FluxPlan, bands_from_nu_bands
from . import grids
from .base import Atmosphere
from .real_gas import B_wavenumber, emission_integrals, get_absorption_coef, optical_depth, transmission
from .real_gas_data import hitran as _hitran
from .real_gas_data.specific_humidity import molecules, ppmv_from_humidity


class RealGas(Atmosphere):
    def __init__(self, nz, ny, molecule_names, T_g=None, q_funcs=None, q_funcs_args=None, n_nu_bands=40,
                 T_star=T_sun, R_star=R_sun, star_planet_dist=AU, albedo=0.3, temp_change=1, delta_temp_change=0.01,
                 T_func=None, p_surface=p_surface_earth, p_toa=p_toa_earth, precision='f64', device=0):
        """Arguments as the reference (real_gas.py:188-244).  Two opt-in extras that default to
        the reference behaviour: ``precision`` ('f64' | 'f32' arithmetic of the wavenumber-resolved
        stage) and the CUDA ``device``."""
        self.precision, self.device = precision, device
        self._plan, self._plan_no_emission, self._plan_key_seen, self._star_seen = None, None, None, None
        self.star = {'T': T_star, 'R': R_star, 'star_planet_dist': star_planet_dist}
        F_stellar_constant = sigma * T_star ** 4 * R_star ** 2 / star_planet_dist ** 2
        super().__init__(nz, ny, F_stellar_constant, albedo, p_surface, p_toa, temp_change, delta_temp_change)
        if T_g is not None:
            self.T_g = T_g
        elif T_func is not None:
            T_g = T_func(self.p_surface)
            self.T_g = T_g
        else:
            self.T_g = self.T0 + 20          # first guess with some greenhouse warming; refined below
        self.molecule_names = molecule_names
        default_args = None
        if q_funcs is None:
            q_funcs = {m: molecules[m]['q'] for m in molecule_names}
            default_args = {m: molecules[m]['q_args'] for m in molecule_names}
        self.q_funcs = q_funcs
        if q_funcs_args is None:
            self.q_funcs_args = default_args
        elif list(q_funcs_args.keys())[:len(q_funcs)] == list(q_funcs.keys()):
            self.q_funcs_args = q_funcs_args
        else:
            raise ValueError("Keys don't match in q_funcs and q_funcs_args")
        table_nu = _hitran.load_table(molecule_names[0])['nu']
        self.d_nu = table_nu[1] - table_nu[0]
        self.n_nu_bands = n_nu_bands
        self._build_spectral_grid()
        self.p_interface = np.sort(self.get_p_grid(), axis=0)
        self.p = np.zeros((self.nz - 1, self.ny))
        for i in range(self.nz - 1):
            self.p[i, :] = np.mean(self.p_interface[i:i + 2, :], 0)
        if T_func is None:
            self.T = np.ones_like(self.p) * self.T_g
            T_interface = np.ones_like(self.p_interface[:, 0]) * self.T_g
        else:
            self.T = T_func(self.p)
            T_interface = T_func(self.p_interface[:, 0])
        self.tau_interface = optical_depth(self.p_interface[:, 0], T_interface, self.nu, self.molecule_names,
                                           self.q_funcs, self.q_funcs_args)
        self._refresh_flux()
        if T_g is None:
            self.inital_Tg_guess()

    # ---- grids ------------------------------------------------------------------------------------
    def _build_spectral_grid(self):
        self.nu, self.nu_lw, nu_overlap, self.nu_sw = self.get_wavenumber_array()
        self.nu_bands = self.get_wavenumber_bands(nu_overlap)

    def get_wavenumber_array(self, fract_to_ignore=0.001, fract_to_ignore_overlap=0.001):
        """(nu, nu_lw, nu_overlap, nu_sw), real_gas.py:300-334."""
        return grids.wavenumber_array(self.d_nu, self.star['T'], self.T_g, fract_to_ignore, fract_to_ignore_overlap)

    def get_wavenumber_bands(self, nu_overlap):
        """nu_bands dict, real_gas.py:336-388."""
        return grids.wavenumber_bands(self.nu_lw, nu_overlap, self.nu_sw, self.n_nu_bands, self.star['T'], self.T_g)

    def get_p_grid(self, min_absorb_coef_use=10e-6, min_log_p_spacing_factor=5000, max_log_p_spacing_factor=50,
                   max_max_log_p_spacing=0.2):
        """Interface pressures [nz x ny] (ny = 1), surface first like the reference returns them
        (real_gas.py:404-503).  Integer nz: fixed spacing law.  nz='auto': log-pressure spacing
        shrinks where q x (mean absorption coefficient) -- i.e. d tau/dp -- is large."""
        if self.nz != 'auto':
            p_interface = grids.fixed_pressure_grid(self.nz, self.p_surface, self.p_toa)[::-1]
            return np.tile(p_interface, (self.ny, 1)).transpose()
        p_fine = np.logspace(np.log10(self.p_surface), np.log10(self.p_toa), int(1e6))
        weight = np.zeros_like(p_fine)          # sum over molecules of q(p) * relative absorption strength
        for name in self.molecule_names:
            table = _hitran.load_table(name)
            k_all = get_absorption_coef(table['p'], np.ones_like(table['p']) * self.T_g, table['nu'], table)
            strong = np.max(k_all, axis=0) > min_absorb_coef_use
            strength = np.mean(k_all[:, strong], axis=1)
            strength = strength / np.max(strength)
            if len(strength) > 1:
                inside = np.where(p_fine >= table['p'].min())[0]
                on_grid = np.ones_like(p_fine)
                on_grid[inside] = interp1d(table['p'], strength)(p_fine[inside])
                on_grid[p_fine < table['p'].min()] = on_grid[inside[-1]]
                strength = on_grid
            weight = weight + self.q_funcs[name](p_fine, *self.q_funcs_args[name]) * strength
        log_p = np.log10(p_fine)
        with np.errstate(divide='ignore'):
            log_w = np.log10(weight)
        log_w[weight == 0] = log_w[weight > 0].min()
        w_lo, w_hi = log_w.min(), log_w.max()
        finest = -w_hi / min_log_p_spacing_factor
        coarsest = np.clip(-w_lo / max_log_p_spacing_factor, finest, max_max_log_p_spacing)
        share_near_max = sum(weight > 0.9 * weight.max()) / len(weight)      # avoid huge nz
        finest = share_near_max * coarsest + (1 - share_near_max) * finest

        def spacing(lw):
            if w_lo == w_hi:
                return finest
            slope = (coarsest - finest) / (w_lo - w_hi)
            return slope * lw + (coarsest - slope * w_lo)

        levels = []
        current = log_p[0]
        while current > log_p[-1]:
            levels.append(current)
            current = levels[-1] - spacing(log_w[np.abs(log_p - current).argmin()])
        levels = np.array(levels)
        steps = np.cumsum(abs(np.ediff1d(levels)))
        steps = steps * ((log_p[0] - log_p[-1]) / steps[-1])                 # stretch to end exactly at p_toa
        levels = np.concatenate((levels[:1], levels[0] - steps))
        self.nz = len(levels)
        return np.tile(10 ** levels, (self.ny, 1)).transpose()

    # ---- the hot path ---------------------------------------------------------------------------------
    def _plan_key(self):
        """What a FluxPlan is built from: the grids (their bytes, not their identities -- callers may edit
        bands or rebuild arrays in place) and the star (re-read on every evaluation, like the reference
        reads self.star in get_flux, real_gas.py:640-642)."""
        import hashlib
        h = hashlib.blake2b(digest_size=16)
        for a in (self.nu, self.p_interface, self.nu_bands['centre'], self.nu_bands['delta'], self.nu_bands['sw'],
                  np.array([len(r) for r in self.nu_bands['range']]), np.array([r[0] for r in self.nu_bands['range']])):
            h.update(np.ascontiguousarray(a).tobytes())
        return h.hexdigest(), (float(self.star['T']), float(self.star['R']), float(self.star['star_planet_dist']))

    def _make_plan(self, bands):
        return FluxPlan(self.p_interface[:, 0], self.nu, bands, precision=self.precision, device=self.device,
                        T_star=self.star['T'], R_star=self.star['R'], star_planet_dist=self.star['star_planet_dist'])

    def _flux_plan(self, no_emission=False):
        grids_key, star = self._plan_key()
        if self._plan is None or grids_key != self._plan_key_seen:
            for old in (self._plan, self._plan_no_emission):
                if old is not None:
                    old.close()
            self._plan = self._make_plan(bands_from_nu_bands(self.nu, self.nu_bands))
            self._plan_no_emission, self._plan_key_seen, self._star_seen = None, grids_key, star
        if no_emission and self._plan_no_emission is None:
            # every band treated as short-wave: the sweep without emission integrals (OLR breakdown)
            bands = bands_from_nu_bands(self.nu, self.nu_bands)
            bands['sw'] = np.ones_like(bands['sw'])
            self._plan_no_emission = self._make_plan(bands)
        if star != self._star_seen:
            for plan in (self._plan, self._plan_no_emission):
                if plan is not None:
                    plan.set_star(*star)
            self._star_seen = star
        return self._plan_no_emission if no_emission else self._plan

    def _evaluate(self, outputs, no_emission=False):
        albedo = float(np.ravel(self.albedo)[0])
        return self._flux_plan(no_emission).flux(np.asarray(self.T, dtype=float)[:, 0], float(np.ravel(self.T_g)[0]),
                                                 tau=self.tau_interface, albedo=albedo, outputs=outputs)

    def get_flux(self, include_olr_breakdown=False):
        """(up_flux, down_flux), each [nz x n_nu_bands] in W m^-2 / cm^-1 (real_gas.py:629-669).
        With include_olr_breakdown the contributions to the top-of-atmosphere up flux are returned as well
        (olr_cont: 'surface' = transmitted surface emission, 'atmos' = the emission integral, :643-667)."""
        out = self._evaluate(('up', 'down'))
        up, down = out['up'][0], out['down'][0]
        if not include_olr_breakdown:
            return up, down
        # surface part = the same sweep with every band treated as short-wave (no emission integrals; the
        # plan for it is built once and kept); the atmosphere's part is the rest (real_gas.py:656-667)
        surf = self._evaluate(('up',), no_emission=True)
        surface = surf['up'][0][0].copy()
        atmos = np.where(self.nu_bands['sw'], 0.0, up[0] - surface)
        return up, down, {'surface': surface, 'atmos': atmos}

    def flux_integrals(self, j, T_interface):
        """(integral_up, integral_down), each [nz]: the emission integrals of long-wave band j
        (real_gas.py:564-627) for interface temperatures T_interface [nz] (or [nz x 1], as get_flux
        passes them).  The band's transmission matrix comes from the GPU (``transmission``), the
        Planck terms from ``B_wavenumber``; the pressure integrals are ``emission_integrals``."""
        P = np.asarray(self.p_interface, dtype=float)[:, 0]
        centre = self.nu_bands['centre'][j]
        trans = transmission(P, P, P, self.nu_bands['range'][j], self.nu_bands['delta'][j], self.nu, self.tau_interface)
        T_int = np.asarray(T_interface, dtype=float).reshape(-1)
        beta = lambda T: np.pi * B_wavenumber(centre, np.asarray(T, dtype=float).reshape(-1))
        return emission_integrals(trans, P, np.asarray(self.p, dtype=float)[:, 0], beta(self.T), beta(T_int),
                                  beta([float(np.ravel(self.T_g)[0])])[0])

    def __str__(self):
        return 'Real Gas'

    def _refresh_flux(self):
        out = self._evaluate(('up', 'down', 'net'))
        self.up_flux, self.down_flux = out['up'][0], out['down'][0]
        delta = self.nu_bands['delta']
        self.net_flux = np.sum(self.up_flux * delta, axis=1) - np.sum(self.down_flux * delta, axis=1)
        return self.net_flux

    def take_time_step(self, t, T_initial=None, changing_tau=False, convective_adjust=False, net_flux_thresh=1e-7,
                       net_flux_percentile=95, conv_thresh=1e-5, conv_t_multiplier=5):
        """One adaptive step (real_gas.py:671-718): fluxes for the current T, then update_temp."""
        if t == 0 and T_initial is not None:
            self.T = T_initial
        self.up_flux, self.down_flux = self.get_flux()
        delta = self.nu_bands['delta']
        net_flux = np.sum(self.up_flux * delta, axis=1) - np.sum(self.down_flux * delta, axis=1)
        return self.update_temp(t, net_flux.reshape(-1, 1), changing_tau, convective_adjust, net_flux_thresh,
                                net_flux_percentile, conv_thresh=conv_thresh, conv_t_multiplier=conv_t_multiplier)

    # ---- ground temperature -----------------------------------------------------------------------------
    def _isothermal_tau(self):
        T_interface = np.ones_like(self.p_interface[:, 0]) * self.T_g
        self.tau_interface = optical_depth(self.p_interface[:, 0], T_interface, self.nu, self.molecule_names,
                                           self.q_funcs, self.q_funcs_args)

    def inital_Tg_guess(self):
        """Newton on T_g (isothermal column) until the net flux summed over levels vanishes, then
        rebuild the spectral grid and optical depth for it (real_gas.py:505-528)."""
        def imbalance(x):
            self.T_g = x
            self.T = np.ones_like(self.p) * self.T_g
            return sum(self._refresh_flux())
        self.T_g = optimize.newton(imbalance, self.T_g)
        self.T = np.ones_like(self.p) * self.T_g
        self._build_spectral_grid()
        self._isothermal_tau()
        self._refresh_flux()

    def find_Tg(self, flux_thresh=0.1, tol=0.5, convective_adjust=False, print_eqb_info=True):
        """T_g for which the equilibrated column has ~zero net flux at the top (real_gas.py:530-562)."""
        print("Finding ground temperature to give top of atmosphere flux balance...")

        def toa_imbalance(x):
            print("Trying T_g = {:.1f} K".format(float(np.ravel(x)[0])))
            self.T_g = x
            self.evolve_to_equilibrium(flux_thresh=flux_thresh, save=False, convective_adjust=convective_adjust)
            return self.net_flux[0]
        if print_eqb_info:
            root = optimize.newton(toa_imbalance, self.T_g, tol=tol)
        else:
            with open(os.devnull, "w") as sink, contextlib.redirect_stdout(sink):
                root = optimize.newton(toa_imbalance, self.T_g, tol=tol)
        return root[0]

    # ---- histories ---------------------------------------------------------------------------------------
    def save_data(self, data_dict, t):
        """Append t, T (and LW/SW flux sums, q in ppmv when those keys exist), real_gas.py:720-746."""
        data_dict['t'].append(t)
        data_dict['T'].append(self.T.copy())
        if "flux" in data_dict:
            sw = self.nu_bands['sw']
            lw = ~sw
            delta = self.nu_bands['delta']
            for key, flux, sel in (('lw_up', self.up_flux, lw), ('lw_down', self.down_flux, lw),
                                   ('sw_up', self.up_flux, sw), ('sw_down', self.down_flux, sw)):
                data_dict['flux'][key].append(np.sum(flux[:, sel] * delta[sel], axis=1))
        if "q" in data_dict:
            for name in data_dict['q'].keys():
                q = self.q_funcs[name](self.p[:, 0], *self.q_funcs_args[name])
                data_dict['q'][name].append(ppmv_from_humidity(q, name))
        return data_dict

    def evolve_change_compos(self, T_g, q_args, data_dict=None, flux_thresh=1e-3, convective_adjust=False, t_end=2.0):
        """Equilibrate for a sequence of (T_g, composition) pairs, switching discontinuously
        (real_gas.py:748-785).  Bands are fixed from the first T_g; tau uses the first T_g's
        isothermal interface temperatures throughout, as in the reference."""
        self.T_g = T_g[0]
        self.T = np.ones_like(self.p) * self.T_g
        self._build_spectral_grid()
        T_interface = np.ones_like(self.p_interface[:, 0]) * self.T_g
        self.q_funcs_args = q_args[0]
        self.tau_interface = optical_depth(self.p_interface[:, 0], T_interface, self.nu, self.molecule_names,
                                           self.q_funcs, self.q_funcs_args)
        self._refresh_flux()
        for i in range(len(T_g)):
            print('Composition ' + str(i + 1) + '/' + str(len(T_g)))
            self.T_g = T_g[i]
            self.q_funcs_args = q_args[i]
            self.tau_interface = optical_depth(self.p_interface[:, 0], T_interface, self.nu, self.molecule_names,
                                               self.q_funcs, self.q_funcs_args)
            data_dict = self.evolve_to_equilibrium(data_dict, flux_thresh=flux_thresh,
                                                   convective_adjust=convective_adjust, t_end=t_end)
            self.time_step_info['DeltaT'] = self.time_step_info['MaxDeltaT']
        return data_dict
