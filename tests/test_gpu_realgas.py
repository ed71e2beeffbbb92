"""GPU: the drop-in ``RealGas`` class against golden runs of the unmodified reference --
construction (grids, bands, optical depth), fluxes, and several adaptive time steps with and
without convective adjustment (tests/golden/make_golden.py wrote step_T / step_t / step_dnf)."""
import contextlib
import io

import numpy as np
import pytest

from util import TABLES, load_case, relmax

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def mods():
    import climatemodel_b200.radiation.real_gas as rg
    import climatemodel_b200.radiation.real_gas_data.hitran as hitran
    import climatemodel_b200.radiation.real_gas_data.specific_humidity as hum
    import climatemodel_b200.radiation.real_gas_data.temperature_profiles as tp
    hitran.LookupTableFolder = TABLES + '/'
    return rg, hum, tp


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def build(mods, name):
    rg, hum, tp = mods
    co2 = dict(q_funcs={'CO2_no_sw': hum.co2}, q_funcs_args={'CO2_no_sw': (370, 80000)})
    if name.startswith('co2_nz50_earth'):
        return rg.RealGas(nz=50, ny=1, molecule_names=['CO2_no_sw'], T_g=288, T_func=tp.earth_temp, **co2)
    if name == 'co2_nz100_iso':
        return rg.RealGas(nz=100, ny=1, molecule_names=['CO2_no_sw'], T_g=288, **co2)
    if name == 'gray_nz60_200b':
        return rg.RealGas(nz=60, ny=1, molecule_names=['gray'], T_g=270, p_toa=10, n_nu_bands=200,
                          q_funcs={'gray': hum.co2}, q_funcs_args={'gray': (370, 80000)})
    if name == 'two_mol_nz41':
        def T_two(p):
            return tp.two_lapse_temp(p, h_tropopause=11, h_top=20, T_ground=295, lapse_trop=6.5)
        return rg.RealGas(nz=41, ny=1, molecule_names=['CO2_no_sw', 'H2O_synth'], T_g=295, T_func=T_two,
                          q_funcs={'CO2_no_sw': hum.co2, 'H2O_synth': hum.h2o},
                          q_funcs_args={'CO2_no_sw': (370, 80000), 'H2O_synth': (1,)})
    raise KeyError(name)


@pytest.mark.parametrize('name', ['co2_nz50_earth', 'co2_nz100_iso', 'gray_nz60_200b', 'two_mol_nz41'])
def test_construction_and_fluxes(mods, name):
    c = load_case(name)
    gas = build(mods, name)
    assert gas.nz == c['P'].size and gas.p_interface.shape == (c['P'].size, 1)
    assert np.array_equal(gas.p_interface[:, 0], c['P'])
    assert np.array_equal(gas.p[:, 0], c['p'])
    assert np.array_equal(gas.nu, c['nu'])
    assert np.array_equal(gas.nu_bands['centre'], c['band_centre'])
    assert np.array_equal(gas.nu_bands['delta'], c['band_delta'])
    assert np.array_equal(gas.nu_bands['sw'], c['band_sw'])
    assert np.array_equal(gas.T[:, 0], c['T'])
    assert gas.tau_interface.shape == (c['P'].size, c['nu'].size)
    assert np.array_equal(gas.tau_interface[:, c['tau_sub_idx']], c['tau_sub'])
    assert gas.up_flux.shape == c['up'].shape
    assert relmax(gas.up_flux, c['up']) < 1e-10
    assert relmax(gas.down_flux, c['down']) < 1e-10
    assert relmax(gas.net_flux, c['net']) < 1e-10
    up, down = gas.get_flux()
    assert relmax(up, c['up']) < 1e-10 and relmax(down, c['down']) < 1e-10


@pytest.mark.parametrize('name', ['co2_nz50_earth', 'co2_nz50_earth_conv', 'co2_nz100_iso', 'two_mol_nz41'])
def test_time_steps_follow_the_reference(mods, name):
    c = load_case(name)
    gas = build(mods, name)
    conv = bool(c['step_conv'])
    t = 0.0
    for i in range(c['step_T'].shape[0]):
        t, dnf = quiet(gas.take_time_step, t, convective_adjust=conv)
        assert abs(t - c['step_t'][i]) <= 1e-7 * c['step_t'][i]
        assert np.abs(gas.T[:, 0] - c['step_T'][i]).max() < 1e-6
        if i > 0:
            assert abs(dnf - c['step_dnf'][i]) <= 1e-6 * max(abs(c['step_dnf'][i]), 1e-3)
    assert relmax(gas.net_flux.reshape(-1), c['step_net_last']) < 1e-8


def test_olr_breakdown_adds_up(mods):
    gas = build(mods, 'co2_nz50_earth')
    up, down, cont = gas.get_flux(include_olr_breakdown=True)
    assert np.allclose(cont['surface'] + cont['atmos'], up[0], rtol=1e-12, atol=1e-300)
    assert np.all(cont['atmos'][gas.nu_bands['sw']] == 0)
    assert np.all(cont['surface'] >= 0)


def test_assigning_tau_and_q_args_like_the_notebooks(mods):
    """centa_presentation/base.py:38-50 recomputes tau after changing q_funcs_args and assigns it."""
    rg, hum, tp = mods
    gas = build(mods, 'co2_nz50_earth')           # temperature falls with height: more CO2, less OLR
    before = gas.get_flux()[0][0].copy()
    gas.q_funcs_args = {'CO2_no_sw': (740, 80000)}
    gas.tau_interface = rg.optical_depth(gas.p_interface[:, 0], np.ones(gas.nz) * gas.T_g, gas.nu,
                                         gas.molecule_names, gas.q_funcs, gas.q_funcs_args)
    after = gas.get_flux()[0][0]
    delta = gas.nu_bands['delta']
    assert np.sum(after * delta) < np.sum(before * delta) - 1.0    # doubling CO2 lowers the OLR (W/m^2)


def test_auto_Tg_guess_balances_the_column(mods):
    rg, hum, tp = mods
    gas = quiet(rg.RealGas, nz=30, ny=1, molecule_names=['CO2_no_sw'], q_funcs={'CO2_no_sw': hum.co2},
                q_funcs_args={'CO2_no_sw': (370, 80000)})
    # the unmodified reference, same call (Newton on the isothermal column, then the spectral grid
    # and optical depth are rebuilt for the new T_g, so the sum is no longer exactly zero)
    assert abs(float(np.ravel(gas.T_g)[0]) - 259.8448678790483) < 1e-8
    assert abs(np.sum(gas.net_flux) - (-7.506481761631903)) < 1e-8


@pytest.mark.parametrize('name', ['co2_nz50_earth', 'two_mol_nz41'])
def test_flux_integrals_method(mods, name):
    """``RealGas.flux_integrals(j, T_interface)`` (real_gas.py:564-627) against the unmodified
    reference's method (tests/golden/flux_integrals.npz): three long-wave bands, the spline
    interface temperatures and a perturbed set, passed as [nz] and as [nz x 1].

    The integrals are sums of beta_k * (Trans(i,k+1) - Trans(i,k)): an error of a few ulp of
    Trans (<= 1 in the upper triangle; the GPU sums the nu-trapezoid in another order and has its
    own exp) moves them by up to ~ eps * max(beta) per term, whatever their size.  Bands 'a' and
    'b' absorb strongly, so the 1e-10 gate alone decides; band 'c' is nearly transparent
    (1 - Trans ~ 1e-13 / 3e-5): its reference integrals are themselves differences of numbers
    equal to 1e-13, and only the rounding floor 16 eps L max(beta) can be asked of it.  The
    pieces are pinned separately too: GPU transmission against the stored matrix, and
    ``emission_integrals`` on the stored matrix against the reference outputs."""
    import os
    from util import GOLDEN
    from climatemodel_b200.radiation.real_gas import B_wavenumber, emission_integrals, transmission
    g = np.load(os.path.join(GOLDEN, 'flux_integrals.npz'))
    gas = build(mods, name)
    assert str(gas) == 'Real Gas'
    P, p = gas.p_interface[:, 0], gas.p[:, 0]
    eps = np.finfo(float).eps
    for tag in ('a', 'b', 'c'):
        key = '%s_%s_' % (name, tag)
        j = int(g[key + 'band'])
        centre = gas.nu_bands['centre'][j]
        beta = lambda T: np.pi * B_wavenumber(centre, np.asarray(T, dtype=float).reshape(-1))
        ref_trans = g[key + 'trans_upper']
        trans = transmission(P, P, P, gas.nu_bands['range'][j], gas.nu_bands['delta'][j], gas.nu, gas.tau_interface)
        assert np.abs(np.triu(trans) - ref_trans).max() <= 1e-13 * np.abs(ref_trans).max()
        for ttag in ('spline', 'pert'):
            T_int = g[key + ttag + '_T_interface']
            ref_up, ref_dn = g[key + ttag + '_up'], g[key + ttag + '_down']
            beta_max = float(max(beta(gas.T).max(), beta(T_int).max()))
            floor = 16 * eps * gas.nz * beta_max * np.abs(ref_trans).max()
            # the host half on the reference's own matrix
            with np.errstate(all='ignore'):
                up0, dn0 = emission_integrals(np.where(np.tri(gas.nz, k=-1, dtype=bool), np.inf, ref_trans), P, p,
                                              beta(gas.T), beta(T_int), beta([float(np.ravel(gas.T_g)[0])])[0])
            assert np.abs(up0 - ref_up).max() <= 1e-10 * np.abs(ref_up).max() + floor
            assert np.abs(dn0 - ref_dn).max() <= 1e-10 * np.abs(ref_dn).max() + floor
            # the method end to end
            up, dn = gas.flux_integrals(j, T_int if ttag == 'pert' else T_int.reshape(-1, 1))
            assert up.shape == (gas.nz,) and dn.shape == (gas.nz,)
            assert np.abs(up - ref_up).max() <= 1e-10 * np.abs(ref_up).max() + floor
            assert np.abs(dn - ref_dn).max() <= 1e-10 * np.abs(ref_dn).max() + floor
            if tag in 'ab':      # well conditioned: the rounding floor is far below the gate
                assert floor < 1e-12 * np.abs(ref_up).max()
                assert relmax(up, ref_up) < 1e-10 and relmax(dn, ref_dn) < 1e-10
