"""CPU: the host-side callers of the flux path -- adaptive time-stepper (base.py:166-246) and dry
convective adjustment (convective_adjustment.py:7-135) -- replayed against vectors written by the
unmodified reference (tests/golden/make_golden_host.py)."""
import os

import numpy as np
import pytest

from util import GOLDEN

from climatemodel_b200.radiation import base
from climatemodel_b200.radiation.convective_adjustment import convective_adjustment, get_enthalpy


@pytest.fixture(scope='module')
def stepper():
    return dict(np.load(os.path.join(GOLDEN, 'host_stepper.npz')))


def test_convective_adjustment_matches_reference():
    c = np.load(os.path.join(GOLDEN, 'host_convective.npz'))
    changed = 0
    for pp, T, want in zip(c['p'], c['T'], c['T_adjusted']):
        got = convective_adjustment(pp, T.reshape(-1, 1).copy())[:, 0]
        assert np.array_equal(got, want)
        changed += int(np.abs(want - T).max() > 1e-3)
        # enthalpy is conserved by construction of the blend
        hp, hT, hw = (pp, T, want) if pp[0] > pp[-1] else (pp[::-1], T[::-1], want[::-1])
        assert abs(get_enthalpy(hw, hp) - get_enthalpy(hT, hp)) < 1e-9 * abs(get_enthalpy(hT, hp))
    assert changed >= 6


def test_convective_adjustment_rejects_2d_pressure():
    with pytest.raises(ValueError):
        convective_adjustment(np.ones((4, 2)), np.ones((4, 2)))


@pytest.mark.parametrize('tag', ['plain', 'conv'])
def test_time_stepper_replays_reference(stepper, tag):
    s = stepper
    nz = s['P'].size
    atm = base.Atmosphere(nz, 1, 1367.0 / 4)
    assert float(np.ravel(atm.T0)[0]) == s['T0']
    P = s['P'].reshape(-1, 1)
    atm.p_interface = P
    atm.p = (P[:-1] + P[1:]) / 2
    atm.T = s['T_init'].reshape(-1, 1).copy()
    atm.net_flux = np.zeros((nz, 1))
    t = 0.0
    for i in range(s[tag + '_t'].size):
        t, dnf = atm.update_temp(t, s[tag + '_net'][i].reshape(-1, 1), convective_adjust=(tag == 'conv'))
        assert t == s[tag + '_t'][i], i
        assert atm.time_step_info['dt'] == s[tag + '_dt'][i], i
        assert atm.time_step_info['DeltaT'] == s[tag + '_DeltaT'][i], i
        assert dnf == s[tag + '_dnf'][i], i
        assert np.array_equal(atm.T[:, 0], s[tag + '_T'][i]), i
    assert list(atm.time_step_info['RemoveInd']) == list(s[tag + '_removed'])
    assert len(s[tag + '_removed']) > 0                     # the 'freeze a level' branch was visited
    assert s[tag + '_DeltaT'].min() < 1 and s[tag + '_DeltaT'].max() == 1


def test_convection_changes_the_trajectory(stepper):
    assert np.abs(stepper['plain_T'] - stepper['conv_T']).max() > 0.1


def test_solar_distribution_and_isothermal_temperature(stepper):
    s = stepper
    assert np.array_equal(base.latitudinal_solar_distribution(s['lat']), s['lat_factor'])
    assert np.array_equal(base.get_isothermal_temp(0.3, 1367.0 / 4, s['lat']), s['iso_T'])
    assert base.latitudinal_solar_distribution(np.array([0.0])) == 1


def test_small_helpers():
    assert base.round_any(37, 5) == 35 and base.round_any(37, 5, 'ceil') == 40 and base.round_any(37, 5, 'floor') == 35
    years, days = base.t_years_days(400 * 24 * 3600)
    assert years == 1 and days == 35
    atm = base.Atmosphere(5, 3, 300.0, albedo=lambda lat: 0.3 + 0 * lat)
    assert np.shape(atm.albedo) == (3,)
    atm.net_flux = np.full((5, 3), 1e-6)
    assert atm.check_equilibrium(1.0) and not base.Atmosphere.check_equilibrium(
        type('A', (), {'net_flux': np.ones((2, 1))})(), 1.0)
