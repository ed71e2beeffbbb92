"""CPU: pin the oracle (oracle/realgas_oracle.py) against golden vectors produced by the
unmodified reference (tests/golden/make_golden.py)."""
import numpy as np
import pytest

import oracle.realgas_oracle as orc
from util import CASES, load_case, load_table, relmax, GOLDEN
import os


def test_planck_and_gather_vectors():
    f = np.load(os.path.join(GOLDEN, 'functions.npz'))
    B = orc.planck_wavenumber(f['planck_nu'], f['planck_T'])
    assert np.array_equal(B, f['planck_B'])
    tab = load_table('CO2_no_sw')
    k = orc.gather_absorption_coef(f['gather_p'], f['gather_T'], f['gather_nu'], tab)
    assert np.array_equal(k, f['gather_k'])
    k2 = orc.absorption_coef_on_grid(f['gather_p'], f['gather_T'], f['load_nu'], tab)
    assert np.array_equal(k2, f['load_k'])
    assert np.all(k2[:, f['load_nu'] > 1500] == 0)


@pytest.mark.parametrize('name', CASES)
def test_optical_depth_bit_exact(name):
    c = load_case(name)
    tau = orc.optical_depth(c['P'], c['T_interface_tau'], c['nu'], c['tables'], list(c['q']))
    assert np.array_equal(tau[:, c['tau_sub_idx']], c['tau_sub'])
    assert np.allclose(tau.sum(axis=1), c['tau_rowsum'], rtol=1e-13, atol=0)


def test_optical_depth_running_sum_equals_reference_loop():
    c = load_case('two_mol_nz41')
    a = orc.optical_depth(c['P'], c['T_interface_tau'], c['nu'], c['tables'], list(c['q']))
    b = orc.optical_depth_reference_loop(c['P'], c['T_interface_tau'], c['nu'], c['tables'], list(c['q']))
    assert np.array_equal(a, b)


@pytest.mark.parametrize('name', CASES)
def test_interface_temperature(name):
    c = load_case(name)
    Ti = orc.interface_temperature(c['p'], c['T'], c['P'], c['T_g'])
    assert np.array_equal(Ti, c['T_interface'])
    A = orc.not_a_knot_matrix(c['p'], c['P'])
    Ti2 = A @ c['T']
    Ti2[-1] = c['T_g']
    assert np.abs(Ti2 - Ti).max() < 1e-10


@pytest.mark.parametrize('name', CASES)
def test_flux_reference_form(name):
    c = load_case(name)
    tau = orc.optical_depth(c['P'], c['T_interface_tau'], c['nu'], c['tables'], list(c['q']))
    up, dn = orc.get_flux(c['P'], c['T'], c['T_g'], tau, c['nu'], c['bands'], c['T_star'], c['R_star'],
                          c['dist'], c['albedo'])
    assert relmax(up, c['up']) < 1e-13
    assert relmax(dn, c['down']) < 1e-13
    net = orc.net_flux(up, dn, c['band_delta'])
    assert relmax(net, c['net']) < 1e-13
    assert relmax(orc.heating_rate(net, c['P']), c['heating']) < 1e-10


@pytest.mark.parametrize('name', CASES)
def test_flux_per_bin_form(name):
    c = load_case(name)
    I = orc.layer_integrand(c['P'], c['T_interface_tau'], c['nu'], c['tables'], list(c['q']))
    up, dn = orc.get_flux_per_bin(c['P'], c['T'], c['T_g'], I, c['nu'], c['bands'], c['T_star'],
                                  c['R_star'], c['dist'], c['albedo'])
    assert relmax(up, c['up']) < 1e-12
    assert relmax(dn, c['down']) < 1e-12
    net = orc.net_flux(up, dn, c['band_delta'])
    assert relmax(net, c['net']) < 1e-12
    assert relmax(orc.heating_rate(net, c['P']), c['heating']) < 1e-10


@pytest.mark.parametrize('name', ['co2_nz50_earth', 'two_mol_nz41'])
def test_transmission_matrix(name):
    c = load_case(name)
    tau = orc.optical_depth(c['P'], c['T_interface_tau'], c['nu'], c['tables'], list(c['q']))
    j = int(c['trans_band'])
    s, n = int(c['band_start'][j]), int(c['band_len'][j])
    Tr = orc.transmission(c['P'], c['P'], c['P'], c['nu'][s:s + n], c['band_delta'][j], c['nu'], tau)
    Tr = np.triu(np.nan_to_num(Tr, nan=0.0, posinf=0.0))
    assert relmax(Tr, c['trans_upper']) < 1e-14
    Tu = orc.band_transmission_matrix(tau, s, n, c['nu'], c['band_delta'][j])
    assert relmax(Tu, c['trans_upper']) < 1e-14


def test_dtransmission_dq():
    c = load_case('two_mol_nz41')
    tau = orc.optical_depth(c['P'], c['T_interface_tau'], c['nu'], c['tables'], list(c['q']))
    j = int(c['dtdq_band'])
    a, b = c['dtdq_rows']
    s, n = int(c['band_start'][j]), int(c['band_len'][j])
    d = orc.dtransmission_dq(c['P'][a:a + 1], c['P'][b:b + 1], c['P'], c['nu'][s:s + n], c['band_delta'][j],
                             c['nu'], tau, c['dtdq_k'])
    assert relmax(d, c['dtdq']) < 1e-14


def test_transmission_raises_on_unknown_level():
    c = load_case('single_line_nz30')
    tau = np.zeros((c['P'].size, c['nu'].size))
    with pytest.raises(ValueError):
        orc.transmission(np.array([1.2345]), c['P'], c['P'], c['nu'][:4], 30.0, c['nu'], tau)
