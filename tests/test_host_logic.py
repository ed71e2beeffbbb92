"""CPU tests: host-side logic of the product package and the C-ABI library surface
(no compute calls -- there is no GPU in the build container)."""
import os
import re

import numpy as np
import pytest

import oracle.realgas_oracle as orc
from util import CASES, load_case

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, 'include', 'climatemodel_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(cmb200_[a-z0-9_]+)\s*\(', text)))


def test_library_loads_and_exports_every_declared_symbol():
    from climatemodel_b200 import _lib
    lib = _lib.load()
    names = header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), 'symbol %s declared in the header is not exported' % n
        assert n in _lib.SIGNATURES, 'symbol %s has no ctypes prototype' % n
    assert set(_lib.SIGNATURES) == set(names)
    assert lib.cmb200_version() >= 100


def test_struct_layout_matches_header_field_order():
    from climatemodel_b200 import _lib
    text = open(os.path.join(ROOT, 'include', 'climatemodel_b200.h')).read()
    body = text[text.index('typedef struct {\n    int64_t n_col;'):text.index('} cmb200_flux_io;')]
    body = re.sub(r'/\*.*?\*/', '', body, flags=re.S)
    fields = re.findall(r'\*?\s*\*?([A-Za-z_]+)\s*;', body)
    assert fields == [f[0] for f in _lib.FluxIO._fields_]


def test_no_cpu_fallback_without_device():
    """In the GPU-less build container every compute entry point must raise, not fall back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    from climatemodel_b200.radiation import real_gas as rg
    from climatemodel_b200._lib import CudaLibraryError
    with pytest.raises(CudaLibraryError):
        rg.B_wavenumber(np.array([100., 200.]), 288.0)
    c = load_case('single_line_nz30')
    from climatemodel_b200.plan import FluxPlan
    with pytest.raises(CudaLibraryError):
        FluxPlan(c['P'], c['nu'], c['bands'])


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'climatemodel_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle', src, flags=re.M), f
                assert '/root/reference' not in src, f


@pytest.mark.parametrize('name', CASES)
def test_wavenumber_grid_and_bands_match_reference(name):
    from climatemodel_b200.radiation import grids
    c = load_case(name)
    nu, nu_lw, nu_ov, nu_sw = grids.wavenumber_array(float(c['d_nu']), c['T_star'], c['T_g'],
                                                     planck=orc.planck_wavenumber)
    assert np.array_equal(nu, c['nu'])
    assert nu_lw.max() == float(c['nu_lw_max']) and nu_sw.min() == float(c['nu_sw_min'])
    bands = grids.wavenumber_bands(nu_lw, nu_ov, nu_sw, len(c['band_centre']), c['T_star'], c['T_g'],
                                   planck=orc.planck_wavenumber)
    from climatemodel_b200.plan import bands_from_nu_bands
    flat = bands_from_nu_bands(nu, bands)
    assert np.array_equal(flat['start'], c['band_start'])
    assert np.array_equal(flat['len'], c['band_len'])
    assert np.array_equal(flat['centre'], c['band_centre'])
    assert np.array_equal(flat['delta'], c['band_delta'])
    assert np.array_equal(flat['sw'].astype(bool), c['band_sw'])


@pytest.mark.parametrize('name', ['co2_nz50_earth', 'co2_nz100_iso', 'gray_nz60_200b', 'two_mol_nz41'])
def test_fixed_pressure_grid_matches_reference(name):
    from climatemodel_b200.radiation import grids
    c = load_case(name)
    p_toa = 10 if name.startswith('gray') else 20
    P = grids.fixed_pressure_grid(c['P'].size, 1.0132e5, p_toa)
    assert np.array_equal(P, c['P'])


def test_bands_from_nu_bands_rejects_non_contiguous():
    from climatemodel_b200.plan import bands_from_nu_bands
    nu = np.arange(0., 100., 10.)
    with pytest.raises(ValueError):
        bands_from_nu_bands(nu, {'range': [np.array([0., 20.])], 'centre': [0.], 'delta': [20.], 'sw': [False]})


def test_emission_integrals_against_reference_vectors():
    """``real_gas.emission_integrals`` (the host half of ``RealGas.flux_integrals``, real_gas.py:564-627)
    against tests/golden/flux_integrals.npz, produced by the unmodified reference's method for two
    long-wave bands of two configurations, with the spline interface temperatures and a perturbed set."""
    import os
    import oracle.realgas_oracle as orc
    from util import GOLDEN
    from climatemodel_b200.radiation.real_gas import emission_integrals
    g = np.load(os.path.join(GOLDEN, 'flux_integrals.npz'))
    n = 0
    for name in ('co2_nz50_earth', 'two_mol_nz41'):
        P, p, T, T_g, centre = (g[name + '_' + k] for k in ('P', 'p', 'T', 'T_g', 'centre'))
        for tag in ('a', 'b'):
            key = '%s_%s_' % (name, tag)
            c = centre[int(g[key + 'band'])]
            beta = lambda x: np.pi * orc.planck_wavenumber(c, np.asarray(x, dtype=float))
            trans = g[key + 'trans_upper'].copy()
            trans[np.tril_indices(P.size, -1)] = np.inf          # the reference's lower triangle overflows: must not be read
            for ttag in ('spline', 'pert'):
                up, dn = emission_integrals(trans, P, p, beta(T), beta(g[key + ttag + '_T_interface']), beta(T_g))
                for got, want in ((up, g[key + ttag + '_up']), (dn, g[key + ttag + '_down'])):
                    assert np.isfinite(got).all()
                    assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max()
                assert up[-1] == 0 and dn[0] == 0
                n += 1
    assert n == 8
