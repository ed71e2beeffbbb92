"""GPU: the line-by-line table builder against a table the unmodified
reference built from the same seeded line list (tests/golden/make_golden.py:make_synthetic_h2o,
reference code hitran.py:223-247)."""
import os

import numpy as np
import pytest

from util import TABLES, load_table

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def lines():
    z = np.load(os.path.join(TABLES, 'H2O_synth_lines.npz'))
    return {k: z[k] for k in z.files}


def test_table_matches_reference_built_table(lines):
    from climatemodel_b200.radiation.real_gas_data import hitran
    ref = load_table('H2O_synth')
    tab = hitran.make_table(dict(lines), p_array=ref['p'], T_array=ref['T'], wavenumber_array=ref['nu'], save=False)
    assert tab['absorption_coef'].shape == ref['absorption_coef'].shape
    assert np.array_equal(tab['p'], ref['p']) and np.array_equal(tab['nu'], ref['nu'])
    got, want = tab['absorption_coef'], ref['absorption_coef']
    assert np.array_equal(got == 0, want == 0)                         # same line windows, bin for bin
    nz = want != 0
    assert np.abs(got[nz] / want[nz] - 1).max() < 1e-12                # element-wise, over 12 decades


def test_single_level_and_window_edges(lines):
    """One level, few line widths kept: most of the grid is outside every window and must be 0."""
    from climatemodel_b200.radiation.real_gas_data import hitran
    nu = np.arange(0, 4500 + 5, 10, dtype=float)
    p, T = np.array([5.0e4]), np.array([260.0])
    k = hitran.get_absorption_coefficient(p, T, nu, 'custom', dict(lines), n_line_widths=300)
    want = np.zeros((1, nu.size))
    for i in range(lines['nu'].size):
        g = hitran.gamma_extrapolate(p, T, lines['gamma_air'][i], lines['n_air'][i])
        s = hitran.s_extrapolate(T, lines['sw'][i], lines['nu'][i], lines['n_air'][i])
        i1, i2 = hitran.wavenumbers_near_line(nu, 10.0, lines['nu'][i], g.max(), 300)
        want[:, i1:i2] += s.reshape(-1, 1) * hitran.lorentzian_profile(nu[i1:i2].reshape(1, -1), lines['nu'][i],
                                                                       g.reshape(-1, 1))
    assert np.array_equal(k == 0, want == 0) and (want == 0).sum() > 50
    nz = want != 0
    assert np.abs(k[nz] / want[nz] - 1).max() < 1e-12


def test_no_lines_and_bad_arguments():
    from climatemodel_b200.radiation.real_gas_data import hitran
    nu = np.arange(0., 100., 10.)
    empty = {k: np.zeros(0) for k in ('nu', 'sw', 'gamma_air', 'n_air')}
    assert np.all(hitran.get_absorption_coefficient(np.array([1e5]), np.array([300.]), nu, 'custom', empty) == 0)
    with pytest.raises(ValueError):
        hitran.get_absorption_coefficient(np.array([1e5, 1e4]), np.array([300.]), nu, 'custom', empty)
    with pytest.raises(ValueError):
        hitran.make_table(empty, wavenumber_array=None, save=False)
