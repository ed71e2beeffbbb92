"""GPU: grey two-stream fluxes against the unmodified reference GreyGas
(tests/golden/make_golden_host.py:grey_cases; grey.py:251-294)."""
import os

import numpy as np
import pytest

from util import GOLDEN, relmax

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('tag', ['lw', 'sw'])
def test_grey_fluxes_match_reference(tag):
    from climatemodel_b200.radiation.grey import grey_fluxes
    g = np.load(os.path.join(GOLDEN, 'host_grey.npz'))
    tau_sw = g[tag + '_tau_sw_interface'] if tag == 'sw' else None
    out = grey_fluxes(g[tag + '_T'], g[tag + '_dtau'], g[tag + '_albedo_mod'], g[tag + '_solar_factor'],
                      float(g[tag + '_F_stellar']), tau_sw_interface=tau_sw)
    for ours, ref in [('up_lw_flux', 'up_lw'), ('down_lw_flux', 'down_lw'), ('up_sw_flux', 'up_sw'),
                      ('down_sw_flux', 'down_sw'), ('net_flux', 'net')]:
        assert relmax(out[ours], g[tag + '_' + ref]) < 1e-13, ours
    assert np.array_equal(out['down_lw_flux'][-1], np.zeros(7))
    iso = grey_fluxes(g[tag + '_T'], g[tag + '_dtau'], g[tag + '_albedo_mod'], g[tag + '_solar_factor'],
                      float(g[tag + '_F_stellar']), tau_sw_interface=tau_sw, isothermal=True)
    assert relmax(iso['up_sw_flux'], g[tag + '_up_sw_iso']) < 1e-15
    assert relmax(iso['down_sw_flux'], g[tag + '_down_sw_iso']) < 1e-15


def test_grey_many_columns_and_errors():
    from climatemodel_b200.radiation.grey import grey_fluxes
    g = np.load(os.path.join(GOLDEN, 'host_grey.npz'))
    rep = 3000
    T, dtau = np.tile(g['sw_T'], (1, rep)), np.tile(g['sw_dtau'], (1, rep))
    out = grey_fluxes(T, dtau, np.tile(g['sw_albedo_mod'], rep), np.tile(g['sw_solar_factor'], rep),
                      float(g['sw_F_stellar']), tau_sw_interface=np.tile(g['sw_tau_sw_interface'], (1, rep)))
    assert np.array_equal(out['net_flux'][:, :7], out['net_flux'][:, -7:])
    assert relmax(out['net_flux'][:, :7], g['sw_net']) < 1e-13
    with pytest.raises(ValueError):
        grey_fluxes(T, dtau[:-1], 0.3, 1.0, 1367.0)
