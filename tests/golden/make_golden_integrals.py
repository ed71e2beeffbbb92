"""Golden vectors for ``RealGas.flux_integrals(j, T_interface)`` (real_gas.py:564-627) from the
UNMODIFIED reference.  Build container only (needs /root/reference):

    python tests/golden/make_golden_integrals.py

For two of the configurations of make_golden.py it calls the reference method for three long-wave
bands with the spline interface temperatures and with a perturbed set, and stores the inputs the
method reads (band transmission matrix -- upper triangle, the lower one overflows --, grids,
temperatures) next to its outputs in ``flux_integrals.npz``.

Band choice: the integrals are sums of beta * (Trans(i,k+1) - Trans(i,k)), so a band whose
whole-column transmission is 1 - 1e-11 yields integrals that are rounding noise of the
reference's own evaluation (and a band with Trans == 1 yields exact zeros).  Tags 'a' and 'b'
are therefore the two long-wave bands with the largest absorption 1 - Trans(0, L-1) (well
conditioned: the 1e-10 gate applies); tag 'c' is the most transparent band that still absorbs
(1 - Trans > 0), kept to exercise the conditioning-aware bound of the test.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from make_golden import quiet, stage_tables_for_reference  # noqa: E402
from oracle.reference_shim import load_reference  # noqa: E402


def main():
    rg = load_reference()
    import Model.radiation.real_gas_data.specific_humidity as hum
    import Model.radiation.real_gas_data.temperature_profiles as tp
    from scipy.interpolate import InterpolatedUnivariateSpline
    stage_tables_for_reference()

    def T_two(p):
        return tp.two_lapse_temp(p, h_tropopause=11, h_top=20, T_ground=295, lapse_trop=6.5)
    cases = {
        'co2_nz50_earth': quiet(rg.RealGas, nz=50, ny=1, molecule_names=['CO2_no_sw'], T_g=288, T_func=tp.earth_temp,
                                q_funcs={'CO2_no_sw': hum.co2}, q_funcs_args={'CO2_no_sw': (370, 80000)}),
        'two_mol_nz41': quiet(rg.RealGas, nz=41, ny=1, molecule_names=['CO2_no_sw', 'H2O_synth'], T_g=295, T_func=T_two,
                              q_funcs={'CO2_no_sw': hum.co2, 'H2O_synth': hum.h2o},
                              q_funcs_args={'CO2_no_sw': (370, 80000), 'H2O_synth': (1,)}),
    }
    out = {}
    rng = np.random.default_rng(3)
    for name, gas in cases.items():
        P = gas.p_interface[:, 0]
        lw = np.where(~gas.nu_bands['sw'])[0]
        Ti = InterpolatedUnivariateSpline(gas.p, gas.T)(gas.p_interface).reshape(-1)
        Ti[-1] = gas.T_g
        Ti_pert = Ti + rng.normal(0, 2.0, Ti.size)
        absorb = np.array([1.0 - rg.transmission(P[:1], P[-1:], P, gas.nu_bands['range'][j], gas.nu_bands['delta'][j],
                                                 gas.nu, gas.tau_interface)[0, 0] for j in lw])
        order = np.argsort(-absorb)
        weakest = [k for k in order if absorb[k] > 0][-1]
        for tag, j in (('a', int(lw[order[0]])), ('b', int(lw[order[1]])), ('c', int(lw[weakest]))):
            Tr = rg.transmission(P, P, P, gas.nu_bands['range'][j], gas.nu_bands['delta'][j], gas.nu, gas.tau_interface)
            key = '%s_%s_' % (name, tag)
            out[key + 'band'] = j
            out[key + 'absorption'] = 1.0 - Tr[0, -1]
            out[key + 'trans_upper'] = np.triu(np.nan_to_num(Tr, nan=0.0, posinf=0.0))
            for ttag, T_int in (('spline', Ti), ('pert', Ti_pert)):
                with np.errstate(all='ignore'):
                    up, dn = gas.flux_integrals(j, T_int.reshape(-1, 1))      # [nz x 1] as get_flux passes it (:634)
                out[key + ttag + '_T_interface'] = T_int
                out[key + ttag + '_up'] = up
                out[key + ttag + '_down'] = dn
        out[name + '_P'] = P.copy()
        out[name + '_p'] = gas.p[:, 0].copy()
        out[name + '_T'] = gas.T[:, 0].copy()
        out[name + '_T_g'] = float(gas.T_g)
        out[name + '_centre'] = gas.nu_bands['centre'].copy()
    np.savez_compressed(os.path.join(HERE, 'flux_integrals.npz'), **out)
    print('flux_integrals.npz written:', len(out), 'arrays')


if __name__ == '__main__':
    main()
