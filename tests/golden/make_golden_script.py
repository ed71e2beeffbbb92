"""Golden run of ``real_gas_script.py`` AS WRITTEN (BASELINE configs[1]) from the UNMODIFIED reference.
Build container only (needs /root/reference):

    python tests/golden/make_golden_script.py [script|compos|findtg ...]

``script``  real_gas_script.py:46-54 with the one shipped table that covers it (the four HITRAN tables of
            the script are stripped from the repository): ``RealGas(nz='auto', ny=1, ['CO2_no_sw'],
            T_g=265.19, p_toa=0.1, temp_change=1, delta_temp_change=0.1)``, ``save_data`` with the four
            flux histories, ``evolve_to_equilibrium(data, flux_thresh=1e-3, t_end=2.0)``.  Stored: the
            automatic pressure grid, the bands, every step's t, the temperature and flux histories
            (every 25th step and the last), the step count and the final state.
``compos``  ``evolve_change_compos`` (real_gas.py:748-785) over two (T_g, CO2) pairs at nz = 30.
``findtg``  ``find_Tg`` (real_gas.py:530-562) at nz = 30.
Outputs: tests/golden/script_*.npz.
"""
import contextlib
import io
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from make_golden import stage_tables_for_reference  # noqa: E402
from oracle.reference_shim import load_reference  # noqa: E402

CO2 = dict(q_funcs_args={'CO2_no_sw': (370, 80000)})
KEEP = 25


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()), np.errstate(all='ignore'):
        return fn(*a, **k)


def bands_of(gas):
    return dict(band_centre=gas.nu_bands['centre'].copy(), band_delta=gas.nu_bands['delta'].copy(),
                band_sw=gas.nu_bands['sw'].copy(), band_len=np.array([len(r) for r in gas.nu_bands['range']]))


def history(data):
    n = len(data['t'])
    keep = sorted(set(list(range(0, n, KEEP)) + [n - 1]))
    out = {'t': np.array(data['t'], dtype=float), 'kept': np.array(keep),
           'T_kept': np.stack([np.asarray(data['T'][i])[:, 0] for i in keep])}
    if 'flux' in data:
        for key, series in data['flux'].items():
            out['flux_' + key] = np.stack([np.asarray(series[i]) for i in keep])
    return out


def run_script(rg, hum):
    t0 = time.time()
    gas = quiet(rg.RealGas, nz='auto', ny=1, molecule_names=['CO2_no_sw'], T_g=265.19, p_toa=0.1, temp_change=1,
                delta_temp_change=0.1, q_funcs={'CO2_no_sw': hum.co2}, **CO2)
    out = dict(nz=gas.nz, P=gas.p_interface[:, 0].copy(), nu_first_last_n=np.array([gas.nu[0], gas.nu[-1], gas.nu.size]),
               T0=gas.T[:, 0].copy(), net0=np.asarray(gas.net_flux).reshape(-1).copy(), **bands_of(gas))
    data = {'t': [], 'T': [], 'flux': {'lw_up': [], 'lw_down': [], 'sw_up': [], 'sw_down': []}}
    data = gas.save_data(data, 0)
    data = quiet(gas.evolve_to_equilibrium, data, flux_thresh=1e-3, convective_adjust=False, t_end=2.0)
    out.update(history(data))
    out.update(n_steps=len(data['t']) - 1, T_final=gas.T[:, 0].copy(), net_final=np.asarray(gas.net_flux).reshape(-1).copy())
    np.savez_compressed(os.path.join(HERE, 'script_real_gas.npz'), **out)
    print('script_real_gas.npz: nz %d, %d steps, t_end %.4g s, %.0f s of reference time' % (gas.nz, out['n_steps'], data['t'][-1], time.time() - t0))


def run_compos(rg, hum):
    gas = quiet(rg.RealGas, nz=30, ny=1, molecule_names=['CO2_no_sw'], T_g=270.0, temp_change=1, delta_temp_change=0.1,
                q_funcs={'CO2_no_sw': hum.co2}, **CO2)
    T_g = [270.0, 276.0]
    q_args = [{'CO2_no_sw': (370, 80000)}, {'CO2_no_sw': (1480, 80000)}]
    data = quiet(gas.evolve_change_compos, T_g, q_args, None, flux_thresh=1e-2, convective_adjust=False, t_end=0.3)
    out = dict(P=gas.p_interface[:, 0].copy(), T_g=np.array(T_g), ppmv=np.array([370.0, 1480.0]), **bands_of(gas))
    out.update(history(data))
    out.update(n_steps=len(data['t']) - 1, T_final=gas.T[:, 0].copy(), net_final=np.asarray(gas.net_flux).reshape(-1).copy())
    np.savez_compressed(os.path.join(HERE, 'script_change_compos.npz'), **out)
    print('script_change_compos.npz: %d steps' % out['n_steps'])


def run_findtg(rg, hum):
    gas = quiet(rg.RealGas, nz=30, ny=1, molecule_names=['CO2_no_sw'], T_g=270.0, temp_change=2, delta_temp_change=0.1,
                q_funcs={'CO2_no_sw': hum.co2}, **CO2)
    tries = []
    orig = gas.evolve_to_equilibrium

    def spy(*a, **k):
        tries.append(float(np.ravel(gas.T_g)[0]))
        return orig(*a, **k)
    gas.evolve_to_equilibrium = spy
    root = quiet(gas.find_Tg, flux_thresh=0.1, tol=0.5, convective_adjust=False, print_eqb_info=False)
    np.savez_compressed(os.path.join(HERE, 'script_find_Tg.npz'), root=float(root), tries=np.array(tries),
                        T_final=gas.T[:, 0].copy(), net_final=np.asarray(gas.net_flux).reshape(-1).copy())
    print('script_find_Tg.npz: root %.6f after %d tries' % (root, len(tries)))


def main():
    which = sys.argv[1:] or ['findtg', 'compos', 'script']
    rg = load_reference()
    import Model.radiation.real_gas_data.specific_humidity as hum
    stage_tables_for_reference()
    for name in which:
        {'script': run_script, 'compos': run_compos, 'findtg': run_findtg}[name](rg, hum)


if __name__ == '__main__':
    main()
