"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

It imports the reference through ``oracle/reference_shim.py`` (stub ``sympl`` constants and an
empty ``matplotlib``), points its ``LookupTableFolder`` at a scratch folder holding the tables
committed under ``tests/golden/tables/`` (re-saved in the reference's pickled-dict ``.npy``
format) and dumps inputs + outputs of the hot-path functions as small ``.npz`` files.

Tables: ``CO2_no_sw``, ``gray`` and ``single_line`` are verbatim copies of the reference's
``HitranData/LookupTables/*.npy`` *data* (stored compressed); ``H2O_synth`` is produced here by
the reference's own line-by-line code (``hitran.get_absorption_coefficient``,
hitran.py:223-247) from a seeded synthetic line list.
"""
import contextlib
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle.reference_shim import load_reference, set_lookup_folder, default_lookup_folder  # noqa: E402

TABLE_DIR = os.path.join(HERE, 'tables')
SCRATCH = os.path.join(ROOT, 'build', 'scratch', 'lookup')


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
        return fn(*a, **k)


def copy_reference_tables():
    os.makedirs(TABLE_DIR, exist_ok=True)
    for name in ['CO2_no_sw', 'gray', 'single_line']:
        d = np.load(default_lookup_folder() + name + '.npy', allow_pickle=True).item()
        np.savez_compressed(os.path.join(TABLE_DIR, name + '.npz'), **d)


def make_synthetic_h2o(hitran):
    """H2O-like table from a seeded line list via the reference's Lorentz line-by-line code."""
    rng = np.random.default_rng(0)
    n_lines = 240
    centres = np.concatenate((rng.uniform(20, 600, 80), rng.uniform(1300, 2000, 80),
                              rng.uniform(3000, 4200, 80)))
    lines = {'nu': np.sort(centres),
             'sw': 10 ** rng.uniform(-3, 3, n_lines),
             'gamma_air': rng.uniform(0.05, 0.1, n_lines),
             'n_air': rng.uniform(0.5, 0.8, n_lines)}
    p_tab = np.logspace(np.log10(1.0132e5), np.log10(20), 24)
    T_tab = np.array([230, 260, 290, 320])
    nu_tab = np.arange(0, 4500 + 5, 10, dtype=float)
    k = np.zeros((p_tab.size, T_tab.size, nu_tab.size))
    for i, T in enumerate(T_tab):
        k[:, i, :] = quiet(hitran.get_absorption_coefficient, p_tab, np.ones_like(p_tab) * T, nu_tab,
                           'custom', lines)
    np.savez_compressed(os.path.join(TABLE_DIR, 'H2O_synth.npz'), p=p_tab, T=T_tab, nu=nu_tab,
                        absorption_coef=k)
    np.savez_compressed(os.path.join(TABLE_DIR, 'H2O_synth_lines.npz'), **lines)


def stage_tables_for_reference():
    os.makedirs(SCRATCH, exist_ok=True)
    for f in os.listdir(TABLE_DIR):
        if f.endswith('.npz') and not f.endswith('_lines.npz'):
            d = dict(np.load(os.path.join(TABLE_DIR, f)))
            np.save(os.path.join(SCRATCH, f[:-4] + '.npy'), d, allow_pickle=True)
    set_lookup_folder(SCRATCH)


def band_arrays(gas):
    nu = gas.nu
    rng = gas.nu_bands['range']
    start = np.array([int(np.where(nu == r[0])[0][0]) for r in rng], dtype=np.int64)
    length = np.array([len(r) for r in rng], dtype=np.int64)
    for s, n, r in zip(start, length, rng):
        assert np.array_equal(nu[s:s + n], r)
    return start, length


def dump_case(name, rg, gas, T_interface_tau, n_steps=0, conv=False, extra=None):
    P = gas.p_interface[:, 0].copy()
    start, length = band_arrays(gas)
    out = dict(
        P=P, p=gas.p[:, 0].copy(), nu=gas.nu.copy(), d_nu=gas.d_nu,
        nu_lw_max=gas.nu_lw.max(), nu_sw_min=gas.nu_sw.min(),
        band_start=start, band_len=length, band_centre=gas.nu_bands['centre'].copy(),
        band_delta=gas.nu_bands['delta'].copy(), band_sw=gas.nu_bands['sw'].copy(),
        T=gas.T[:, 0].copy(), T_g=float(gas.T_g), albedo=float(np.ravel(gas.albedo)[0]),
        T_star=float(gas.star['T']), R_star=float(gas.star['R']), dist=float(gas.star['star_planet_dist']),
        T_interface_tau=np.asarray(T_interface_tau, dtype=float).reshape(-1),
        molecule_names=np.array(gas.molecule_names),
        q=np.stack([gas.q_funcs[m](np.sort(P), *gas.q_funcs_args[m]) for m in gas.molecule_names]),
    )
    tau = gas.tau_interface
    sub = np.unique(np.concatenate((np.arange(0, tau.shape[1], 7), [tau.shape[1] - 1])))
    out['tau_sub_idx'] = sub
    out['tau_sub'] = tau[:, sub].copy()
    out['tau_rowsum'] = tau.sum(axis=1)
    from scipy.interpolate import InterpolatedUnivariateSpline
    Ti = InterpolatedUnivariateSpline(gas.p, gas.T)(gas.p_interface).reshape(-1)
    Ti[-1] = gas.T_g
    out['T_interface'] = Ti
    up, dn = gas.get_flux()
    net = np.sum(up * gas.nu_bands['delta'], axis=1) - np.sum(dn * gas.nu_bands['delta'], axis=1)
    from Model.constants import g, c_p_dry
    out.update(up=up, down=dn, net=net,
               heating=g / c_p_dry * (net[1:] - net[:-1]) / (P[1:] - P[:-1]))
    # one long-wave band: transmission matrix (upper triangle) and d/dq
    j = int(np.where(~gas.nu_bands['sw'])[0][-1])
    Tr = rg.transmission(P, P, P, gas.nu_bands['range'][j], gas.nu_bands['delta'][j], gas.nu, tau)
    out['trans_band'] = j
    out['trans_upper'] = np.triu(np.nan_to_num(Tr, nan=0.0, posinf=0.0))
    if n_steps:
        t = 0.0
        Ts, ts, dnf = [], [], []
        for _ in range(n_steps):
            t, d = quiet(gas.take_time_step, t, convective_adjust=conv)
            Ts.append(gas.T[:, 0].copy())
            ts.append(t)
            dnf.append(d)
        out.update(step_T=np.array(Ts), step_t=np.array(ts), step_dnf=np.array(dnf), step_conv=conv,
                   step_net_last=gas.net_flux.reshape(-1).copy())
    if extra:
        out.update(extra)
    np.savez_compressed(os.path.join(HERE, name + '.npz'), **out)
    print(name, 'L=%d N=%d B=%d' % (P.size, gas.nu.size, gas.n_nu_bands),
          'olr=%.6f' % float(np.sum(up[0] * gas.nu_bands['delta'])))


def main():
    rg = load_reference()
    import Model.radiation.real_gas_data.hitran as hitran
    import Model.radiation.real_gas_data.specific_humidity as hum
    import Model.radiation.real_gas_data.temperature_profiles as tp
    copy_reference_tables()
    make_synthetic_h2o(hitran)
    stage_tables_for_reference()

    co2 = dict(q_funcs={'CO2_no_sw': hum.co2}, q_funcs_args={'CO2_no_sw': (370, 80000)})

    gas = quiet(rg.RealGas, nz=50, ny=1, molecule_names=['CO2_no_sw'], T_g=288, T_func=tp.earth_temp, **co2)
    dump_case('co2_nz50_earth', rg, gas, tp.earth_temp(gas.p_interface[:, 0]), n_steps=6)

    gas = quiet(rg.RealGas, nz=50, ny=1, molecule_names=['CO2_no_sw'], T_g=288, T_func=tp.earth_temp, **co2)
    dump_case('co2_nz50_earth_conv', rg, gas, tp.earth_temp(gas.p_interface[:, 0]), n_steps=6, conv=True)

    gas = quiet(rg.RealGas, nz=100, ny=1, molecule_names=['CO2_no_sw'], T_g=288, **co2)
    dump_case('co2_nz100_iso', rg, gas, np.ones(100) * 288.0, n_steps=3)

    gas = quiet(rg.RealGas, nz=60, ny=1, molecule_names=['gray'], T_g=270, p_toa=10, n_nu_bands=200,
                q_funcs={'gray': hum.co2}, q_funcs_args={'gray': (370, 80000)})
    dump_case('gray_nz60_200b', rg, gas, np.ones(60) * 270.0)

    gas = quiet(rg.RealGas, nz=30, ny=1, molecule_names=['single_line'], T_g=280, n_nu_bands=40,
                q_funcs={'single_line': hum.co2}, q_funcs_args={'single_line': ()})
    dump_case('single_line_nz30', rg, gas, np.ones(30) * 280.0)

    def T_two(p):
        return tp.two_lapse_temp(p, h_tropopause=11, h_top=20, T_ground=295, lapse_trop=6.5)
    gas = quiet(rg.RealGas, nz=41, ny=1, molecule_names=['CO2_no_sw', 'H2O_synth'], T_g=295, T_func=T_two,
                q_funcs={'CO2_no_sw': hum.co2, 'H2O_synth': hum.h2o},
                q_funcs_args={'CO2_no_sw': (370, 80000), 'H2O_synth': (1,)})
    # dtransmission_dq for the synthetic molecule at one (p1, p2) pair
    j = 5
    k_row = rg.load_absorption_coef(gas.p_interface[20:21, 0], np.array([T_two(gas.p_interface[20, 0])]),
                                    gas.nu, 'H2O_synth')[0]
    P = gas.p_interface[:, 0]
    dT = rg.dtransmission_dq(P[10:11], P[25:26], P, gas.nu_bands['range'][j], gas.nu_bands['delta'][j],
                             gas.nu, gas.tau_interface, k_row)
    dump_case('two_mol_nz41', rg, gas, T_two(gas.p_interface[:, 0]), n_steps=4,
              extra=dict(dtdq_band=j, dtdq_rows=np.array([10, 25]), dtdq_k=k_row, dtdq=dT))

    # function-level vectors
    nu = np.array([10., 50., 667., 1500., 2590., 10000., 52050.])
    T = np.array([150., 200., 255., 288., 320., 5778.]).reshape(-1, 1)
    tab = np.load(os.path.join(SCRATCH, 'CO2_no_sw.npy'), allow_pickle=True).item()
    pq = np.array([5., 20., 21., 3000., 50000., 101320., 2e5])
    Tq = np.array([100., 259.9, 260., 260.1, 300., 340., 400.])
    nuq = np.array([0., 4.9, 5.0, 5.1, 666., 1495., 1500.])
    np.savez_compressed(os.path.join(HERE, 'functions.npz'),
                        planck_nu=nu, planck_T=T, planck_B=rg.B_wavenumber(nu, T),
                        gather_p=pq, gather_T=Tq, gather_nu=nuq,
                        gather_k=rg.get_absorption_coef(pq, Tq, nuq, tab),
                        load_nu=np.arange(1400., 1600., 10.),
                        load_k=rg.load_absorption_coef(pq, Tq, np.arange(1400., 1600., 10.), 'CO2_no_sw'))
    print('functions.npz written')


if __name__ == '__main__':
    main()
