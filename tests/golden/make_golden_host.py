"""Golden vectors for the host-side callers of the flux path, from the
UNMODIFIED reference.  Build container only:

    python tests/golden/make_golden_host.py

* ``host_convective.npz``: random unstable profiles -> reference ``convective_adjustment``.
* ``host_stepper.npz``:   the reference ``Atmosphere.update_temp`` (base.py:166-196) driven by a
  scripted, seeded sequence of net-flux arrays (a relaxing column with oscillating levels so that
  every branch of ``update_time_step`` -- shrink, grow, freeze a level -- is visited), with and
  without convective adjustment.  The CPU tests replay the same sequence through
  ``climatemodel_b200.radiation.base.Atmosphere``.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle.reference_shim import load_reference  # noqa: E402


def scripted_net_flux(step, T, p_interface, rng):
    """A made-up 'radiation scheme': net flux relaxing T towards a target profile, plus a level
    whose sign flips every step and seeded noise.  Pure function of its arguments."""
    nz = p_interface.shape[0]
    x = p_interface[:-1, 0] / p_interface[0, 0]
    target = np.maximum(310 * x ** 0.45, 205.0)                    # super-adiabatic troposphere
    target[-3:] += np.array([10.0, 20.0, 30.0])                     # stable lid (see main())
    drive = np.concatenate(([0.0], np.cumsum((T[:, 0] - target) * 0.05)))
    drive[nz // 3] += 3.0 * (-1) ** step
    drive += rng.normal(0, 1e-3, nz)
    if step > 40:
        drive[: nz // 4] = 0.0          # levels dropping below net_flux_thresh
    return drive.reshape(nz, 1)


def main():
    load_reference()
    import importlib
    base = importlib.import_module('Model.radiation.base')
    conv = importlib.import_module('Model.radiation.convective_adjustment')

    rng = np.random.default_rng(7)
    nz = 40
    p = np.logspace(np.log10(101320.0), np.log10(50.0), nz)
    profiles, adjusted, ps = [], [], []
    i = 0
    while len(profiles) < 24:
        i += 1
        T = 210 + 80 * (p / p[0]) ** rng.uniform(0.15, 0.6) + rng.normal(0, rng.uniform(0, 6), nz)
        T[-3:] += np.array([15.0, 30.0, 45.0])      # a stable lid: the reference indexes one level above an unstable run
        pp, TT = (p, T) if i % 2 == 0 else (p[::-1].copy(), T[::-1].copy())
        try:
            out = conv.convective_adjustment(pp, TT.reshape(-1, 1).copy())
        except IndexError:
            continue
        profiles.append(TT)
        ps.append(pp)
        adjusted.append(out[:, 0].copy())
    np.savez_compressed(os.path.join(HERE, 'host_convective.npz'), p=np.array(ps), T=np.array(profiles),
                        T_adjusted=np.array(adjusted))

    out = {}
    for tag, do_conv in [('plain', False), ('conv', True)]:
        nz = 31
        atm = base.Atmosphere(nz, 1, 1367.0 / 4)
        P = np.logspace(np.log10(101320.0), np.log10(20.0), nz).reshape(-1, 1)
        atm.p_interface = P
        atm.p = (P[:-1] + P[1:]) / 2
        x = atm.p / P[0]
        atm.T = np.maximum(300 * x ** 0.4, 210.0)                      # already super-adiabatic low down
        atm.T[-3:, 0] += np.array([5.0, 10.0, 15.0])
        T_init = atm.T[:, 0].copy()
        atm.net_flux = np.zeros((nz, 1))
        rng = np.random.default_rng(11)
        t, n_steps = 0.0, 90
        Ts, ts, dts, dnfs, nets, deltas = [], [], [], [], [], []
        for step in range(n_steps):
            net = scripted_net_flux(step, atm.T, P, rng)
            t, dnf = atm.update_temp(t, net, convective_adjust=do_conv)
            Ts.append(atm.T[:, 0].copy()); ts.append(t); dts.append(atm.time_step_info['dt'])
            dnfs.append(dnf); nets.append(net[:, 0].copy()); deltas.append(atm.time_step_info['DeltaT'])
        out.update({tag + '_T': np.array(Ts), tag + '_t': np.array(ts), tag + '_dt': np.array(dts),
                    tag + '_dnf': np.array(dnfs), tag + '_net': np.array(nets), tag + '_DeltaT': np.array(deltas),
                    tag + '_removed': np.array(atm.time_step_info['RemoveInd'], dtype=np.int64)})
        print(tag, 'removed', atm.time_step_info['RemoveInd'], 'DeltaT range', min(deltas), max(deltas))
    out.update(P=P[:, 0], T_init=T_init, T0=float(np.ravel(atm.T0)[0]))
    lat = np.linspace(-90, 90, 19)
    out.update(lat=lat, lat_factor=base.latitudinal_solar_distribution(lat),
               iso_T=base.get_isothermal_temp(0.3, 1367.0 / 4, lat))
    np.savez_compressed(os.path.join(HERE, 'host_stepper.npz'), **out)


def grey_cases():
    """``host_grey.npz``: inputs and outputs of the reference grey two-stream fluxes
    (GreyGas.get_lw_flux / get_sw_flux, grey.py:251-294) for a long-wave-only atmosphere and one
    with a short-wave absorber, 7 latitudes, perturbed temperatures."""
    import importlib
    load_reference()
    grey = importlib.import_module('Model.radiation.grey')
    god = importlib.import_module('Model.radiation.grey_optical_depth')
    out = {}
    rng = np.random.default_rng(5)
    p_s = 101320.0
    for tag, sw in [('lw', None), ('sw', god.exponential)]:
        gas = grey.GreyGas(nz=36, ny=7, tau_lw_func=god.scale_height, tau_lw_func_args=[0.22 * p_s, 4, p_s],
                           tau_sw_func=sw, tau_sw_func_args=None if sw is None else [0.3 * p_s, 0.4, p_s],
                           albedo=np.linspace(0.2, 0.5, 7))
        gas.T = gas.T + rng.normal(0, 12, gas.T.shape)
        gas.get_lw_flux()
        up_sw, down_sw = gas.get_sw_flux()
        iso = gas.get_sw_flux(isothermal=True)
        net = gas.up_lw_flux - gas.down_lw_flux + up_sw - down_sw
        out.update({tag + '_T': gas.T.copy(), tag + '_dtau': gas.dtau.copy(), tag + '_F_sw0': np.asarray((1 - gas.albedo_mod) * gas.solar_latitude_factor * gas.F_stellar_constant / 4),
                    tag + '_albedo_mod': np.asarray(gas.albedo_mod, dtype=float) * np.ones(7),
                    tag + '_solar_factor': np.asarray(gas.solar_latitude_factor, dtype=float) * np.ones(7),
                    tag + '_F_stellar': float(gas.F_stellar_constant),
                    tag + '_tau_sw_interface': np.zeros_like(gas.tau_interface) if sw is None else gas.tau_sw_interface.copy(),
                    tag + '_up_lw': gas.up_lw_flux.copy(), tag + '_down_lw': gas.down_lw_flux.copy(),
                    tag + '_up_sw': up_sw.copy(), tag + '_down_sw': down_sw.copy(), tag + '_net': net.copy(),
                    tag + '_up_sw_iso': iso[0].copy(), tag + '_down_sw_iso': iso[1].copy()})
    np.savez_compressed(os.path.join(HERE, 'host_grey.npz'), **out)
    print('grey', {k: np.shape(v) for k, v in out.items() if k.startswith('sw_')})


if __name__ == '__main__':
    if len(sys.argv) > 1 and sys.argv[1] == 'grey':
        grey_cases()
    else:
        main()
        grey_cases()
