"""BASELINE configs[0] -- grey radiation, `radiation_script.py`: the reference's `GreyGas` timed on the host cores,
and its two flux routines next to the B200 grey kernel on the same arrays.

    python tools/bench_grey.py [--out profiles/r02_grey.json] [--no-gpu]

(1) reference time-stepping, SURVEY 8(d) "Config 1": `GreyGas(nz=100, ny, exponential long-wave + peak short-wave
    absorber).evolve_to_equilibrium(flux_thresh=0.1)` (`radiation_script.py:24-26, 43-46`; `nz='auto'` does not
    terminate in this container), ny = 1 and ny = 30: steps, seconds, steps/s.  The UNMODIFIED reference,
    from `oracle/_ref` (or `/root/reference`), through `oracle/reference_shim.py`.
(2) one flux evaluation of that model's state, `get_lw_flux` + `get_sw_flux` (`grey.py:251-294`), reference (NumPy,
    vectorised over latitude) against `cmb200_grey_flux_host` (host buffers, copies inside the call), at the
    model's ny and tiled to many columns; the two results are compared (parity of the grey kernel).

Test / baseline infrastructure: the product never imports `oracle`.
"""
import argparse
import contextlib
import io
import json
import os
import sys
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def reference_model(ny, nz=100):
    import importlib
    from oracle import reference_shim as shim
    shim.load_reference()
    grey = importlib.import_module('Model.radiation.grey')
    od = importlib.import_module('Model.radiation.grey_optical_depth')
    with contextlib.redirect_stdout(io.StringIO()), np.errstate(all='ignore'):
        return grey.GreyGas(nz=nz, ny=ny, tau_lw_func=od.exponential, tau_lw_func_args=[100000, 4],
                            tau_sw_func=od.peak_in_atmosphere, tau_sw_func_args=[30000, 2000, 0.5])


def time_stepping(ny):
    gas = reference_model(ny)
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()), np.errstate(all='ignore'):
        data = gas.evolve_to_equilibrium(flux_thresh=0.1)
    dt = time.perf_counter() - t0
    steps = len(data['t']) - 1
    return {'ny': ny, 'nz': 100, 'steps': steps, 'seconds': dt, 'steps_per_s': steps / dt,
            'column_steps_per_s': steps * ny / dt}


def reference_fluxes(gas):
    gas.get_lw_flux()
    up_sw, down_sw = gas.get_sw_flux()
    return gas.up_lw_flux - gas.down_lw_flux + up_sw - down_sw


def flux_evaluation(ny, tile, use_gpu, reps=20):
    """One flux evaluation of the reference model's state at ny * tile columns, both sides."""
    gas = reference_model(ny)
    rng = np.random.default_rng(3)
    gas.T = gas.T + rng.normal(0, 8, gas.T.shape)
    if tile > 1:                                    # more latitudes: the reference is vectorised over its last axis
        for name, v in list(vars(gas).items()):
            if isinstance(v, np.ndarray) and v.ndim == 2 and v.shape[1] == ny:
                setattr(gas, name, np.ascontiguousarray(np.tile(v, (1, tile))))
            elif isinstance(v, np.ndarray) and v.ndim == 1 and v.size == ny:
                setattr(gas, name, np.tile(v, tile))
        gas.ny = ny * tile
    n_col = gas.T.shape[1]
    with np.errstate(all='ignore'):
        net_ref = reference_fluxes(gas)
        t0 = time.perf_counter()
        for _ in range(reps):
            reference_fluxes(gas)
        ref_s = (time.perf_counter() - t0) / reps
    rec = {'columns': n_col, 'levels': int(gas.T.shape[0]) + 1, 'reference_ms': ref_s * 1e3,
           'reference_column_evaluations_per_s': n_col / ref_s}
    if use_gpu:
        from climatemodel_b200.radiation.grey import grey_fluxes
        args = (gas.T, gas.dtau, np.asarray(gas.albedo_mod, dtype=float) * np.ones(n_col),
                np.asarray(gas.solar_latitude_factor, dtype=float) * np.ones(n_col), float(gas.F_stellar_constant))
        kw = dict(tau_sw_interface=gas.tau_sw_interface)
        out = grey_fluxes(*args, **kw)
        t0 = time.perf_counter()
        for _ in range(reps):
            grey_fluxes(*args, **kw)
        gpu_s = (time.perf_counter() - t0) / reps
        rec.update({'b200_host_call_ms': gpu_s * 1e3, 'b200_column_evaluations_per_s': n_col / gpu_s,
                    'net_flux_rel_err_vs_reference': float(np.abs(out['net_flux'] - net_ref).max() / np.abs(net_ref).max()),
                    'note': 'cmb200_grey_flux_host: host arrays in, five [nz, ny] host arrays out, copies and allocation inside the call'})
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default=None)
    ap.add_argument('--no-gpu', action='store_true')
    a = ap.parse_args()
    warnings.filterwarnings('ignore')
    from oracle import ref_runner
    rec = {'workload': 'BASELINE configs[0]: grey radiation single-column radiative-equilibrium time-stepping (radiation_script.py)',
           'reference': ref_runner.which(), 'host_cores_used': 1,
           'time_stepping': [time_stepping(1), time_stepping(30)],
           'flux_evaluation': [flux_evaluation(30, 1, not a.no_gpu), flux_evaluation(30, 4000, not a.no_gpu, reps=5)]}
    print(json.dumps(rec))
    if a.out:
        with open(a.out, 'w') as f:
            json.dump(rec, f, indent=1)


if __name__ == '__main__':
    main()
