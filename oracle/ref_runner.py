"""TEST / BASELINE INFRASTRUCTURE ONLY -- never imported by the product package.

Drives the UNMODIFIED reference (``/root/reference`` in the build container, the staged copy
``oracle/_ref`` on the GPU box -- oracle/make_ref.py) over columns of bench.py's synthetic workload:
for every column the reference's own ``optical_depth`` (real_gas.py:86-127) rebuilds tau from the
column's interface temperatures and q profiles, ``RealGas.get_flux`` (real_gas.py:629-669) evaluates
the up / down band fluxes, and the net flux (:712-713) and heating rate (base.py:166-168) follow.
This is SURVEY.md 8(d)'s CPU baseline C1 (one process) / C2 (one process per host core).

Only bench.py (``--impl reference`` and the ``cpu_baseline`` leg) and tests call this.
"""
import contextlib
import io
import os
import tempfile
import time
import warnings

import numpy as np

from . import reference_shim as shim

NAMES = ['CO2_synth_bench', 'H2O_synth_bench']
_STATE = {}


def available():
    return shim.reference_available()


def which():
    """Where the reference that will be timed lives, and the staging digest if it is the staged copy."""
    root = shim.REFERENCE_ROOT
    note = os.path.join(root, 'STAGED.txt')
    if os.path.exists(note):
        return root + ' (' + open(note).read().strip().replace('\n', '; ') + ')'
    return root


def _array_q(p, values):
    """q(p) callable handed to the reference (real_gas.py:118 calls q_funcs[m](p, *args)): the column's
    profile on the interface pressures, which optical_depth passes in ascending order (:106).  The
    constructor also evaluates q on a fine construction grid (get_p_grid, :458) whose result an integer
    nz never uses: any other length gets zeros."""
    p = np.asarray(p)
    return values if p.shape == np.shape(values) else np.zeros_like(p, dtype=float)


def setup(w, T_g=288.0):
    """One reference ``RealGas`` on the workload's grids (the column state is assigned per evaluation).
    The two synthetic tables are written in the reference's own on-disk format (pickled dict .npy,
    hitran.py:315-356) to a scratch LookupTables folder."""
    if _STATE.get('key') == id(w):
        return _STATE
    warnings.filterwarnings('ignore')
    rg = shim.load_reference()
    folder = tempfile.mkdtemp(prefix='cmb200_ref_tables_')
    for name, table in zip(NAMES, w['tables']):
        np.save(os.path.join(folder, name + '.npy'), {k: np.asarray(v) for k, v in table.items()}, allow_pickle=True)
    shim.set_lookup_folder(folder)
    L = int(np.asarray(w['P']).size)
    zero_q = np.zeros(L)
    with contextlib.redirect_stdout(io.StringIO()), np.errstate(all='ignore'):
        gas = rg.RealGas(nz=L, ny=1, molecule_names=list(NAMES), T_g=T_g,
                         q_funcs={n: _array_q for n in NAMES}, q_funcs_args={n: (zero_q,) for n in NAMES},
                         n_nu_bands=int(np.asarray(w['bands']['centre']).size))
    # the reference must have built the grids the GPU plan was given
    assert np.array_equal(gas.p_interface[:, 0], np.asarray(w['P'])), 'pressure grids differ'
    assert np.array_equal(gas.nu, np.asarray(w['nu'])), 'wavenumber grids differ'
    assert np.array_equal(gas.nu_bands['centre'], np.asarray(w['bands']['centre'])), 'bands differ'
    assert np.array_equal(gas.nu_bands['sw'], np.asarray(w['bands']['sw']).astype(bool)), 'bands differ'
    import importlib
    const = importlib.import_module('Model.constants')
    _STATE.update(key=id(w), rg=rg, gas=gas, g=const.g, c_p=const.c_p_dry, folder=folder)
    return _STATE


def column(w, T, T_g, T_tau, q, albedo):
    """(net [L], heating [L-1], up [L x B], down [L x B]) of one column from the reference's code."""
    st = setup(w)
    rg, gas = st['rg'], st['gas']
    P = gas.p_interface[:, 0]
    with np.errstate(all='ignore'):
        gas.T = np.asarray(T, dtype=float).reshape(-1, 1).copy()
        gas.T_g = float(T_g)
        gas.albedo = np.array([float(albedo)])
        gas.tau_interface = rg.optical_depth(P, np.asarray(T_tau, dtype=float), gas.nu, gas.molecule_names,
                                             gas.q_funcs, {n: (np.asarray(q[m], dtype=float),) for m, n in enumerate(NAMES)})
        up, down = gas.get_flux()
    delta = gas.nu_bands['delta']
    net = np.sum(up * delta, axis=1) - np.sum(down * delta, axis=1)           # real_gas.py:712-713
    heating = st['g'] / st['c_p'] * np.diff(net) / np.diff(P)                  # base.py:166-168
    return net, heating, up, down


_W = {}


def _init(w):
    _W['w'] = w
    setup(w)


def _job(args):
    net, heating, _, _ = column(_W['w'], *args)
    return net, heating


def time_columns(w, cols, pool=None):
    """Seconds the reference needs for the given columns (T, T_g, T_tau, q [M, n, L], albedo) -- in this
    process, or spread over a multiprocessing pool created with ``initializer=_init, initargs=(w,)`` --
    and the (net, heating) results."""
    T, T_g, T_tau, q, albedo = cols
    jobs = [(T[i], T_g[i], T_tau[i], q[:, i], albedo[i]) for i in range(T.shape[0])]
    if pool is None:
        _init(w)
        t0 = time.perf_counter()
        res = [_job(j) for j in jobs]
    else:
        t0 = time.perf_counter()
        res = pool.map(_job, jobs, chunksize=1)
    return time.perf_counter() - t0, res
