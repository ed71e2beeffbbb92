"""Drop-in mirror of the reference's ``Model/radiation/real_gas.py`` module-level API
(real_gas.py:14-184): same names, argument meaning, return shapes and error behaviour, with the
arithmetic done by the CUDA library through its C ABI.  There is no NumPy fallback -- without
the library or a GPU these functions raise.

The ``RealGas`` class lives in ``real_gas_model.py`` and is re-exported here.
"""
import ctypes as C

import numpy as np

from .. import _lib
from .._lib import as_f64, check, dptr
from ..constants import g
from .real_gas_data import hitran as _hitran


def _i32p(a):
    return a.ctypes.data_as(_lib.c_int32_p)


# ---- A1 -----------------------------------------------------------------------------------
def B_wavenumber(nu, T):
    """Planck function per wavenumber (cm^-1); broadcasts like the reference
    (real_gas.py:30-42)."""
    nu_b, T_b = np.broadcast_arrays(np.asarray(nu, dtype=float), np.asarray(T, dtype=float))
    shape = nu_b.shape
    nu_f = np.ascontiguousarray(nu_b, dtype=np.float64).reshape(-1)
    T_f = np.ascontiguousarray(T_b, dtype=np.float64).reshape(-1)
    # element-wise call: one "temperature" per element keeps arbitrary broadcasting exact
    out = np.empty(nu_f.size)
    if nu_f.size:
        cst = _lib.constants_struct()
        lib = _lib.load()
        if np.all(T_f == T_f[0]):
            T1 = as_f64([T_f[0]])
            check(lib.cmb200_planck_wavenumber_host(C.byref(cst), dptr(nu_f), nu_f.size, dptr(T1), 1, dptr(out)))
        else:
            # general case: evaluate the outer product row by row over distinct temperatures
            Tu, inv = np.unique(T_f, return_inverse=True)
            for i, Tv in enumerate(Tu):
                sel = np.where(inv == i)[0]
                nus = np.ascontiguousarray(nu_f[sel])
                o = np.empty(sel.size)
                T1 = as_f64([Tv])
                check(lib.cmb200_planck_wavenumber_host(C.byref(cst), dptr(nus), nus.size, dptr(T1), 1, dptr(o)))
                out[sel] = o
    out = out.reshape(shape)
    return out if out.ndim else float(out)


def B_freq(freq, T):
    """Planck function per frequency (s^-1) (real_gas.py:14-27): B_wavenumber / (100 c) at
    nu = freq / (100 c)."""
    from ..constants import speed_of_light
    return B_wavenumber(np.asarray(freq, dtype=float) / (100 * speed_of_light), T) / (100 * speed_of_light)


# ---- A2 -----------------------------------------------------------------------------------
def _gather(p, T, nu, table, zero_outside):
    p = as_f64(p).reshape(-1)
    T = as_f64(T).reshape(-1)
    nu = as_f64(nu).reshape(-1)
    if T.size != p.size:
        raise ValueError('p and T must have the same number of levels')
    keep = []
    tab = _lib.table_struct(table, keep)
    out = np.zeros((p.size, nu.size))
    check(_lib.load().cmb200_absorption_coef_host(C.byref(tab), dptr(p), dptr(T), p.size, dptr(nu), nu.size,
                                                  1 if zero_outside else 0, dptr(out)))
    return out


def get_absorption_coef(p, T, nu, absorb_coef_dict):
    """Nearest-neighbour lookup k[np_crop x n_wavenumber] (real_gas.py:45-62)."""
    return _gather(p, T, nu, absorb_coef_dict, zero_outside=False)


def load_absorption_coef(p, T, wavenumber, molecule_name):
    """As above from the molecule's table file, zero outside its wavenumber range
    (real_gas.py:65-83)."""
    return _gather(p, T, wavenumber, _hitran.load_table(molecule_name), zero_outside=True)


# ---- A3 -----------------------------------------------------------------------------------
def optical_depth(p, T, wavenumber, molecule_names, q_funcs, q_funcs_args):
    """tau[np x n_wavenumber], tau[-1] the surface value (real_gas.py:86-127).  The q_funcs are
    host callables exactly as in the reference; everything per (level, bin) runs on the GPU and
    is bit-identical to the reference's trapezoid."""
    p = np.sort(as_f64(p).reshape(-1))
    T = as_f64(T).reshape(-1)
    nu = as_f64(wavenumber).reshape(-1)
    names = list(molecule_names)
    tables = [_hitran.load_table(m) for m in names]
    q = np.zeros((max(len(names), 1), p.size))
    for i, m in enumerate(names):
        q[i] = np.asarray(q_funcs[m](p, *q_funcs_args[m]), dtype=float).reshape(-1)
    return optical_depth_from_tables(p, T, nu, tables, q[:len(names)])


def optical_depth_from_tables(p_sorted, T, nu, tables, q):
    """optical_depth with the table dicts and q[M, np] passed in directly."""
    p = as_f64(p_sorted).reshape(-1)
    T = as_f64(T).reshape(-1)
    nu = as_f64(nu).reshape(-1)
    q = as_f64(q).reshape(len(tables), p.size) if len(tables) else np.zeros((1, p.size))
    keep = []
    arr = (_lib.Table * max(len(tables), 1))(*[_lib.table_struct(t, keep) for t in tables])
    tau = np.zeros((p.size, nu.size))
    cst = _lib.constants_struct()
    check(_lib.load().cmb200_optical_depth_host(C.byref(cst), len(tables), arr, dptr(p), dptr(T), p.size,
                                                dptr(nu), nu.size, dptr(q), dptr(tau)))
    return tau


# ---- A4 / A5 -----------------------------------------------------------------------------
def _locate(p1, p2, p_all, nu_band, nu_all):
    """Index lookups by exact float equality, rows returned in p_all / nu_all order, and the
    ValueError of real_gas.py:143-147."""
    p_all = np.asarray(p_all, dtype=float).reshape(-1, 1)
    nu_all = np.asarray(nu_all, dtype=float).reshape(-1, 1)
    i1 = np.where(p_all == np.asarray(p1, dtype=float).reshape(1, -1))[0]
    i2 = np.where(p_all == np.asarray(p2, dtype=float).reshape(1, -1))[0]
    inu = np.where(nu_all == np.asarray(nu_band, dtype=float).reshape(1, -1))[0]
    if len(i1) == 0 or len(i2) == 0 or len(inu) == 0:
        raise ValueError('p1, p2 or nu have no values in p_all or nu_all')
    return (np.ascontiguousarray(i1, dtype=np.int32), np.ascontiguousarray(i2, dtype=np.int32),
            np.ascontiguousarray(inu, dtype=np.int32))


def transmission(p1, p2, p_all, nu_band, delta_nu_band, nu_all, tau):
    """Band-averaged transmission [np1 x np2] (real_gas.py:130-154)."""
    i1, i2, inu = _locate(p1, p2, p_all, nu_band, nu_all)
    tau = as_f64(tau)
    nu_band = as_f64(nu_band).reshape(-1)
    if inu.size != nu_band.size:
        raise ValueError('nu_band has %d values but %d of them are in nu_all' % (nu_band.size, inu.size))
    out = np.zeros((np.size(p1), np.size(p2)))
    if i1.size != out.shape[0] or i2.size != out.shape[1]:
        raise ValueError('every p1 / p2 value must match exactly one level of p_all')
    with np.errstate(all='ignore'):
        check(_lib.load().cmb200_transmission_host(dptr(tau), tau.shape[0], tau.shape[1], _i32p(i1), i1.size,
                                                   _i32p(i2), i2.size, _i32p(inu), dptr(nu_band), nu_band.size,
                                                   float(delta_nu_band), dptr(out)))
    return out


def dtransmission_dq(p1, p2, p_all, nu_band, delta_nu_band, nu_all, tau, absorption_coef):
    """d(transmission)/dq of one molecule [np1 x np2] (real_gas.py:157-184).  The reference's
    factor (p1 - p2) broadcasts against the *bin* axis, so -- as in all its call sites
    (centa_presentation/base.py:184) -- p1 and p2 must be single values."""
    i1, i2, inu = _locate(p1, p2, p_all, nu_band, nu_all)
    if np.size(p1) != 1 or np.size(p2) != 1:
        raise ValueError('dtransmission_dq supports single p1 and p2 values')
    tau = as_f64(tau)
    nu_band = as_f64(nu_band).reshape(-1)
    k_band = np.ascontiguousarray(as_f64(absorption_coef).reshape(-1)[inu])
    scale = as_f64(np.asarray(p1, dtype=float).reshape(-1)[:1] - np.asarray(p2, dtype=float).reshape(-1)[:1])
    out = np.zeros((1, 1))
    check(_lib.load().cmb200_dtransmission_dq_host(dptr(tau), tau.shape[0], tau.shape[1], _i32p(i1), 1, _i32p(i2), 1,
                                                   _i32p(inu), dptr(nu_band), nu_band.size, float(delta_nu_band),
                                                   dptr(k_band), dptr(scale), float(g), dptr(out)))
    return out


def emission_integrals(trans, P, p, beta_centre, beta_interface, beta_ground):
    """The pressure integrals of ``RealGas.flux_integrals`` (real_gas.py:564-627) from a band's
    transmission matrix, for all levels at once.

    trans [L x L] = Trans_j(a, b) on the interfaces (only a <= b is read: the other triangle
    overflows in the reference too); P [L] interfaces, p [L-1] cell centres, ascending;
    beta_centre [L-1], beta_interface [L], beta_ground = pi B(band centre, .) at the cell centres,
    at the interfaces and at T_g.  With D_up[i, k] = (Trans(i, k+1) - Trans(i, k)) / (P_k+1 - P_k)
    the upward integral from interface i runs over [P_i, p_i .. p_L-2, P_L-1] with the integrand
    beta_interface[i] D_up[i, i] at its first point, beta_centre[k] D_up[i, k] at the centres and
    beta_ground D_up[i, L-2] at the surface; the downward one to interface m over
    [P_0, p_0 .. p_m-1, P_m] with D_dn[k, m] = (Trans(k+1, m) - Trans(k, m)) / (P_k+1 - P_k),
    beta_interface[0] at the top and beta_interface[m] at the end.  Trapezoid rule, as np.trapz.
    Returns (integral_up [L], integral_down [L]); integral_up[L-1] = integral_down[0] = 0."""
    trans = np.asarray(trans, dtype=float)
    P = np.asarray(P, dtype=float).reshape(-1)
    p = np.asarray(p, dtype=float).reshape(-1)
    bc = np.asarray(beta_centre, dtype=float).reshape(-1)
    bi = np.asarray(beta_interface, dtype=float).reshape(-1)
    bg = float(np.ravel(beta_ground)[0])
    L = P.size
    if trans.shape != (L, L) or p.size != L - 1 or bc.size != L - 1 or bi.size != L:
        raise ValueError('emission_integrals: inconsistent shapes')
    dP, dpc = np.diff(P), np.diff(p)
    ones = np.ones((L, L), dtype=bool)
    with np.errstate(all='ignore'):
        D_up = np.where(np.triu(ones[:, :L - 1]), np.diff(trans, axis=1) / dP[None, :], 0.0)       # [i, k], k >= i
        D_dn = np.where(np.triu(ones[:L - 1, :], 1), np.diff(trans, axis=0) / dP[:, None], 0.0)    # [k, m], k < m
    rows = np.arange(L - 1)
    # upward
    G = D_up * bc[None, :]
    first = 0.5 * (bi[:-1] * D_up[rows, rows] + G[rows, rows]) * (p - P[:-1])
    inner = np.where(np.triu(ones[:, :L - 2]), 0.5 * (G[:, :-1] + G[:, 1:]) * dpc[None, :], 0.0).sum(axis=1)
    last = 0.5 * (G[:, -1] + bg * D_up[:, -1]) * (P[-1] - p[-1])
    up = np.zeros(L)
    up[:-1] = -(first + inner[:-1] + last[:-1])
    # downward
    G = D_dn * bc[:, None]
    first = 0.5 * (bi[0] * D_dn[0, :] + G[0, :]) * (p[0] - P[0])
    inner = np.where(np.triu(ones[:L - 2, :], 2), 0.5 * (G[:-1, :] + G[1:, :]) * dpc[:, None], 0.0).sum(axis=0)
    cols = np.arange(1, L)
    last = 0.5 * (G[cols - 1, cols] + bi[1:] * D_dn[cols - 1, cols]) * (P[1:] - p)
    down = np.zeros(L)
    down[1:] = first[1:] + inner[1:] + last
    return up, down


def __getattr__(name):
    # RealGas is defined in real_gas_model (which imports this module); resolve lazily so that
    # ``from climatemodel_b200.radiation.real_gas import RealGas`` works like the reference import.
    if name == 'RealGas':
        from .real_gas_model import RealGas
        return RealGas
    raise AttributeError(name)
