"""Grey two-stream fluxes on the GPU: the numeric core of the reference's ``GreyGas``
(``get_lw_flux`` grey.py:251-275, ``get_sw_flux`` :277-294), batched over latitude columns.  The
grid builder and the sympy optical-depth closures of the grey model are not part of this package;
callers pass the arrays ``GreyGas`` holds (``T``, ``dtau``, ``tau_sw_interface``, ``albedo_mod``,
``solar_latitude_factor``), all in its layout [level, column] with level 0 at the surface."""
import numpy as np

from .. import _lib
from ..constants import sigma


def _opt(a):
    return None if a is None else _lib.dptr(a)


def grey_fluxes(T, dtau, albedo_mod, solar_latitude_factor, F_stellar_constant, tau_sw_interface=None,
                isothermal=False):
    """Long-wave and short-wave streams and the net flux for ny columns.

    T, dtau: [nz-1, ny]; albedo_mod, solar_latitude_factor: scalars or [ny]; tau_sw_interface:
    [nz, ny] or None (no short-wave absorber).  Returns a dict of [nz, ny] arrays:
    up_lw_flux, down_lw_flux, up_sw_flux, down_sw_flux, net_flux."""
    T, dtau = _lib.as_f64(T), _lib.as_f64(dtau)
    if T.ndim != 2 or T.shape != dtau.shape:
        raise ValueError('T and dtau must both be [nz-1, ny]')
    nz, ny = T.shape[0] + 1, T.shape[1]
    albedo_mod = np.broadcast_to(_lib.as_f64(albedo_mod).reshape(-1), (ny,))
    factor = np.broadcast_to(_lib.as_f64(solar_latitude_factor).reshape(-1), (ny,))
    toa = _lib.as_f64((1 - albedo_mod) * factor * F_stellar_constant / 4)            # grey.py:265
    up0 = _lib.as_f64(np.ones(ny) * albedo_mod * factor * F_stellar_constant / 4)    # :287-288
    down0 = _lib.as_f64(np.ones(ny) * factor * F_stellar_constant / 4)               # :289-290
    tau_sw = None
    if tau_sw_interface is not None and not isothermal:
        tau_sw = _lib.as_f64(tau_sw_interface)
        if tau_sw.shape != (nz, ny):
            raise ValueError('tau_sw_interface must be [nz, ny]')
    out = {k: np.zeros((nz, ny)) for k in ('up_lw_flux', 'down_lw_flux', 'up_sw_flux', 'down_sw_flux', 'net_flux')}
    _lib.check(_lib.load().cmb200_grey_flux_host(
        nz, ny, float(sigma), _lib.dptr(T), _lib.dptr(dtau), _lib.dptr(toa), _lib.dptr(up0), _lib.dptr(down0),
        _opt(tau_sw), *[_lib.dptr(out[k]) for k in ('up_lw_flux', 'down_lw_flux', 'up_sw_flux', 'down_sw_flux',
                                                     'net_flux')]))
    return out
