"""ctypes binding of ``libclimatemodel_b200.so`` (the C ABI declared in
``include/climatemodel_b200.h``).

There is deliberately no fallback: if the shared library is missing or a call fails, an
exception is raised.  ``ValueError`` is used for argument errors so that callers see the same
exception type the reference raises (``real_gas.py:147, :176``).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libclimatemodel_b200.so')

OK, ERR_ARG, ERR_CUDA, ERR_STATE, ERR_LIMIT = 0, -1, -2, -3, -4
F64, F32 = 0, 1

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)
c_uint8_p = C.POINTER(C.c_uint8)


class Constants(C.Structure):
    _fields_ = [('g', C.c_double), ('c_p_dry', C.c_double), ('h_planck', C.c_double),
                ('speed_of_light', C.c_double), ('k_boltzmann', C.c_double)]


class Table(C.Structure):
    _fields_ = [('n_p', C.c_int32), ('n_T', C.c_int32), ('n_nu', C.c_int32),
                ('p', c_double_p), ('T', c_double_p), ('nu', c_double_p), ('k', c_double_p)]


class FluxIO(C.Structure):
    _fields_ = [('n_col', C.c_int64),
                ('T', C.c_void_p), ('T_g', C.c_void_p), ('albedo', C.c_void_p),
                ('albedo_scalar', C.c_double),
                ('tau', C.c_void_p), ('tau_id', C.c_void_p), ('n_tau', C.c_int32),
                ('T_tau', C.c_void_p), ('q', C.c_void_p), ('q_per_column', C.c_int32),
                ('up', C.c_void_p), ('down', C.c_void_p), ('net', C.c_void_p),
                ('heating', C.c_void_p), ('T_interface', C.c_void_p)]


class CudaLibraryError(RuntimeError):
    pass


_lib = None

# name -> (restype, argtypes); must list every symbol of include/climatemodel_b200.h
SIGNATURES = {
    'cmb200_last_error': (C.c_char_p, []),
    'cmb200_version': (C.c_int, []),
    'cmb200_device_count': (C.c_int, [C.POINTER(C.c_int)]),
    'cmb200_planck_wavenumber_host': (C.c_int, [C.POINTER(Constants), c_double_p, C.c_int64, c_double_p,
                                                C.c_int64, c_double_p]),
    'cmb200_absorption_coef_host': (C.c_int, [C.POINTER(Table), c_double_p, c_double_p, C.c_int64,
                                              c_double_p, C.c_int64, C.c_int, c_double_p]),
    'cmb200_optical_depth_host': (C.c_int, [C.POINTER(Constants), C.c_int32, C.POINTER(Table), c_double_p,
                                            c_double_p, C.c_int64, c_double_p, C.c_int64, c_double_p,
                                            c_double_p]),
    'cmb200_transmission_host': (C.c_int, [c_double_p, C.c_int64, C.c_int64, c_int32_p, C.c_int32, c_int32_p,
                                           C.c_int32, c_int32_p, c_double_p, C.c_int32, C.c_double,
                                           c_double_p]),
    'cmb200_dtransmission_dq_host': (C.c_int, [c_double_p, C.c_int64, C.c_int64, c_int32_p, C.c_int32,
                                               c_int32_p, C.c_int32, c_int32_p, c_double_p, C.c_int32,
                                               C.c_double, c_double_p, c_double_p, C.c_double, c_double_p]),
    'cmb200_line_absorption_host': (C.c_int, [C.POINTER(Constants), C.c_double, C.c_double, C.c_int64, c_double_p,
                                              c_double_p, C.c_int64, c_double_p, C.c_int64, c_double_p, c_double_p,
                                              c_double_p, c_double_p, C.c_double, c_double_p]),
    'cmb200_stepper_create': (C.c_int, [C.POINTER(C.c_void_p), C.c_void_p, C.c_int64, C.c_double, C.c_double]),
    'cmb200_stepper_reset': (C.c_int, [C.c_void_p]),
    'cmb200_stepper_destroy': (C.c_int, [C.c_void_p]),
    'cmb200_stepper_update': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double,
                                        C.c_double, C.c_int, C.c_void_p]),
    'cmb200_convective_adjust': (C.c_int, [C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'cmb200_stepper_set_convection': (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    'cmb200_stepper_update_convective': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                   C.c_double, C.c_double, C.c_int, C.c_double, C.c_double,
                                                   C.c_void_p]),
    'cmb200_stepper_set_frozen': (C.c_int, [C.c_void_p, C.c_void_p]),
    'cmb200_stepper_set_level_order': (C.c_int, [C.c_void_p, C.c_int]),
    'cmb200_stepper_state': (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_int32_p, c_int32_p]),
    'cmb200_grey_flux_host': (C.c_int, [C.c_int32, C.c_int64, C.c_double] + [c_double_p] * 11),
    'cmb200_plan_create': (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.POINTER(Constants), C.c_int32,
                                     c_double_p, C.c_int64, c_double_p, C.c_int32, c_int32_p, c_int32_p,
                                     c_double_p, c_double_p, c_uint8_p]),
    'cmb200_plan_destroy': (C.c_int, [C.c_void_p]),
    'cmb200_plan_set_star': (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double]),
    'cmb200_plan_set_tables': (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(Table)]),
    'cmb200_plan_set_band_range': (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    'cmb200_plan_segments': (C.c_int, [C.c_void_p, c_int32_p, C.POINTER(C.c_int64)]),
    'cmb200_plan_set_segment_range': (C.c_int, [C.c_void_p, C.c_int64, C.c_int64]),
    'cmb200_plan_info': (C.c_int, [C.c_void_p, c_int32_p, C.POINTER(C.c_int64), c_int32_p, c_int32_p,
                                   c_int32_p, c_int32_p, C.POINTER(C.c_int64)]),
    'cmb200_plan_optical_depth': (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_int, C.c_void_p, C.c_void_p]),
    'cmb200_plan_optical_depth_noalloc': (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_int, C.c_void_p, C.c_void_p]),
    'cmb200_flux': (C.c_int, [C.c_void_p, C.POINTER(FluxIO), C.c_void_p]),
    'cmb200_flux_host': (C.c_int, [C.c_void_p, C.POINTER(FluxIO)]),
    'cmb200_flux_workspace_bytes': (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    'cmb200_launch_count': (C.c_int64, []),
    'cmb200_test_exp_neg_host': (C.c_int, [c_double_p, C.c_int64, c_double_p]),
    'cmb200_test_exp_neg_table_host': (C.c_int, [c_double_p, C.c_int64, c_double_p]),
    'cmb200_measure_peaks': (C.c_int, [C.c_int, c_double_p, c_double_p, c_double_p]),
    'cmb200_measure_peaks2': (C.c_int, [C.c_int, c_double_p, c_double_p, c_double_p]),
}


def load():
    """Load the shared library (once) and attach the prototypes.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CudaLibraryError(
            '%s is missing -- build it with `python -m climatemodel_b200.csrc.build` '
            '(there is no CPU fallback)' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc == OK:
        return
    msg = load().cmb200_last_error().decode('utf-8', 'replace')
    if rc == ERR_ARG:
        raise ValueError(msg)
    raise CudaLibraryError('climatemodel_b200 error %d: %s' % (rc, msg))


def constants_struct():
    from . import constants as k
    return Constants(k.g, k.c_p_dry, k.h_planck, k.speed_of_light, k.k_boltzmann)


def as_f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def dptr(a):
    return a.ctypes.data_as(c_double_p)


def table_struct(table, keep):
    """cmb200_table view of a reference-format table dict; arrays are appended to ``keep`` so
    that they outlive the call."""
    p, T, nu, k = (as_f64(table[x]) for x in ('p', 'T', 'nu', 'absorption_coef'))
    if k.shape != (p.size, T.size, nu.size):
        raise ValueError('absorption_coef has shape %s, expected %s' % (k.shape, (p.size, T.size, nu.size)))
    keep.extend([p, T, nu, k])
    return Table(p.size, T.size, nu.size, dptr(p), dptr(T), dptr(nu), dptr(k))


def device_count():
    n = C.c_int(0)
    check(load().cmb200_device_count(C.byref(n)))
    return n.value


def launch_count():
    return int(load().cmb200_launch_count())


def measure_peaks(device=0):
    a, b, c = C.c_double(), C.c_double(), C.c_double()
    check(load().cmb200_measure_peaks(device, C.byref(a), C.byref(b), C.byref(c)))
    d, e, f = C.c_double(), C.c_double(), C.c_double()
    check(load().cmb200_measure_peaks2(device, C.byref(d), C.byref(e), C.byref(f)))
    return {'fp64_tflops': a.value, 'fp32_tflops': b.value, 'copy_gbs': c.value,
            'fp64_reg_tflops': d.value, 'fp32_reg_tflops': e.value, 'l2_stream_gbs': f.value}
