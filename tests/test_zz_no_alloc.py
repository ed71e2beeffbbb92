"""GPU: the allocation-free entry point cmb200_plan_optical_depth_noalloc (no counterpart in the reference)
against the existing cmb200_plan_optical_depth, through the
C ABI, on two calls in a row that share the library's fixed scratch array.

Collected last on purpose.  Optical depth is integer / separately
rounded arithmetic: the tolerance this suite uses for it is bit-exactness (test_optical_depth_bit_exact),
and the two entry points launch the same kernels, so the results must be EQUAL.
"""
import ctypes as C

import numpy as np
import pytest

from util import load_case

pytestmark = pytest.mark.gpu


def _call(lib, name, plan, n_col, T_tau, T_g, q, tau, stream):
    from climatemodel_b200._lib import check
    check(getattr(lib, name)(plan._h, n_col, T_tau.data_ptr() if T_tau is not None else None,
                             T_g.data_ptr() if T_g is not None else None, q.data_ptr(), 1 if q.dim() == 3 else 0,
                             tau.data_ptr(), C.c_void_p(stream)))


@pytest.mark.parametrize('precision', ['f64', 'f32'])
def test_no_alloc_optical_depth_equals_the_regular_entry_point(precision):
    import torch
    from climatemodel_b200 import _lib
    from climatemodel_b200.plan import FluxPlan
    c = load_case('two_mol_nz41')
    plan = FluxPlan(c['P'], c['nu'], c['bands'], tables=c['tables'], precision=precision)
    lib = _lib.load()
    dev = torch.device('cuda', 0)
    L, N, M = plan.L, plan.N, plan.M
    rng = np.random.default_rng(31)
    stream = torch.cuda.current_stream(dev).cuda_stream

    def inputs(n_col, per_column_q, isothermal):
        T_tau = None if isothermal else torch.tensor(c['T_interface_tau'][None] + rng.uniform(-40, 40, (n_col, L)), device=dev)
        T_g = torch.tensor(c['T_g'] + rng.uniform(-30, 30, n_col), device=dev)
        q = np.asarray(c['q'], dtype=float)
        if per_column_q:
            q = q[:, None, :] * rng.uniform(0.5, 2.0, (M, n_col, 1))
        return T_tau, T_g, torch.tensor(np.ascontiguousarray(q), device=dev)

    # two calls in a row on the shared scratch: many columns first, then fewer columns with other table rows
    # (a call must not see the previous call's indices), then the isothermal form (T_tau = NULL)
    for n_col, per_col, iso in ((23, True, False), (5, False, False), (9, True, True)):
        T_tau, T_g, q = inputs(n_col, per_col, iso)
        want = torch.full((n_col, L, N), -1.0, dtype=torch.float64, device=dev)
        got = torch.full((n_col, L, N), -2.0, dtype=torch.float64, device=dev)
        _call(lib, 'cmb200_plan_optical_depth', plan, n_col, T_tau, T_g, q, want, stream)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(dev)[0]
        _call(lib, 'cmb200_plan_optical_depth_noalloc', plan, n_col, T_tau, T_g, q, got, stream)
        torch.cuda.synchronize()
        if (n_col, per_col, iso) != (23, True, False):      # the first call may still load the module lazily
            assert torch.cuda.mem_get_info(dev)[0] == free0           # nothing was allocated on the device
        a, b = got.cpu().numpy(), want.cpu().numpy()
        assert np.isfinite(a).all() and a.min() >= 0.0 and a.max() > 0.0
        assert np.array_equal(a, b)
    # the Python wrapper goes through the same entry point
    T_tau, T_g, q = inputs(4, True, False)
    out = plan.optical_depth_noalloc(q, torch.empty((4, L, N), dtype=torch.float64, device=dev), T_tau=T_tau)
    assert np.array_equal(out.cpu().numpy(), plan.optical_depth_device(q, T_tau=T_tau).cpu().numpy())


def test_no_alloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    from climatemodel_b200 import _lib
    from climatemodel_b200.plan import FluxPlan
    c = load_case('two_mol_nz41')
    plan = FluxPlan(c['P'], c['nu'], c['bands'], tables=c['tables'])
    lib = _lib.load()
    dev = torch.device('cuda', 0)
    limit = (1 << 20) // (plan.M * plan.L)                      # columns the 1 MiB scratch holds for this plan
    assert limit < 65535
    q = torch.tensor(np.ascontiguousarray(c['q'], dtype=float), device=dev)
    T_g = torch.full((limit + 1,), 280.0, dtype=torch.float64, device=dev)
    dummy = torch.empty(8, dtype=torch.float64, device=dev)     # never written: the call is refused before any launch
    with pytest.raises(ValueError, match='exceeds the fixed scratch'):
        _call(lib, 'cmb200_plan_optical_depth_noalloc', plan, limit + 1, None, T_g, q, dummy, 0)
    torch.cuda.synchronize()
