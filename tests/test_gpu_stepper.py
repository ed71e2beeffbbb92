"""GPU: the batched device-side time step (cmb200_stepper_update) against the host-side
``Atmosphere.update_temp`` -- itself replayed against the unmodified reference in
tests/test_host_stepper.py -- fed with the same net fluxes: temperatures, times, dt, DeltaT, the
level that sets the step, frozen levels and delta_net_flux must be EQUAL, column by column."""
import numpy as np
import pytest

from util import load_case

pytestmark = pytest.mark.gpu


ORDERS = ['top_first', 'surface_first']


def host_columns(P, T0, level_order):
    """One host Atmosphere per column, its arrays in `level_order`: 'top_first' as RealGas keeps
    them (p_interface ascending, real_gas.py:281), 'surface_first' as the grey model does.
    P [L] ascending, T0 [n_col, L-1] top first (the device layout)."""
    from climatemodel_b200.radiation import base
    cols = []
    for T in T0:
        atm = base.Atmosphere(P.size, 1, 1367.0 / 4)
        atm.p_interface = to_host(P, level_order).reshape(-1, 1).copy()
        atm.p = 0.5 * (atm.p_interface[:-1] + atm.p_interface[1:])
        atm.T = to_host(T, level_order).reshape(-1, 1).copy()
        atm.net_flux = np.zeros((P.size, 1))
        cols.append(atm)
    return cols


def to_host(x, level_order):
    """Device (top first) level order -> the host model's, and back (the flip is its own inverse)."""
    return x if level_order == 'top_first' else x[::-1]


@pytest.mark.parametrize('level_order', ORDERS)
@pytest.mark.parametrize('name,n_steps,temp_change', [('co2_nz50_earth', 40, 1.0), ('co2_nz50_earth', 150, 12.0), ('two_mol_nz41', 60, 6.0)])
def test_device_steps_equal_host_steps(name, n_steps, temp_change, level_order):
    import torch
    from climatemodel_b200.plan import FluxPlan
    from climatemodel_b200.radiation.columns import ColumnStepper
    c = load_case(name)
    order = np.argsort(c['P'])
    P = c['P'][order]
    L = P.size
    plan = FluxPlan(P, c['nu'], c['bands'], tables=c['tables'])
    rng = np.random.default_rng(4)
    n_col = 6
    T_base = c['T'][np.argsort(c['p'])]
    T0 = T_base + rng.normal(0, 4, (n_col, L - 1))
    T0[1] = 250.0                                                     # an isothermal column
    T_g = c['T_g'] + rng.normal(0, 3, n_col)
    q = np.ascontiguousarray(c['q'][:, order])
    dev = torch.device('cuda', 0)
    T = torch.tensor(T0, device=dev)
    cols = ColumnStepper(plan, T, torch.tensor(T_g, device=dev), q=torch.tensor(q, device=dev), albedo=0.3,
                         temp_change=temp_change, level_order=level_order)
    host = host_columns(P, T0, level_order)
    for atm in host:
        atm.time_step_info.update(DeltaT=temp_change, MaxDeltaT=temp_change)
    t_host = np.zeros(n_col)
    seen_shrink = seen_removed = False
    for step in range(n_steps):
        dnf = cols.step().cpu().numpy()
        net = cols.net_flux.cpu().numpy()
        T_dev, t_dev, st = cols.T.cpu().numpy(), cols.t.cpu().numpy(), cols.state()
        for i, atm in enumerate(host):
            t_host[i], dnf_host = atm.update_temp(t_host[i], to_host(net[i], level_order).reshape(-1, 1).copy())
            info = atm.time_step_info
            assert np.array_equal(to_host(atm.T[:, 0], level_order), T_dev[i]), (step, i)
            assert t_host[i] == t_dev[i] and info['dt'] == st['dt'][i], (step, i)
            assert info['DeltaT'] == st['DeltaT'][i] and info['MaxTendInd'] == st['MaxTendInd'][i], (step, i)
            assert len(info['RemoveInd']) == st['n_removed'][i], (step, i)
            assert dnf_host == dnf[i], (step, i)
            seen_shrink |= info['DeltaT'] < temp_change
            seen_removed |= len(info['RemoveInd']) > 0
    if temp_change > 1:
        assert seen_shrink                     # the oscillation branch was exercised
    assert np.all(t_dev > 0) and len(set(np.round(t_dev, 3))) > 1     # every column on its own clock


def test_changing_tau_and_reset():
    import torch
    from climatemodel_b200.plan import FluxPlan
    from climatemodel_b200.radiation.columns import ColumnStepper
    c = load_case('co2_nz50_earth')
    order = np.argsort(c['P'])
    P = c['P'][order]
    plan = FluxPlan(P, c['nu'], c['bands'], tables=c['tables'])
    dev = torch.device('cuda', 0)
    T0 = np.tile(c['T'][np.argsort(c['p'])], (3, 1))
    T = torch.tensor(T0, device=dev)
    cols = ColumnStepper(plan, T, torch.full((3,), c['T_g'], device=dev, dtype=torch.float64),
                         q=torch.tensor(np.ascontiguousarray(c['q'][:, order]), device=dev))
    assert np.all(cols.step().cpu().numpy() == 1e6)                   # first step: no previous net flux
    assert np.all(cols.step().cpu().numpy() < 1e6)
    assert np.all(cols.step(changing_tau=True).cpu().numpy() == 1e6)
    t3 = cols.t.cpu().numpy().copy()
    assert np.all(t3 > 0) and np.all(t3 == t3[0])                     # identical columns stay identical
    cols.reset()
    assert np.all(cols.t.cpu().numpy() == 0) and np.all(cols.state()['MaxTendInd'] == -1)


# ---- dry convective adjustment on the device -------------------------------------
def test_convective_adjustment_equals_reference_vectors():
    """The 24 profiles of tests/golden/host_convective.npz (unmodified reference, both pressure
    orders): the device kernel must return the same bits."""
    import os
    import torch
    from util import GOLDEN
    from climatemodel_b200.radiation.columns import convective_adjustment_device
    c = np.load(os.path.join(GOLDEN, 'host_convective.npz'))
    dev = torch.device('cuda', 0)
    changed = 0
    for order in (True, False):                       # group the profiles by pressure order, batch each group
        idx = [i for i in range(c['p'].shape[0]) if (c['p'][i][0] > c['p'][i][1]) == order]
        T = torch.tensor(np.ascontiguousarray(c['T'][idx]), device=dev)
        out, status = convective_adjustment_device(c['p'][idx[0]], T)
        assert np.all(status.cpu().numpy() == 0)
        assert np.array_equal(out.cpu().numpy(), c['T_adjusted'][idx])
        changed += int((np.abs(c['T_adjusted'][idx] - c['T'][idx]).max(axis=1) > 1e-3).sum())
    assert changed >= 6


def test_convective_adjustment_random_profiles_equal_host():
    import torch
    from climatemodel_b200.radiation.columns import convective_adjustment_device
    from climatemodel_b200.radiation.convective_adjustment import convective_adjustment
    rng = np.random.default_rng(12)
    for nz in (9, 40, 99, 150):
        p = np.logspace(np.log10(101320.0), np.log10(30.0), nz)
        T = 215 + 75 * (p / p[0]) ** rng.uniform(0.1, 0.7, (64, 1)) + rng.normal(0, rng.uniform(0, 5, (64, 1)), (64, nz))
        lid = max(3, nz // 8)
        T[:, -lid:] = T[:, -lid - 1:-lid] + np.linspace(5.0, 60.0, lid)   # a stable lid (the reference indexes above unstable runs)
        want = convective_adjustment(p, T.T.copy()).T
        got, status = convective_adjustment_device(p, torch.tensor(T, device=torch.device('cuda', 0)))
        assert np.all(status.cpu().numpy() == 0)
        assert np.array_equal(got.cpu().numpy(), want), nz
        if nz >= 40:
            assert (np.abs(want - T).max(axis=1) > 1e-3).sum() >= 5


@pytest.mark.parametrize('level_order', ORDERS)
def test_device_steps_with_convection_equal_host_steps(level_order):
    import torch
    from climatemodel_b200.plan import FluxPlan
    from climatemodel_b200.radiation.columns import ColumnStepper
    c = load_case('co2_nz50_earth')
    order = np.argsort(c['P'])
    P = c['P'][order]
    L = P.size
    plan = FluxPlan(P, c['nu'], c['bands'], tables=c['tables'])
    rng = np.random.default_rng(9)
    n_col = 5
    T0 = c['T'][np.argsort(c['p'])] + rng.normal(0, 2, (n_col, L - 1))
    T0[:, -12:] += np.linspace(0, 14, 12)                      # warm lower troposphere: super-adiabatic
    T_g = c['T_g'] + rng.normal(0, 3, n_col)
    dev = torch.device('cuda', 0)
    cols = ColumnStepper(plan, torch.tensor(T0, device=dev), torch.tensor(T_g, device=dev),
                         q=torch.tensor(np.ascontiguousarray(c['q'][:, order]), device=dev), albedo=0.3,
                         level_order=level_order)
    host = host_columns(P, T0, level_order)
    t_host = np.zeros(n_col)
    moved = 0
    for step in range(40):
        T_before = cols.T.cpu().numpy().copy()
        dnf = cols.step(convective_adjust=True, conv_thresh=1e-9, conv_t_multiplier=3).cpu().numpy()
        net = cols.net_flux.cpu().numpy()
        T_dev, t_dev, st = cols.T.cpu().numpy(), cols.t.cpu().numpy(), cols.state()
        for i, atm in enumerate(host):
            t_host[i], dnf_host = atm.update_temp(t_host[i], to_host(net[i], level_order).reshape(-1, 1).copy(),
                                                   convective_adjust=True, conv_thresh=1e-9, conv_t_multiplier=3)
            assert np.array_equal(to_host(atm.T[:, 0], level_order), T_dev[i]), (step, i)
            assert t_host[i] == t_dev[i] and atm.time_step_info['dt'] == st['dt'][i], (step, i)
            assert dnf_host == dnf[i], (step, i)
            moved += len(atm.time_step_info['convective_levels']) > 0
    assert moved >= 5                                           # every column convected at least once


@pytest.mark.parametrize('use_graph', [False, True])
@pytest.mark.parametrize('level_order', ORDERS)
@pytest.mark.parametrize('convective', [False, True])
def test_batched_evolve_to_equilibrium_equals_per_column_host_loop(convective, level_order, use_graph):
    """ColumnStepper.evolve_to_equilibrium against Atmosphere.evolve_to_equilibrium's loop
    (base.py:266-335) run column by column on the host with the same net fluxes: every column stops
    at its own step -- by the flux test, by the tightened threshold of the second step or by t_end --
    and is frozen there while the others go on; final temperatures, times and step counts are EQUAL."""
    import torch
    from climatemodel_b200.plan import FluxPlan
    from climatemodel_b200.radiation.base import t_years_days
    from climatemodel_b200.radiation.columns import ColumnStepper
    c = load_case('co2_nz50_earth')
    order = np.argsort(c['P'])
    P = c['P'][order]
    L = P.size
    plan = FluxPlan(P, c['nu'], c['bands'], tables=c['tables'])
    rng = np.random.default_rng(21)
    n_col = 6
    T0 = c['T'][np.argsort(c['p'])] + rng.normal(0, 1.5, (n_col, L - 1)) * np.linspace(0.2, 3.0, n_col)[:, None]
    if convective:
        T0[:, -10:] += np.linspace(0, 10, 10)
    T_g = c['T_g'] + rng.normal(0, 3, n_col)
    q = np.ascontiguousarray(c['q'][:, order])
    kw = dict(flux_thresh=0.5, t_end=0.02, convective_adjust=convective, conv_thresh=1e-9, conv_t_multiplier=3)
    dev = torch.device('cuda', 0)
    cols = ColumnStepper(plan, torch.tensor(T0, device=dev), torch.tensor(T_g, device=dev), q=torch.tensor(q, device=dev),
                         albedo=0.3, temp_change=4.0, level_order=level_order)
    n_steps = cols.evolve_to_equilibrium(check_every=5, max_steps=600, use_graph=use_graph, **kw).cpu().numpy()

    # host: one Atmosphere per column; the net fluxes come from the same plan and the same batch shape
    host = host_columns(P, T0, level_order)
    for atm in host:
        atm.time_step_info.update(DeltaT=4.0, MaxDeltaT=4.0)
    T_host, t_host = T0.copy(), np.zeros(n_col)
    thr, done, steps = np.full(n_col, kw['flux_thresh']), np.zeros(n_col, bool), np.zeros(n_col, int)
    i = 0
    while not done.all():
        assert i < 600
        net = plan.flux(T_host, T_g, q=q, albedo=0.3, outputs=('net',))['net']
        for k, atm in enumerate(host):
            if done[k]:
                continue
            t_host[k], dnf = atm.update_temp(t_host[k], to_host(net[k], level_order).reshape(-1, 1).copy(),
                                             convective_adjust=convective, conv_thresh=kw['conv_thresh'],
                                             conv_t_multiplier=kw['conv_t_multiplier'])
            T_host[k] = to_host(atm.T[:, 0], level_order)
            steps[k] += 1
            if i == 1:
                thr[k] = min(thr[k], 0.99 * dnf)
            years, days = t_years_days(t_host[k])
            done[k] = atm.check_equilibrium(dnf, thr[k]) or years + days / 365 > kw['t_end']
        i += 1
    assert np.array_equal(n_steps, steps)
    assert np.array_equal(cols.T.cpu().numpy(), T_host)
    assert np.array_equal(cols.t.cpu().numpy(), t_host)
    assert len(set(steps)) > 1 and steps.max() > 5            # the columns stopped at different steps


# ---- the reference's own RealGas step histories through the device stepper --------------------------------
@pytest.mark.parametrize('name', ['co2_nz50_earth', 'co2_nz50_earth_conv', 'co2_nz100_iso', 'two_mol_nz41'])
def test_device_stepper_replays_reference_realgas_steps(name):
    """tests/golden/make_golden.py recorded T, t and delta_net_flux of the UNMODIFIED reference's
    ``RealGas.take_time_step`` (real_gas.py:671-718 -> base.py:130-195) for 4-6 steps, with and
    without convective adjustment.  One column through ``ColumnStepper`` (flux kernel + device
    time step + device convective adjustment, RealGas level numbering) must follow them to the
    same tolerances the host ``RealGas`` is held to in test_gpu_realgas.py."""
    import torch
    from climatemodel_b200.plan import FluxPlan
    from climatemodel_b200.radiation.columns import ColumnStepper
    c = load_case(name)
    assert np.all(np.diff(c['P']) > 0)                       # RealGas order: index 0 = top of atmosphere
    plan = FluxPlan(c['P'], c['nu'], c['bands'], tables=c['tables'], T_star=c['T_star'], R_star=c['R_star'],
                    star_planet_dist=c['dist'])
    dev = torch.device('cuda', 0)
    conv = bool(c['step_conv'])
    n_col = 3                                                 # identical columns must stay identical
    T = torch.tensor(np.tile(c['T'], (n_col, 1)), device=dev)
    cols = ColumnStepper(plan, T, torch.full((n_col,), c['T_g'], dtype=torch.float64, device=dev),
                         q=torch.tensor(np.ascontiguousarray(c['q']), device=dev),
                         T_tau=torch.tensor(np.tile(c['T_interface_tau'], (n_col, 1)), device=dev), albedo=c['albedo'])
    for i in range(c['step_T'].shape[0]):
        dnf = cols.step(convective_adjust=conv).cpu().numpy()
        t, T_now = cols.t.cpu().numpy(), cols.T.cpu().numpy()
        assert np.all(t == t[0]) and np.all(T_now == T_now[0])
        assert abs(t[0] - c['step_t'][i]) <= 1e-7 * c['step_t'][i], i
        assert np.abs(T_now[0] - c['step_T'][i]).max() < 1e-6, i
        if i > 0:
            assert abs(dnf[0] - c['step_dnf'][i]) <= 1e-6 * max(abs(c['step_dnf'][i]), 1e-3), i
        else:
            assert dnf[0] == 1e6
