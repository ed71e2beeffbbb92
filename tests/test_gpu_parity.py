"""GPU parity tests: the CUDA path, called through the C ABI (ctypes), against the golden
vectors produced by the unmodified reference and against the CPU oracle on seeded inputs.

Tolerances (BASELINE.json north_star): FP64 mode max|d|/max|ref| <= 1e-10 on fluxes, net flux
and heating rate; FP32 mode <= 1e-5.  Integer / gather work and optical depth: bit-exact.
"""
import os

import numpy as np
import pytest

import oracle.realgas_oracle as orc
from util import CASES, GOLDEN, load_case, load_table, relmax

pytestmark = pytest.mark.gpu

TOL64 = 1e-10
TOL32 = 1e-5


@pytest.fixture(scope='module')
def rg():
    import climatemodel_b200.radiation.real_gas as m
    return m


def make_plan(c, precision='f64', tables=False):
    from climatemodel_b200.plan import FluxPlan
    return FluxPlan(c['P'], c['nu'], c['bands'], tables=c['tables'] if tables else None, precision=precision,
                    T_star=c['T_star'], R_star=c['R_star'], star_planet_dist=c['dist'])


def oracle_tau(c):
    return orc.optical_depth(c['P'], c['T_interface_tau'], c['nu'], c['tables'], list(c['q']))


# ---- A1 / A2 ---------------------------------------------------------------------------------
def test_planck_and_gather_vectors(rg):
    f = np.load(os.path.join(GOLDEN, 'functions.npz'))
    B = rg.B_wavenumber(f['planck_nu'], f['planck_T'])
    assert B.shape == f['planck_B'].shape
    np.testing.assert_allclose(B, f['planck_B'], rtol=2e-13, atol=0)
    tab = load_table('CO2_no_sw')
    k = rg.get_absorption_coef(f['gather_p'], f['gather_T'], f['gather_nu'], tab)
    assert np.array_equal(k, f['gather_k'])
    from climatemodel_b200.radiation import real_gas as m
    k2 = m._gather(f['gather_p'], f['gather_T'], f['load_nu'], tab, zero_outside=True)
    assert np.array_equal(k2, f['load_k'])


def test_gather_long_axis_bisection_matches_linear_scan(rg):
    rng = np.random.default_rng(1)
    tab = {'p': np.array([1e5, 1e4, 1e3]), 'T': np.array([250., 300.]),
           'nu': np.arange(0, 3000, 2.5), 'absorption_coef': rng.random((3, 2, 1200))}
    nu = np.concatenate((rng.uniform(-10, 3010, 500), tab['nu'][:50] + 1.25, tab['nu'][:50]))
    p = np.array([2e5, 3e4, 5.5e3, 10.])
    T = np.array([100., 275., 275.1, 400.])
    k = rg.get_absorption_coef(p, T, nu, tab)
    assert np.array_equal(k, orc.gather_absorption_coef(p, T, nu, tab))


def test_layer_transmissivity_exp_accuracy():
    """The kernel's own exp(-x) (degree-11 polynomial, exponent-field scaling) against NumPy:
    <= 2 ulp over the range layer optical thicknesses take, clamp to ~0 beyond 708."""
    from climatemodel_b200 import _lib
    import ctypes as C
    rng = np.random.default_rng(0)
    x = np.concatenate((10 ** rng.uniform(-12, 2.8, 200000), rng.uniform(0, 40, 200000), [0.0, 1e-300, 700.0, 707.5]))
    out = np.empty_like(x)
    _lib.check(_lib.load().cmb200_test_exp_neg_host(_lib.dptr(x), x.size, _lib.dptr(out)))
    ref = np.exp(-x)
    assert np.max(np.abs(out - ref) / ref) < 2 * np.finfo(float).eps
    big = np.array([709.0, 1e3, 1e6, 1e300])
    _lib.check(_lib.load().cmb200_test_exp_neg_host(_lib.dptr(big), big.size, _lib.dptr(out[:4])))
    assert np.all(out[:4] < 1e-300) and np.all(out[:4] >= 0)


def test_layer_transmissivity_table_exp_accuracy():
    """The shared-memory table routine the FP64 default's short-wave passes use (2^(j/512) table, one-FMA range
    reduction, cubic): relative error <= 1e-13 up to the clamp at 708 (the one-FMA reduction loses n' 2^-63, n' =
    x 512/ln2) and <= 1.1e-14 for x <= 40 (cubic truncation r^4 / 24 <= 8.7e-15), where a transmissivity still matters (exp(-40) = 4e-18) -- four orders of
    magnitude inside the 1e-10 gate, so 99 layers of it stay below 1e-12 -- and ~0 beyond the clamp."""
    from climatemodel_b200 import _lib
    rng = np.random.default_rng(1)
    x = np.concatenate((10 ** rng.uniform(-12, 2.8, 200000), rng.uniform(0, 40, 200000), [0.0, 1e-300, 700.0, 707.5]))
    out = np.empty_like(x)
    _lib.check(_lib.load().cmb200_test_exp_neg_table_host(_lib.dptr(x), x.size, _lib.dptr(out)))
    ref = np.exp(-x)
    rel = np.abs(out - ref) / ref
    assert rel.max() < 1e-13
    assert rel[x <= 40].max() < 1.2e-14
    big = np.array([709.0, 1e3, 1e6, 1e300])
    _lib.check(_lib.load().cmb200_test_exp_neg_table_host(_lib.dptr(big), big.size, _lib.dptr(out[:4])))
    assert np.all(out[:4] < 1e-300) and np.all(out[:4] >= 0)


# ---- A3 ------------------------------------------------------------------------------------
@pytest.mark.parametrize('name', CASES)
def test_optical_depth_bit_exact(rg, name):
    c = load_case(name)
    tau = rg.optical_depth_from_tables(c['P'], c['T_interface_tau'], c['nu'], c['tables'], c['q'])
    assert np.array_equal(tau[:, c['tau_sub_idx']], c['tau_sub'])
    assert np.array_equal(tau, oracle_tau(c))


def test_optical_depth_empty_and_unsorted(rg):
    c = load_case('single_line_nz30')
    tau = rg.optical_depth_from_tables(c['P'], c['T_interface_tau'], c['nu'][:0], c['tables'], c['q'])
    assert tau.shape == (30, 0)
    with pytest.raises(ValueError):
        rg.optical_depth_from_tables(c['P'][::-1], c['T_interface_tau'], c['nu'], c['tables'], c['q'])


# ---- A4 / A5 ---------------------------------------------------------------------------------
@pytest.mark.parametrize('name', ['co2_nz50_earth', 'two_mol_nz41'])
def test_transmission_matrix(rg, name):
    c = load_case(name)
    tau = oracle_tau(c)
    j = int(c['trans_band'])
    s, n = int(c['band_start'][j]), int(c['band_len'][j])
    Tr = rg.transmission(c['P'], c['P'], c['P'], c['nu'][s:s + n], c['band_delta'][j], c['nu'], tau)
    low = np.tril_indices(c['P'].size, -1)
    ref_full = orc.transmission(c['P'], c['P'], c['P'], c['nu'][s:s + n], c['band_delta'][j], c['nu'], tau)
    # lower triangle: overflowing entries overflow on both sides
    assert np.array_equal(np.isinf(Tr[low]), np.isinf(ref_full[low]))
    Tr = np.triu(np.nan_to_num(Tr, nan=0.0, posinf=0.0))
    assert relmax(Tr, c['trans_upper']) < 1e-13


def test_transmission_subset_and_errors(rg):
    c = load_case('co2_nz50_earth')
    tau = oracle_tau(c)
    j = 3
    s, n = int(c['band_start'][j]), int(c['band_len'][j])
    a = rg.transmission(c['P'][5:9], c['P'][-1:], c['P'], c['nu'][s:s + n], c['band_delta'][j], c['nu'], tau)
    b = orc.transmission(c['P'][5:9], c['P'][-1:], c['P'], c['nu'][s:s + n], c['band_delta'][j], c['nu'], tau)
    assert a.shape == (4, 1) and relmax(a, b) < 1e-13
    with pytest.raises(ValueError):
        rg.transmission(np.array([1.2345]), c['P'], c['P'], c['nu'][:4], 30.0, c['nu'], tau)


def test_dtransmission_dq(rg):
    c = load_case('two_mol_nz41')
    tau = oracle_tau(c)
    j = int(c['dtdq_band'])
    a, b = c['dtdq_rows']
    s, n = int(c['band_start'][j]), int(c['band_len'][j])
    d = rg.dtransmission_dq(c['P'][a:a + 1], c['P'][b:b + 1], c['P'], c['nu'][s:s + n], c['band_delta'][j],
                            c['nu'], tau, c['dtdq_k'])
    assert relmax(d, c['dtdq']) < 1e-13


# ---- A6-A9: resident tau ----------------------------------------------------------------------
def heating_ok(got, ref, up, dn, delta, P, tol, eps=np.finfo(float).eps):
    """Heating rate = g/c_p d(net)/dp with net = sum_j up_j delta_j - sum_j dn_j delta_j: a
    difference of differences of nearly equal numbers, so even the reference's own FP64
    evaluation carries rounding noise of a few ulp of those band sums divided by dp.  Pass when
    the norm-wise error is below ``tol`` (north_star: 1e-10) or below that intrinsic floor,
    16 eps max(sum_j up_j delta_j) g/c_p / dp per layer.  ``eps`` is the unit
    roundoff of the arithmetic the wavenumber-resolved stage ran in: FP32 mode carries every chain in
    single precision, so its floor is the FP32 analogue (the band sums themselves are added in FP64)."""
    err = np.abs(np.asarray(got) - ref)
    mag = max(np.sum(np.abs(up) * delta, axis=1).max(), np.sum(np.abs(dn) * delta, axis=1).max())
    floor = 16 * eps * mag * orc.g / orc.c_p_dry / np.diff(P)
    return bool(np.all((err <= tol * np.abs(ref).max()) | (err <= floor)))


def check_flux(out, c, tol):
    assert relmax(out['up'][0], c['up']) < tol
    assert relmax(out['down'][0], c['down']) < tol
    assert relmax(out['net'][0], c['net']) < tol
    assert heating_ok(out['heating'][0], c['heating'], c['up'], c['down'], c['band_delta'], c['P'], tol)


@pytest.mark.parametrize('name', CASES)
def test_flux_from_tau_fp64(name):
    c = load_case(name)
    plan = make_plan(c)
    out = plan.flux(c['T'], c['T_g'], tau=oracle_tau(c), albedo=c['albedo'],
                    outputs=('up', 'down', 'net', 'heating', 'T_interface'))
    assert np.abs(out['T_interface'][0] - c['T_interface']).max() < 1e-9
    check_flux(out, c, TOL64)


@pytest.mark.parametrize('name', CASES)
def test_flux_fused_from_tables_fp64(name):
    c = load_case(name)
    plan = make_plan(c, tables=True)
    out = plan.flux(c['T'], c['T_g'], T_tau=c['T_interface_tau'], q=c['q'], albedo=c['albedo'])
    check_flux(out, c, TOL64)


@pytest.mark.parametrize('name', CASES)
def test_flux_fp32_mode(name):
    c = load_case(name)
    plan = make_plan(c, precision='f32', tables=True)
    out = plan.flux(c['T'], c['T_g'], tau=oracle_tau(c), albedo=c['albedo'])
    assert relmax(out['up'][0], c['up']) < TOL32
    assert relmax(out['down'][0], c['down']) < TOL32
    assert relmax(out['net'][0], c['net']) < TOL32
    eps32 = float(np.finfo(np.float32).eps)
    assert heating_ok(out['heating'][0], c['heating'], c['up'], c['down'], c['band_delta'], c['P'], TOL32, eps=eps32)
    out = plan.flux(c['T'], c['T_g'], T_tau=c['T_interface_tau'], q=c['q'], albedo=c['albedo'])
    assert relmax(out['up'][0], c['up']) < TOL32
    assert relmax(out['down'][0], c['down']) < TOL32
    assert relmax(out['net'][0], c['net']) < TOL32
    # heating rate (north_star: <= 1e-5 in FP32 mode): norm-wise, or below the FP32 rounding floor of the
    # difference of band sums it is made of -- and never worse than a few FP32 ulp of the net flux per layer
    assert heating_ok(out['heating'][0], c['heating'], c['up'], c['down'], c['band_delta'], c['P'], TOL32, eps=eps32)
    dnet = np.abs(np.diff(out['net'][0]) - np.diff(c['net']))
    assert dnet.max() <= 32 * eps32 * np.abs(c['net']).max()


# ---- batched columns ------------------------------------------------------------------------------
def batch_inputs(c, n_col, seed=0):
    rng = np.random.default_rng(seed)
    L = c['P'].size
    T = c['T'][None, :] + rng.normal(0, 3.0, (n_col, L - 1))
    T_g = c['T_g'] + rng.normal(0, 3.0, n_col)
    T_tau = c['T_interface_tau'][None, :] + rng.uniform(-25, 25, (n_col, L))
    albedo = rng.uniform(0.1, 0.5, n_col)
    q = c['q'][:, None, :] * rng.uniform(0.5, 2.0, (c['q'].shape[0], n_col, 1))
    return T, T_g, T_tau, albedo, np.ascontiguousarray(q)


def oracle_column(c, T, T_g, T_tau, albedo, q):
    I = orc.layer_integrand(c['P'], T_tau, c['nu'], c['tables'], list(q))
    up, dn = orc.get_flux_per_bin(c['P'], T, T_g, I, c['nu'], c['bands'], c['T_star'], c['R_star'], c['dist'], albedo)
    net = orc.net_flux(up, dn, c['band_delta'])
    return up, dn, net, orc.heating_rate(net, c['P'])


@pytest.mark.parametrize('name,n_col', [('two_mol_nz41', 37), ('co2_nz50_earth', 5)])
def test_batched_fused_columns_against_oracle(name, n_col):
    c = load_case(name)
    plan = make_plan(c, tables=True)
    T, T_g, T_tau, albedo, q = batch_inputs(c, n_col)
    out = plan.flux(T, T_g, T_tau=T_tau, q=q, albedo=albedo)
    for i in range(n_col):
        up, dn, net, heat = oracle_column(c, T[i], T_g[i], T_tau[i], albedo[i], q[:, i])
        assert relmax(out['up'][i], up) < TOL64
        assert relmax(out['down'][i], dn) < TOL64
        assert relmax(out['net'][i], net) < TOL64
        assert heating_ok(out['heating'][i], heat, up, dn, c['band_delta'], c['P'], TOL64)


def test_whole_column_mode_equals_split_mode():
    """Few columns are split over several CTAs (partial sums + finalize kernel); many columns run
    one CTA per column with the epilogue fused.  Same column, both ways: the band sums are added in
    a different order, so fluxes agree to rounding and the heating rate (a difference of net
    fluxes) to rounding relative to the fluxes."""
    c = load_case('two_mol_nz41')
    plan = make_plan(c, tables=True)
    T, T_g, T_tau, albedo, q = batch_inputs(c, 3, seed=5)
    few = plan.flux(T, T_g, T_tau=T_tau, q=q, albedo=albedo)
    n_rep = 1200                                     # > resident CTAs -> one CTA per column
    rep = lambda a: np.ascontiguousarray(np.tile(a, (n_rep,) + (1,) * (a.ndim - 1)))
    many = plan.flux(rep(T), rep(T_g), T_tau=rep(T_tau), albedo=rep(albedo),
                     q=np.ascontiguousarray(np.tile(q, (1, n_rep, 1))))
    for k in ('up', 'down', 'net', 'heating'):
        assert relmax(many[k][:3], few[k]) < (1e-10 if k == 'heating' else 1e-13)
        assert np.array_equal(many[k][:3], many[k][-3:])         # deterministic across CTAs


@pytest.mark.parametrize('name', ['two_mol_nz41', 'co2_nz50_earth'])
def test_fp32_pair_passes_match_single_segment_passes(name, monkeypatch):
    """FP32 default: a short-wave pass carries two consecutive segments per lane in packed FP32 arithmetic
    (FFMA2 / FMUL2 / FADD2, 8 warps); CMB200_FLUX_VARIANT=28675 is the same kernel with one segment per pass on 12
    warps.  Same chains, other summation order of the lanes: a few FP32 ulp of the largest flux apart, in split
    mode (few columns: parts end inside bands, odd shares) and in whole-column mode."""
    c = load_case(name)
    plan = make_plan(c, precision='f32', tables=True)
    for n_col in (2, 1300):
        T, T_g, T_tau, albedo, q = batch_inputs(c, n_col, seed=23)
        monkeypatch.setenv('CMB200_FLUX_VARIANT', '28675')
        one = plan.flux(T, T_g, T_tau=T_tau, q=q, albedo=albedo)
        monkeypatch.delenv('CMB200_FLUX_VARIANT', raising=False)
        two = plan.flux(T, T_g, T_tau=T_tau, q=q, albedo=albedo)
        for k in ('up', 'down', 'net'):
            assert np.isfinite(two[k]).all()
            assert relmax(two[k], one[k]) < 4e-6
        for i in range(min(n_col, 16)):
            assert heating_ok(two['heating'][i], one['heating'][i], two['up'][i], two['down'][i], c['bands']['delta'], c['P'],
                              1e-5, eps=np.finfo(np.float32).eps)


@pytest.mark.parametrize('precision', ['f64', 'f32'])
def test_tensor_memory_tstore_equals_shared_memory_tstore(precision, monkeypatch):
    """CMB200_FLUX_VARIANT=3 keeps the per-lane t-store in tensor memory (LDTM / STTM) and stages the
    long-wave coefficient rows in shared memory; CMB200_FLUX_VARIANT=0 selects the plain kernel
    (t-store in shared memory, coefficients broadcast with shuffles).  Same arithmetic in the same
    order: the results are equal bit for bit, in both optical-depth sources, split and whole-column
    mode.  The defaults differ from them in the table-fused path only, by rounding: the direct short-wave
    passes sum the lanes in another order (through tensor memory), FP64 adds the shared-memory exp table
    (error <= 1.1e-14 per layer) and has to stay within 1e-12 of variant 0, FP32 evaluates exp with a bare
    ex2.approx and has to stay within a few FP32 ulp (4e-6 of the largest flux; heating: the FP32 rounding
    floor of the flux differences).  On resident tau the FP32 default IS variant 3."""
    c = load_case('two_mol_nz41')
    plan = make_plan(c, precision=precision, tables=True)
    tau = oracle_tau(c)
    for n_col in (3, 1300):
        T, T_g, T_tau, albedo, q = batch_inputs(c, n_col, seed=11)
        res = {}
        for variant in ('0', '3', None):
            if variant is None:
                monkeypatch.delenv('CMB200_FLUX_VARIANT', raising=False)
            else:
                monkeypatch.setenv('CMB200_FLUX_VARIANT', variant)
            res[variant] = (plan.flux(T, T_g, T_tau=T_tau, q=q, albedo=albedo),
                            plan.flux(T, T_g, tau=tau, albedo=albedo))
        for a, b, d in zip(res['0'], res['3'], res[None]):
            for k in ('up', 'down', 'net', 'heating'):
                assert np.isfinite(b[k]).all() and np.isfinite(d[k]).all()
                assert np.array_equal(a[k], b[k])
                if precision == 'f32':
                    if a is res['0'][1]:                       # resident tau: same kernel
                        assert np.array_equal(a[k], d[k])
                    elif k != 'heating':
                        assert relmax(d[k], a[k]) < 4e-6
                    else:
                        for i in range(min(n_col, 16)):
                            assert heating_ok(d[k][i], a[k][i], d['up'][i], d['down'][i], c['bands']['delta'], c['P'], 1e-5,
                                              eps=np.finfo(np.float32).eps)
                elif k != 'heating':
                    assert relmax(d[k], a[k]) < 1e-12
                else:
                    for i in range(min(n_col, 16)):
                        assert heating_ok(d[k][i], a[k][i], d['up'][i], d['down'][i], c['bands']['delta'], c['P'], 1e-10)


def test_tau_id_selects_table():
    c = load_case('co2_nz50_earth')
    plan = make_plan(c)
    tau = oracle_tau(c)
    taus = np.stack([tau, 0.5 * tau, 2.0 * tau])
    ids = np.array([2, 0, 1, 0], dtype=np.int32)
    T = np.tile(c['T'], (4, 1))
    out = plan.flux(T, np.full(4, c['T_g']), tau=taus, tau_id=ids, albedo=c['albedo'])
    for i, t in enumerate(ids):
        one = plan.flux(c['T'], c['T_g'], tau=taus[t], albedo=c['albedo'])
        assert relmax(out['net'][i], one['net'][0]) < 1e-13
    assert relmax(out['up'][1], c['up']) < TOL64


def test_band_sharding_partial_nets_add_up():
    c = load_case('two_mol_nz41')
    full = make_plan(c, tables=True).flux(c['T'], c['T_g'], T_tau=c['T_interface_tau'], q=c['q'], albedo=c['albedo'])
    B = len(c['band_centre'])
    cuts = [0, B // 3, B // 2, B]
    net = np.zeros_like(full['net'])
    up = np.zeros_like(full['up'])
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        plan = make_plan(c, tables=True)
        plan.set_band_range(lo, hi)
        part = plan.flux(c['T'], c['T_g'], T_tau=c['T_interface_tau'], q=c['q'], albedo=c['albedo'])
        net += part['net']
        up += part['up']
        assert np.all(part['up'][:, :, :lo] == 0) and np.all(part['up'][:, :, hi:] == 0)
    assert relmax(net, full['net']) < 1e-13
    assert relmax(up, full['up']) < 1e-13


def test_fp32_pair_passes_on_shards_and_without_segment_table(monkeypatch):
    """The FP32 pair passes on partial plans: band ranges and wavenumber-segment ranges (shares that start and end
    inside bands, at odd segment offsets) add up to the whole plan's fluxes within FP32 rounding, and so does the
    bin-space table (CMB200_SEGMENT_TABLE=0: no pair passes, one segment per pass)."""
    c = load_case('two_mol_nz41')
    args = dict(T_tau=c['T_interface_tau'], q=c['q'], albedo=c['albedo'])
    full = make_plan(c, precision='f32', tables=True).flux(c['T'], c['T_g'], **args)
    n_seg = int(make_plan(c, precision='f32', tables=True).segments()[1].size)
    cuts = sorted({0, 1, n_seg // 3, n_seg // 3 + 3, n_seg // 2 + 1, n_seg})
    net = np.zeros_like(full['net'])
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        plan = make_plan(c, precision='f32', tables=True)
        plan.set_segment_range(lo, hi)
        net += plan.flux(c['T'], c['T_g'], **args)['net']
    assert relmax(net, full['net']) < 4e-6
    monkeypatch.setenv('CMB200_SEGMENT_TABLE', '0')
    binspace = make_plan(c, precision='f32', tables=True).flux(c['T'], c['T_g'], **args)
    for k in ('up', 'down', 'net'):
        assert relmax(binspace[k], full[k]) < 4e-6


# ---- size-independent properties at a larger size ----------------------------------------------------
def test_properties_larger_grid():
    """10k-bin synthetic table, 100 levels, 64 columns: (i) zero absorber => up flux is the
    surface emission at every level and the down flux the TOA value; (ii) fluxes are linear in
    the star term: doubling (1 - albedo) doubles down-flux of the short-wave bands;
    (iii) two random columns agree with the oracle at this size."""
    rng = np.random.default_rng(3)
    L, N, B = 100, 10401, 40
    P = np.sort(10 ** (np.log10(101320.) + 1 - 10 ** (np.log10(np.log10(101320.) - np.log10(20.) + 1) / (L - 1) * np.arange(L))))
    nu = np.arange(N) * 5.0
    edges = np.linspace(0, N - 1, B + 1).astype(int)
    bands = {'start': edges[:-1].astype(np.int32), 'len': (np.diff(edges) + 1).astype(np.int32),
             'centre': nu[(edges[:-1] + edges[1:]) // 2], 'delta': nu[edges[1:]] - nu[edges[:-1]],
             'sw': (nu[edges[:-1]] > 3000).astype(np.uint8)}
    tab = {'p': np.logspace(5.1, 1, 30), 'T': np.array([200., 250., 300.]), 'nu': nu,
           'absorption_coef': 10 ** rng.uniform(-6, 2, (30, 3, N))}
    from climatemodel_b200.plan import FluxPlan
    plan = FluxPlan(P, nu, bands, tables=[tab])
    n_col = 64
    T = np.full((n_col, L - 1), 260.0)
    T_g = np.full(n_col, 260.0)
    # (i)
    out = plan.flux(T, T_g, q=np.zeros((1, L)), albedo=0.3, outputs=('up', 'down'))
    assert np.abs(out['up'] - out['up'][:, -1:, :]).max() <= 1e-12 * np.abs(out['up']).max()
    assert np.abs(out['down'] - out['down'][:, :1, :]).max() <= 1e-12 * np.abs(out['down']).max()
    # (ii)
    q = np.full((1, L), 1e-3)
    a = plan.flux(T, T_g, q=q, albedo=0.4, outputs=('down',))['down']
    b = plan.flux(T, T_g, q=q, albedo=0.7, outputs=('down',))['down']
    sw = bands['sw'].astype(bool)
    assert relmax(a[:, :, sw], 2.0 * b[:, :, sw]) < 1e-13
    # (iii) two perturbed columns at this size against the oracle's per-bin form
    T2 = 260.0 + rng.normal(0, 5, (2, L - 1))
    Tg2 = np.array([262.0, 255.0])
    Ttau = 250.0 + rng.uniform(-60, 60, (2, L))
    q2 = 10 ** rng.uniform(-5, -2, (1, 2, L))
    out = plan.flux(T2, Tg2, T_tau=Ttau, q=q2, albedo=0.3)
    c = {'P': P, 'nu': nu, 'tables': [tab], 'bands': bands, 'T_star': 5778., 'R_star': 6.96340e8,
         'dist': 1.495978707e11, 'band_delta': bands['delta']}
    for i in range(2):
        up, dn, net, heat = oracle_column(c, T2[i], Tg2[i], Ttau[i], 0.3, q2[:, i])
        assert relmax(out['up'][i], up) < TOL64
        assert relmax(out['down'][i], dn) < TOL64
        assert relmax(out['net'][i], net) < TOL64
        assert heating_ok(out['heating'][i], heat, up, dn, c['band_delta'], c['P'], TOL64)


def test_errors():
    c = load_case('single_line_nz30')
    plan = make_plan(c)
    with pytest.raises(ValueError):
        plan.flux(c['T'][:-1], c['T_g'], tau=oracle_tau(c))
    with pytest.raises(Exception):
        plan.flux(c['T'], c['T_g'], q=c['q'])         # no tables on the plan
    from climatemodel_b200.plan import FluxPlan
    with pytest.raises(ValueError):
        FluxPlan(c['P'][::-1], c['nu'], c['bands'])
    with pytest.raises(ValueError):
        FluxPlan(c['P'][:4], c['nu'], c['bands'])


# ---- shapes at the edges of the kernel's decomposition -------------------------------------------------
@pytest.mark.parametrize('precision', ['f64', 'f32'])
@pytest.mark.parametrize('L,M,N,B', [(5, 1, 70, 3), (9, 2, 333, 7), (10, 3, 97, 2), (17, 4, 500, 11),
                                      (150, 2, 640, 9), (230, 1, 200, 5)])
def test_shapes_against_oracle(L, M, N, B, precision):
    """Level counts below / at / off the 8-layer block size and beyond what 8 warps' t-stores fit
    (fewer warps per CTA), 1-4 molecules, ragged bands (1-bin bands, bands that are not a
    multiple of 32 bins, a long-wave band spanning several segments), few and many columns, both
    optical-depth sources: against the oracle's evaluation of the reference algorithm.  In FP32 mode the
    same shapes exercise the 12-warp direct passes (three warps per lane quarter of tensor memory) and their
    fall-backs when the t-stores of 12 warps do not fit (L = 150, 230)."""
    tol = TOL64 if precision == 'f64' else TOL32
    eps = np.finfo(float).eps if precision == 'f64' else np.finfo(np.float32).eps
    rng = np.random.default_rng(100 * L + M)
    P = np.sort(np.logspace(np.log10(101320.), np.log10(30.), L))
    nu = 10.0 + np.arange(N) * 7.0
    cuts = np.sort(rng.choice(np.arange(1, N - 1), B - 1, replace=False))
    edges = np.concatenate(([0], cuts, [N - 1]))
    if B >= 3:
        edges[1] = edges[0] + 1                     # a band of two bins
    start, stop = edges[:-1], edges[1:]
    bands = {'start': start.astype(np.int32), 'len': (stop - start + 1).astype(np.int32),
             'centre': 0.5 * (nu[start] + nu[stop]), 'delta': nu[stop] - nu[start],
             'sw': (np.arange(B) >= max(1, (2 * B) // 3)).astype(np.uint8)}
    tabs = [{'p': np.logspace(5.1, 1, 12 + m), 'T': np.linspace(180., 320., 3 + m), 'nu': nu,
             'absorption_coef': 10 ** rng.uniform(-7, 1, (12 + m, 3 + m, N))} for m in range(M)]
    from climatemodel_b200.plan import FluxPlan
    plan = FluxPlan(P, nu, bands, tables=tabs, precision=precision)
    c = {'P': P, 'nu': nu, 'tables': tabs, 'bands': bands, 'T_star': 5778., 'R_star': 6.96340e8,
         'dist': 1.495978707e11, 'band_delta': bands['delta']}
    for n_col in (1, 3, 400):
        T = 250.0 + rng.normal(0, 8, (n_col, L - 1))
        T_g = 260.0 + rng.normal(0, 5, n_col)
        T_tau = 250.0 + rng.uniform(-80, 80, (n_col, L))
        q = 10 ** rng.uniform(-6, -2, (M, n_col, L))
        albedo = rng.uniform(0.1, 0.6, n_col)
        out = plan.flux(T, T_g, T_tau=T_tau, q=q, albedo=albedo)
        for i in sorted({0, n_col // 2, n_col - 1}):
            up, dn, net, heat = oracle_column(c, T[i], T_g[i], T_tau[i], albedo[i], q[:, i])
            assert relmax(out['up'][i], up) < tol
            assert relmax(out['down'][i], dn) < tol
            assert relmax(out['net'][i], net) < tol
            assert heating_ok(out['heating'][i], heat, up, dn, c['band_delta'], c['P'], tol, eps=eps)
        # the resident-tau source with the tau of column 0 for every column
        tau0 = orc.optical_depth(P, T_tau[0], nu, tabs, list(q[:, 0]))
        res = plan.flux(T, T_g, tau=tau0, albedo=albedo, outputs=('net',))['net']
        up, dn = orc.get_flux(P, T[-1], T_g[-1], tau0, nu, bands, albedo=albedo[-1])
        assert relmax(res[-1], orc.net_flux(up, dn, bands['delta'])) < tol
