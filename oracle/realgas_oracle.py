"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the real-gas radiative-transfer hot path.

NumPy restatement of the algorithm in the reference's ``Model/radiation/real_gas.py`` and
``Model/radiation/base.py`` (jduffield65/ClimateModel).  It exists to *check* the CUDA path:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it.  The product package ``climatemodel_b200`` never does.

Parity status: **pinned** -- ``tests/test_oracle_golden.py`` checks every function here against
golden vectors produced by running the *unmodified* reference in the build container
(``tests/golden/make_golden.py`` via ``oracle/reference_shim.py``).  Third-party arithmetic the
reference leans on and that is not part of its tree: ``sympl`` constants (unpinned; see
``SYMPL`` below) and SciPy FITPACK ``InterpolatedUnivariateSpline`` (``real_gas.py:634``;
re-used here as the same call, and cross-checked against an independent not-a-knot solve).

Two forms of the flux solver are given:

* ``get_flux``            -- the reference's own discretisation (band-averaged transmission
                             matrix -> finite differences in p -> trapezoid in p), following
                             ``real_gas.py:564-669`` step by step.
* ``get_flux_per_bin``    -- the algebraically identical per-wavenumber-bin O(L) recurrence that
                             the CUDA kernels implement ("per-bin form").
                             Checked against ``get_flux`` and against the golden vectors.

Index convention (``real_gas.py:281``): interface 0 = top of atmosphere, L-1 = surface.
"""
import numpy as np
from scipy.interpolate import InterpolatedUnivariateSpline

# --- constants: Model/constants.py:3-19 (sympl 0.4.x values; same table as reference_shim) ----
SYMPL = {
    "g": 9.80665, "c_p_dry": 1004.64, "sigma": 5.670367e-8, "p_surface_earth": 1.0132e5,
    "speed_of_light": 299792458.0, "h_planck": 6.62607004e-34, "k_boltzmann": 1.38064852e-23,
    "R_specific": 287.0,
}
g = SYMPL["g"]
c_p_dry = SYMPL["c_p_dry"]
speed_of_light = SYMPL["speed_of_light"]
h_planck = SYMPL["h_planck"]
k_boltzmann = SYMPL["k_boltzmann"]
p_surface_earth = SYMPL["p_surface_earth"]
R_specific = SYMPL["R_specific"]
AU = 1.495978707e11
R_sun = 6.96340e8
T_sun = 5778

_trapz = getattr(np, "trapezoid", None) or np.trapz


# ------------------------------------------------------------------------------------------
# A1  Planck function  (real_gas.py:14-42)
# ------------------------------------------------------------------------------------------
def planck_frequency(freq, T):
    """B(f, T) per unit frequency, no overflow guard (real_gas.py:26-27)."""
    u = h_planck * freq / (k_boltzmann * T)
    return (2. * h_planck * freq ** 3 / speed_of_light ** 2) / (np.exp(u) - 1.)


def planck_wavenumber(nu, T):
    """B(nu, T) per cm^-1 (real_gas.py:40-42)."""
    return (100 * speed_of_light) * planck_frequency(100 * nu * speed_of_light, T)


# ------------------------------------------------------------------------------------------
# A2  nearest-neighbour gather from the k[p, T, nu] table  (real_gas.py:45-83)
# ------------------------------------------------------------------------------------------
def nearest_index(values, axis_values):
    """argmin |v - axis| with NumPy's first-index tie-break (real_gas.py:57-60)."""
    values = np.asarray(values, dtype=float).reshape(-1, 1)
    axis_values = np.asarray(axis_values).reshape(1, -1)
    return np.abs(values - axis_values).argmin(axis=1)


def gather_absorption_coef(p, T, nu, table):
    """k[level, bin] for levels (p, T) and bins nu inside the table range (real_gas.py:45-62)."""
    ip = nearest_index(p, table['p'])
    it = nearest_index(T, table['T'])
    inu = nearest_index(nu, table['nu'])
    return table['absorption_coef'][ip, it][:, inu]


def absorption_coef_on_grid(p, T, nu, table):
    """As ``load_absorption_coef`` (real_gas.py:65-83) but the table dict is passed in: zero
    outside [table nu.min(), table nu.max()], nearest-neighbour inside."""
    p = np.asarray(p, dtype=float)
    nu = np.asarray(nu, dtype=float)
    out = np.zeros((p.size, nu.size))
    inside = np.where((nu >= table['nu'].min()) & (nu <= table['nu'].max()))[0]
    out[:, inside] = gather_absorption_coef(p, T, nu[inside], table)
    return out


# ------------------------------------------------------------------------------------------
# A3  optical depth  (real_gas.py:86-127)
# ------------------------------------------------------------------------------------------
def optical_depth(p, T, nu, tables, q_values):
    """tau[level, bin]; ``tables`` is a list of table dicts, ``q_values`` the matching list of
    q(p) arrays evaluated on the *sorted* p (real_gas.py:106, :118).

    tau_i = trapz(integrand[:i+2], [0, p_0..p_i]) (real_gas.py:123-126), evaluated as the
    running sum it is algebraically equal to -- each term d*(y1+y0)/2 exactly as np.trapz forms it.
    """
    p = np.sort(np.asarray(p, dtype=float))
    integrand = np.zeros((p.size, np.size(nu)))
    for table, q in zip(tables, q_values):
        integrand += absorption_coef_on_grid(p, T, nu, table) * np.asarray(q).reshape(-1, 1)
    integrand = integrand / g
    y = np.concatenate((np.zeros((1, integrand.shape[1])), integrand))
    x = np.concatenate(([0.0], p))
    terms = np.diff(x).reshape(-1, 1) * (y[1:] + y[:-1]) / 2.0
    tau = np.empty_like(integrand)
    running = np.zeros(integrand.shape[1])
    for i in range(p.size):
        running = running + terms[i]
        tau[i] = running
    return tau


def optical_depth_reference_loop(p, T, nu, tables, q_values):
    """Literal O(L^2 N) form of real_gas.py:125-126 (used once in the tests to show the running
    sum above is bit-identical to what the reference loop produces)."""
    p = np.sort(np.asarray(p, dtype=float))
    integrand = np.zeros((p.size, np.size(nu)))
    for table, q in zip(tables, q_values):
        integrand += absorption_coef_on_grid(p, T, nu, table) * np.asarray(q).reshape(-1, 1)
    integrand = integrand / g
    y = np.concatenate((np.zeros((1, integrand.shape[1])), integrand))
    x = np.concatenate(([0.0], p))
    tau = np.zeros_like(integrand)
    for i in range(p.size):
        tau[i, :] = _trapz(y[:i + 2], x[:i + 2], axis=0)
    return tau


# ------------------------------------------------------------------------------------------
# A4 / A5  band-averaged transmission and its q-derivative  (real_gas.py:130-184)
# ------------------------------------------------------------------------------------------
def _locate(all_values, wanted, what):
    """Row/bin lookup by exact float equality, ValueError when nothing matches (real_gas.py:143-147)."""
    idx = np.where(np.asarray(all_values).reshape(-1, 1) == np.asarray(wanted).reshape(1, -1))[0]
    if idx.size == 0:
        raise ValueError('p1, p2 or nu have no values in p_all or nu_all')
    return idx


def transmission(p1, p2, p_all, nu_band, delta_nu_band, nu_all, tau):
    """Trans[a, b] = trapz_nu exp(tau[a] - tau[b]) / delta (real_gas.py:130-154).  Lower-triangle
    entries overflow exactly as they do in the reference."""
    ia = _locate(p_all, p1, 'p1')
    ib = _locate(p_all, p2, 'p2')
    inu = _locate(nu_all, nu_band, 'nu')
    t = tau[:, inu]
    with np.errstate(over='ignore', invalid='ignore'):
        e = np.exp(t[ia][:, None, :] - t[ib][None, :, :])
        return _trapz(e, nu_band, axis=2) / delta_nu_band


def dtransmission_dq(p1, p2, p_all, nu_band, delta_nu_band, nu_all, tau, absorption_coef):
    """d Trans / d q for one molecule (real_gas.py:157-184).  NB the factor (p1 - p2) broadcasts
    against the trailing (bin) axis exactly as written in the reference, so p1/p2 must be
    scalars or length-|band| compatible there; callers use scalars."""
    ia = _locate(p_all, p1, 'p1')
    ib = _locate(p_all, p2, 'p2')
    inu = _locate(nu_all, nu_band, 'nu')
    t = tau[:, inu]
    k = np.asarray(absorption_coef)[inu]
    with np.errstate(over='ignore', invalid='ignore'):
        e = (p1 - p2) * k * np.exp(t[ia][:, None, :] - t[ib][None, :, :]) / g
        return _trapz(e, nu_band, axis=2) / delta_nu_band


# ------------------------------------------------------------------------------------------
# A7 (first line)  interface temperatures  (real_gas.py:634-636)
# ------------------------------------------------------------------------------------------
def interface_temperature(p_centre, T_centre, p_interface, T_g):
    """Interpolating cubic (FITPACK, not-a-knot) spline through cell centres, evaluated --
    and at the top extrapolated -- at the interfaces; surface value replaced by T_g."""
    spline = InterpolatedUnivariateSpline(np.asarray(p_centre).reshape(-1), np.asarray(T_centre).reshape(-1))
    Ti = spline(np.asarray(p_interface).reshape(-1))
    Ti[-1] = T_g
    return Ti


def not_a_knot_matrix(x, xq):
    """Independent construction of the same spline as a linear operator: returns A with
    spline(xq) = A @ y for data y on knots x (cubic, not-a-knot end conditions, natural
    extrapolation of the end polynomials).  Used to cross-check FITPACK and, in the product,
    to build the device operator."""
    x = np.asarray(x, dtype=float)
    n = x.size
    h = np.diff(x)
    # unknowns: second derivatives m_0..m_{n-1};  M m = R y
    M = np.zeros((n, n))
    R = np.zeros((n, n))
    for i in range(1, n - 1):
        M[i, i - 1] = h[i - 1]
        M[i, i] = 2 * (h[i - 1] + h[i])
        M[i, i + 1] = h[i]
        R[i, i - 1] = 6 / h[i - 1]
        R[i, i] = -6 / h[i - 1] - 6 / h[i]
        R[i, i + 1] = 6 / h[i]
    # not-a-knot: third derivative continuous at x_1 and x_{n-2}
    M[0, 0], M[0, 1], M[0, 2] = h[1], -(h[0] + h[1]), h[0]
    M[-1, -3], M[-1, -2], M[-1, -1] = h[-1], -(h[-2] + h[-1]), h[-2]
    S = np.linalg.solve(M, R)            # m = S y
    xq = np.asarray(xq, dtype=float)
    A = np.zeros((xq.size, n))
    seg = np.clip(np.searchsorted(x, xq, side='right') - 1, 0, n - 2)
    for r, (s, q) in enumerate(zip(seg, xq)):
        hs = h[s]
        a = (x[s + 1] - q) / hs
        b = (q - x[s]) / hs
        A[r, s] += a
        A[r, s + 1] += b
        A[r] += ((a ** 3 - a) * S[s] + (b ** 3 - b) * S[s + 1]) * hs * hs / 6.0
    return A


# ------------------------------------------------------------------------------------------
# A6 / A7  fluxes, reference discretisation  (real_gas.py:564-669)
# ------------------------------------------------------------------------------------------
def band_transmission_matrix(tau, start, length, nu, delta):
    """Upper-triangular Trans[a, b] (a <= b) for one band; entries with a > b are left at 0
    (the reference computes them -- they overflow -- and never reads them, real_gas.py:605,617)."""
    L = tau.shape[0]
    t = tau[:, start:start + length]
    x = nu[start:start + length]
    out = np.zeros((L, L))
    for a in range(L):
        e = np.exp(t[a][None, :] - t[a:])            # rows b = a..L-1
        out[a, a:] = _trapz(e, x, axis=1) / delta
    return out


def flux_integrals(Trans, P, p, beta_centre, beta_interface, beta_ground):
    """Emission integrals of one long-wave band (real_gas.py:564-627).

    Trans: [L, L] upper triangle; P: interfaces [L]; p: centres [L-1];
    beta_*: pi*B(band centre, T) on centres [L-1], interfaces [L] and at T_g."""
    L = P.size
    dP = np.diff(P)
    D_up = np.zeros((L, L - 1))
    D_dn = np.zeros((L - 1, L))
    for a in range(L):
        D_up[a, a:] = np.diff(Trans[a, a:]) / dP[a:]                # real_gas.py:584
    for b in range(L):
        D_dn[:b, b] = np.diff(Trans[:b + 1, b]) / dP[:b]            # real_gas.py:595
    up = np.zeros(L)
    dn = np.zeros(L)
    for i in range(L - 1):
        # upward: x = [P_i, p_i..p_{L-2}, P_{L-1}]   (real_gas.py:605-613)
        x = np.concatenate((P[i:i + 1], p[i:], P[-1:]))
        f = np.concatenate(([beta_interface[i] * D_up[i, i]],
                            beta_centre[i:] * D_up[i, i:],
                            [beta_ground * D_up[i, L - 2]]))
        up[i] = -_trapz(f, x)
        # downward at interface m: x = [P_0, p_0..p_{m-1}, P_m]   (real_gas.py:616-626)
        m = L - 1 - i
        x = np.concatenate((P[:1], p[:m], P[m:m + 1]))
        f = np.concatenate(([beta_interface[0] * D_dn[0, m]],
                            beta_centre[:m] * D_dn[:m, m],
                            [beta_interface[m] * D_dn[m - 1, m]]))
        dn[m] = _trapz(f, x)
    return up, dn


def get_flux(P, T, T_g, tau, nu, bands, T_star=T_sun, R_star=R_sun, dist=AU, albedo=0.3,
             T_interface=None):
    """up, down [L, B] as ``RealGas.get_flux`` (real_gas.py:629-669).

    bands: dict with integer arrays 'start', 'len' (bin runs in nu), float arrays 'centre',
    'delta' and bool array 'sw'."""
    P = np.asarray(P, dtype=float).reshape(-1)
    T = np.asarray(T, dtype=float).reshape(-1)
    L = P.size
    p = 0.5 * (P[:-1] + P[1:])          # real_gas.py:282-284 (np.mean of two == (a+b)/2)
    if T_interface is None:
        T_interface = interface_temperature(p, T, P, T_g)
    B = len(bands['centre'])
    solar = R_star ** 2 / dist ** 2 * (1 - albedo) / 4
    up = np.zeros((L, B))
    dn = np.zeros((L, B))
    for j in range(B):
        c = bands['centre'][j]
        s0, n0, dj = int(bands['start'][j]), int(bands['len'][j]), bands['delta'][j]
        if bands['sw'][j]:
            # short-wave band: the reference evaluates only Trans(i, L-1) and Trans(0, i) (:647-655),
            # two length-L vectors, not the L x L matrix
            t, x = tau[:, s0:s0 + n0], nu[s0:s0 + n0]
            up[:, j] = np.pi * planck_wavenumber(c, T_g) * _trapz(np.exp(t - t[L - 1]), x, axis=1) / dj
            dn[:, j] = np.pi * planck_wavenumber(c, T_star) * solar * _trapz(np.exp(t[0] - t), x, axis=1) / dj
            continue
        Tr = band_transmission_matrix(tau, s0, n0, nu, dj)
        up[:, j] = np.pi * planck_wavenumber(c, T_g) * Tr[:, L - 1]                 # :638, :647
        dn[:, j] = np.pi * planck_wavenumber(c, T_star) * solar * Tr[0, :]          # :640-642, :652
        if not bands['sw'][j]:
            with np.errstate(over='ignore', invalid='ignore', divide='ignore'):
                bc = np.pi * planck_wavenumber(c, T)
                bi = np.pi * planck_wavenumber(c, T_interface)
                bg = np.pi * planck_wavenumber(c, T_g)
            iu, idn = flux_integrals(Tr, P, p, bc, bi, bg)
            up[:, j] += iu
            dn[:, j] += idn
    return up, dn


# ------------------------------------------------------------------------------------------
# A8 / A9  net flux and heating rate  (real_gas.py:712-713, base.py:166-168)
# ------------------------------------------------------------------------------------------
def net_flux(up, dn, delta):
    return np.sum(up * delta, axis=1) - np.sum(dn * delta, axis=1)


def heating_rate(net, P):
    P = np.asarray(P, dtype=float).reshape(-1)
    return g / c_p_dry * (net[1:] - net[:-1]) / (P[1:] - P[:-1])


# ------------------------------------------------------------------------------------------
# Per-bin O(L) recurrence -- the form the CUDA kernels use
# ------------------------------------------------------------------------------------------
def emission_coefficients(P, beta_centre, beta_interface, beta_ground):
    """Pressure-quadrature coefficients of one (column, long-wave band), already divided by the
    layer thickness.  Returns (dn_last, dn_gen, up_first, up_gen), each [L-1]:

      down, interface m=k+1:  I_dn(m) = S + dn_last[k]*(1-t_k);  carry H = S + dn_gen[k]*(1-t_k)
      up,   interface i:      I_up(i) = t_i*S + up_first[i]*(1-t_i); carry S = t_i*S + up_gen[i]*(1-t_i)

    Derived from the trapezoid weights of real_gas.py:603-626."""
    L = P.size
    p = 0.5 * (P[:-1] + P[1:])
    dP = np.diff(P)
    bc, bi, bg = beta_centre, beta_interface.copy(), beta_ground
    bi[L - 1] = bg
    n = L - 1
    gen = np.zeros(n)                     # generic interior weight 0.5*beta_k*(p_{k+1}-p_{k-1})
    gen[1:n - 1] = 0.5 * bc[1:n - 1] * (p[2:] - p[:n - 2])
    dn_gen = gen.copy()
    up_gen = gen.copy()
    dn_last = np.zeros(n)
    up_first = np.zeros(n)
    if n == 1:
        both = 0.5 * bi[0] * (p[0] - P[0]) + 0.5 * bc[0] * (P[1] - P[0]) + 0.5 * bg * (P[1] - p[0])
        dn_last[0] = both
        up_first[0] = both
    else:
        dn_gen[0] = 0.5 * bi[0] * (p[0] - P[0]) + 0.5 * bc[0] * (p[1] - P[0])
        dn_gen[n - 1] = 0.0               # never used
        dn_last[0] = 0.5 * bi[0] * (p[0] - P[0]) + 0.5 * bc[0] * (P[1] - P[0]) + 0.5 * bi[1] * (P[1] - p[0])
        k = np.arange(1, n)
        dn_last[1:] = 0.5 * bc[k] * (P[k + 1] - p[k - 1]) + 0.5 * bi[k + 1] * (P[k + 1] - p[k])
        i = np.arange(0, n - 1)
        up_first[:n - 1] = 0.5 * bi[i] * (p[i] - P[i]) + 0.5 * bc[i] * (p[i + 1] - P[i])
        up_first[n - 1] = (0.5 * bi[n - 1] * (p[n - 1] - P[n - 1]) + 0.5 * bc[n - 1] * (P[n] - P[n - 1])
                           + 0.5 * bg * (P[n] - p[n - 1]))
        up_gen[0] = 0.0                   # never used
        up_gen[n - 1] = 0.5 * bc[n - 1] * (P[n] - p[n - 2]) + 0.5 * bg * (P[n] - p[n - 1])
    return dn_last / dP, dn_gen / dP, up_first / dP, up_gen / dP


def band_bin_weights(nu, start, length, delta):
    """Trapezoid weights over the bins of a band divided by the band width (real_gas.py:154)."""
    x = np.asarray(nu[start:start + length], dtype=float)
    w = np.zeros(length)
    d = np.diff(x)
    w[:-1] += 0.5 * d
    w[1:] += 0.5 * d
    return w / delta


def get_flux_per_bin(P, T, T_g, integrand, nu, bands, T_star=T_sun, R_star=R_sun, dist=AU,
                     albedo=0.3, T_interface=None, segment_range=None, bins_per_segment=32):
    """Same up, down [L, B] as ``get_flux`` but from the layer integrand I[L, N]
    (= sum_mol k q / g on the interfaces) by one exp per (layer, bin) and two linear
    recurrences over the layers; band sums with the trapezoid nu-weights.

    ``segment_range`` = (lo, hi) keeps only the chains of segments lo..hi-1 (runs of
    ``bins_per_segment`` bins inside a band, numbered band after band) -- the partial sum one rank
    of a wavenumber-sharded evaluation holds; the boundary terms of a band go with
    its first segment.  Partial sums over a tiling of the segments add up to the full result."""
    P = np.asarray(P, dtype=float).reshape(-1)
    T = np.asarray(T, dtype=float).reshape(-1)
    L = P.size
    p = 0.5 * (P[:-1] + P[1:])
    dP = np.diff(P)
    if T_interface is None:
        T_interface = interface_temperature(p, T, P, T_g)
    B = len(bands['centre'])
    solar = R_star ** 2 / dist ** 2 * (1 - albedo) / 4
    up = np.zeros((L, B))
    dn = np.zeros((L, B))
    for j in range(B):
        s, n = int(bands['start'][j]), int(bands['len'][j])
        w = band_bin_weights(nu, s, n, bands['delta'][j])
        w_all = np.sum(w)
        if segment_range is not None:
            first = int(np.sum(-(-np.asarray(bands['len'][:j], dtype=np.int64) // bins_per_segment)))
            seg = first + np.arange(n) // bins_per_segment
            keep = (seg >= segment_range[0]) & (seg < segment_range[1])
            w = np.where(keep, w, 0.0)
            if not (segment_range[0] <= first < segment_range[1]):
                w_all = 0.0
        I = integrand[:, s:s + n]
        dtau = dP.reshape(-1, 1) * (I[1:] + I[:-1]) / 2.0       # layer optical thickness
        t = np.exp(-dtau)
        om = 1.0 - t
        c = bands['centre'][j]
        bg = np.pi * planck_wavenumber(c, T_g)
        bs = np.pi * planck_wavenumber(c, T_star) * solar
        lw = not bands['sw'][j]
        if lw:
            with np.errstate(over='ignore', invalid='ignore', divide='ignore'):
                bc = np.pi * planck_wavenumber(c, T)
                bi = np.pi * planck_wavenumber(c, T_interface)
            dn_last, dn_gen, up_first, up_gen = emission_coefficients(P, bc, bi, bg)
        # downward sweep, TOA -> surface
        E = np.ones(n)
        H = np.zeros(n)
        dn[0, j] = bs * w_all
        for k in range(L - 1):
            E = E * t[k]
            val = bs * E
            if lw:
                S = t[k] * H
                val = val + S + dn_last[k] * om[k]
                H = S + dn_gen[k] * om[k]
            dn[k + 1, j] = np.sum(w * val)
        # upward sweep, surface -> TOA
        E = np.ones(n)
        S = np.zeros(n)
        up[L - 1, j] = bg * w_all
        for i in range(L - 2, -1, -1):
            E = E * t[i]
            val = bg * E
            if lw:
                y = t[i] * S
                val = val + y + up_first[i] * om[i]
                S = y + up_gen[i] * om[i]
            up[i, j] = np.sum(w * val)
    return up, dn


def layer_integrand(p, T, nu, tables, q_values):
    """I[L, N] = sum_mol k_mol q_mol / g on sorted p (the array real_gas.py:119-121 builds)."""
    p = np.sort(np.asarray(p, dtype=float))
    integrand = np.zeros((p.size, np.size(nu)))
    for table, q in zip(tables, q_values):
        integrand += absorption_coef_on_grid(p, T, nu, table) * np.asarray(q).reshape(-1, 1)
    return integrand / g
