// Plan object: everything RealGas.__init__ (real_gas.py:275-292) derives from the grids, kept
// resident in HBM -- pressure quadrature weights, spline operator, band/chain/segment tables and
// the absorption-coefficient level table.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstring>

#include "internal.cuh"

namespace cmb {

static thread_local std::string t_error;
std::atomic<long long> g_launches{0};

void set_error(const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    t_error = buf;
}

// Not-a-knot cubic spline through the cell centres p_k = (P_k + P_{k+1})/2, evaluated at (and,
// for the top interface, extrapolated to) the interfaces P_i, as a linear operator on the data:
// T_int = A T.  This is what InterpolatedUnivariateSpline(p, T)(p_interface) computes
// (real_gas.py:634-635; FITPACK k=3 interpolating spline == not-a-knot cubic).  Returned
// transposed, At[k][i].
int spline_operator_host(const double *P, int L, std::vector<double> &At) {
    const int n = L - 1;  // knots
    std::vector<double> x(n), h(n - 1);
    for (int k = 0; k < n; ++k) x[k] = 0.5 * (P[k] + P[k + 1]);
    for (int k = 0; k < n - 1; ++k) h[k] = x[k + 1] - x[k];
    // unknown second derivatives m: Mm = R y.  Augmented system [M | R], n x 2n.
    const int w = 2 * n;
    std::vector<double> a((size_t)n * w, 0.0);
    auto Mx = [&](int r, int c) -> double & { return a[(size_t)r * w + c]; };
    auto Rx = [&](int r, int c) -> double & { return a[(size_t)r * w + n + c]; };
    for (int i = 1; i < n - 1; ++i) {
        Mx(i, i - 1) = h[i - 1];
        Mx(i, i) = 2.0 * (h[i - 1] + h[i]);
        Mx(i, i + 1) = h[i];
        Rx(i, i - 1) = 6.0 / h[i - 1];
        Rx(i, i) = -6.0 / h[i - 1] - 6.0 / h[i];
        Rx(i, i + 1) = 6.0 / h[i];
    }
    // third derivative continuous across the second and the last-but-one knot
    Mx(0, 0) = h[1];
    Mx(0, 1) = -(h[0] + h[1]);
    Mx(0, 2) = h[0];
    Mx(n - 1, n - 3) = h[n - 2];
    Mx(n - 1, n - 2) = -(h[n - 3] + h[n - 2]);
    Mx(n - 1, n - 1) = h[n - 3];
    // Gauss-Jordan with partial pivoting
    for (int c = 0; c < n; ++c) {
        int piv = c;
        for (int r = c + 1; r < n; ++r)
            if (std::fabs(Mx(r, c)) > std::fabs(Mx(piv, c))) piv = r;
        if (Mx(piv, c) == 0.0) {
            set_error("pressure grid gives a singular spline system");
            return CMB200_ERR_ARG;
        }
        if (piv != c)
            for (int k = 0; k < w; ++k) std::swap(a[(size_t)piv * w + k], a[(size_t)c * w + k]);
        const double inv = 1.0 / Mx(c, c);
        for (int r = 0; r < n; ++r) {
            if (r == c) continue;
            const double f = Mx(r, c) * inv;
            if (f == 0.0) continue;
            for (int k = c; k < w; ++k) a[(size_t)r * w + k] -= f * a[(size_t)c * w + k];
        }
    }
    for (int r = 0; r < n; ++r) {
        const double inv = 1.0 / Mx(r, r);
        for (int k = 0; k < n; ++k) Rx(r, k) *= inv;  // now S = M^-1 R
    }
    At.assign((size_t)n * L, 0.0);
    for (int i = 0; i < L; ++i) {
        const double q = P[i];
        int s = (int)(std::upper_bound(x.begin(), x.end(), q) - x.begin()) - 1;
        s = std::min(std::max(s, 0), n - 2);
        const double hs = h[s];
        const double ca = (x[s + 1] - q) / hs, cb = (q - x[s]) / hs;
        const double fa = (ca * ca * ca - ca) * hs * hs / 6.0, fb = (cb * cb * cb - cb) * hs * hs / 6.0;
        for (int k = 0; k < n; ++k) At[(size_t)k * L + i] = fa * Rx(s, k) + fb * Rx(s + 1, k);
        At[(size_t)s * L + i] += ca;
        At[(size_t)(s + 1) * L + i] += cb;
    }
    return CMB200_OK;
}

// Trapezoid weights in pressure of flux_integrals (real_gas.py:603-626), regrouped per layer
// and divided by the layer thickness; see DESIGN.md section 3 for the derivation.
// geo[10][L-1]: dn_last = g0 bi_k + g1 bc_k + g2 bi_{k+1};  dn_gen = g3 bi_k + g4 bc_k;
//               up_first = g5 bi_k + g6 bc_k + g7 bi_{k+1}; up_gen = g8 bc_k + g9 bi_{k+1}.
static void pressure_quadrature(const std::vector<double> &P, std::vector<double> &geo) {
    const int L = (int)P.size(), n = L - 1;
    std::vector<double> p(n);
    for (int k = 0; k < n; ++k) p[k] = 0.5 * (P[k] + P[k + 1]);
    geo.assign((size_t)10 * n, 0.0);
    auto G = [&](int a, int k) -> double & { return geo[(size_t)a * n + k]; };
    for (int k = 0; k < n; ++k) {
        const double dP = P[k + 1] - P[k];
        const bool first = (k == 0), last = (k == n - 1);
        // downward, value at interface k+1 (last layer of the path)
        if (first) {
            G(0, k) = 0.5 * (p[0] - P[0]) / dP;
            G(1, k) = 0.5 * (P[1] - P[0]) / dP;
            G(2, k) = 0.5 * (P[1] - p[0]) / dP;
        } else {
            G(1, k) = 0.5 * (P[k + 1] - p[k - 1]) / dP;
            G(2, k) = 0.5 * (P[k + 1] - p[k]) / dP;
        }
        // downward, carried to deeper interfaces
        if (!last) {
            if (first) {
                G(3, k) = 0.5 * (p[0] - P[0]) / dP;
                G(4, k) = 0.5 * (p[1] - P[0]) / dP;
            } else {
                G(4, k) = 0.5 * (p[k + 1] - p[k - 1]) / dP;
            }
        }
        // upward, value at interface k (first layer of the path)
        G(5, k) = 0.5 * (p[k] - P[k]) / dP;
        if (!last) {
            G(6, k) = 0.5 * (p[k + 1] - P[k]) / dP;
        } else {
            G(6, k) = 0.5 * (P[n] - P[n - 1]) / dP;
            G(7, k) = 0.5 * (P[n] - p[n - 1]) / dP;
        }
        // upward, carried to higher interfaces
        if (!first) {
            if (!last) {
                G(8, k) = 0.5 * (p[k + 1] - p[k - 1]) / dP;
            } else {
                G(8, k) = 0.5 * (P[n] - p[n - 2]) / dP;
                G(9, k) = 0.5 * (P[n] - p[n - 1]) / dP;
            }
        }
    }
}

int build_segments(cmb200_plan *p) {
    std::vector<int32_t> sb, sc;
    const int span = 32 * kChains;
    int64_t g = 0;  // segment index over the whole band list
    for (int j = 0; j < p->B; ++j)
        for (int c0 = 0; c0 < p->hlen[j]; c0 += span, ++g) {
            if (j < p->band_lo || j >= p->band_hi || g < p->seg_lo || g >= p->seg_hi) continue;
            sb.push_back(j);
            sc.push_back(c0);
        }
    p->n_seg_total = g;
    p->n_seg = (int)sb.size();
    p->hseg_band = sb;
    p->hseg_c0 = sc;
    CMB_CUDA(p->seg_band.upload(sb.data(), sb.size()));
    CMB_CUDA(p->seg_c0.upload(sc.data(), sc.size()));
    // per-lane chain of every segment, flat, so that the kernel fetches a segment with two coalesced loads
    std::vector<int32_t> bin(sb.size() * 32 + 32, 0);
    std::vector<double> w64(sb.size() * 32 + 32, 0.0);
    std::vector<float> w32(sb.size() * 32 + 32, 0.f);
    for (size_t s = 0; s < sb.size(); ++s)
        for (int lane = 0; lane < 32; ++lane) {
            const int j = sb[s], ci = sc[s] + lane;
            const bool act = ci < p->hlen[j];
            bin[s * 32 + lane] = p->hstart[j] + (act ? ci : 0);
            if (act && !p->hwchain.empty()) w64[s * 32 + lane] = p->hwchain[p->hchain_off[j] + ci];
            w32[s * 32 + lane] = (float)w64[s * 32 + lane];
        }
    CMB_CUDA(p->seg_bin.upload(bin.data(), bin.size()));
    if (p->precision == CMB200_F64)
        CMB_CUDA(p->seg_w64.upload(w64.data(), w64.size()));
    else
        CMB_CUDA(p->seg_w32.upload(w32.data(), w32.size()));
    CMB_CUDA(cudaDeviceSynchronize());
    p->partition.n_col = -1;
    p->ks_valid = false;
    return CMB200_OK;
}

}  // namespace cmb

using namespace cmb;

extern "C" const char *cmb200_last_error(void) { return t_error.c_str(); }
extern "C" int cmb200_version(void) { return 100; }
extern "C" int64_t cmb200_launch_count(void) { return (int64_t)g_launches.load(); }
extern "C" int cmb200_device_count(int *count) {
    CMB_CHECK_ARG(count, "null pointer");
    *count = 0;
    CMB_CUDA(cudaGetDeviceCount(count));
    return CMB200_OK;
}

extern "C" int cmb200_plan_create(cmb200_plan **plan, int device, int precision,
                                  const cmb200_constants *c, int32_t L, const double *P, int64_t N,
                                  const double *nu, int32_t B, const int32_t *band_start,
                                  const int32_t *band_len, const double *band_centre,
                                  const double *band_delta, const uint8_t *band_sw) {
    CMB_CHECK_ARG(plan && c && P && nu && band_start && band_len && band_centre && band_delta && band_sw,
                  "null pointer");
    CMB_CHECK_ARG(precision == CMB200_F64 || precision == CMB200_F32, "precision must be CMB200_F64 or CMB200_F32");
    CMB_CHECK_ARG(L >= 5, "need at least 5 pressure interfaces (cubic spline through L-1 cell centres), got %d", L);
    CMB_CHECK_ARG(N >= 1 && N < (int64_t)1 << 31, "bad number of wavenumber bins");
    CMB_CHECK_ARG(B >= 1, "need at least one band");
    for (int i = 1; i < L; ++i) CMB_CHECK_ARG(P[i] > P[i - 1], "interface pressures must be strictly ascending");
    CMB_CHECK_ARG(P[0] > 0, "top interface pressure must be positive");
    for (int j = 0; j < B; ++j) {
        CMB_CHECK_ARG(band_len[j] >= 1 && band_start[j] >= 0 && (int64_t)band_start[j] + band_len[j] <= N,
                      "band %d bin run [%d, +%d) outside the wavenumber grid", j, band_start[j], band_len[j]);
        CMB_CHECK_ARG(band_delta[j] != 0.0, "band %d has zero width", j);
    }
    CMB_CUDA(cudaSetDevice(device));
    cmb200_plan *p = new cmb200_plan();
    p->device = device;
    p->precision = precision;
    p->c = *c;
    p->L = L;
    p->N = N;
    p->B = B;
    p->hP.assign(P, P + L);
    p->hnu.assign(nu, nu + N);
    p->hstart.assign(band_start, band_start + B);
    p->hlen.assign(band_len, band_len + B);
    p->hcentre.assign(band_centre, band_centre + B);
    p->hdelta.assign(band_delta, band_delta + B);
    p->hsw.assign(band_sw, band_sw + B);
    p->band_lo = 0;
    p->band_hi = B;

    std::vector<double> dx(L), hdP(L - 1), geo, At;
    for (int i = 0; i < L; ++i) dx[i] = P[i] - (i ? P[i - 1] : 0.0);
    for (int k = 0; k < L - 1; ++k) hdP[k] = (P[k + 1] - P[k]) / 2.0;
    pressure_quadrature(p->hP, geo);
    int rc = spline_operator_host(P, L, At);
    if (rc) {
        delete p;
        return rc;
    }
    // nu-trapezoid weights of every (band, bin) chain, divided by the band width (real_gas.py:154)
    std::vector<int64_t> off(B + 1, 0);
    for (int j = 0; j < B; ++j) off[j + 1] = off[j] + band_len[j];
    p->n_chains = off[B];
    std::vector<double> w64(off[B]), sumw(B);
    std::vector<float> w32(off[B]);
    for (int j = 0; j < B; ++j) {
        const double *x = nu + band_start[j];
        const int len = band_len[j];
        double s = 0.0;
        for (int i = 0; i < len; ++i) {
            double wi = 0.0;
            if (i > 0) wi += 0.5 * (x[i] - x[i - 1]);
            if (i < len - 1) wi += 0.5 * (x[i + 1] - x[i]);
            wi /= band_delta[j];
            w64[off[j] + i] = wi;
            w32[off[j] + i] = (float)wi;
            s += wi;
        }
        sumw[j] = s;
    }
    cudaError_t e = cudaSuccess;
    auto ok = [&](cudaError_t x) {
        if (e == cudaSuccess) e = x;
    };
    ok(p->P.upload(P, L));
    ok(p->dx.upload(dx.data(), L));
    ok(p->hdP.upload(hdP.data(), L - 1));
    ok(p->geo.upload(geo.data(), geo.size()));
    ok(p->splineT.upload(At.data(), At.size()));
    ok(p->band_start.upload(band_start, B));
    ok(p->band_len.upload(band_len, B));
    ok(p->band_centre.upload(band_centre, B));
    ok(p->band_delta.upload(band_delta, B));
    ok(p->band_sw.upload(band_sw, B));
    ok(p->band_sumw.upload(sumw.data(), B));
    ok(p->chain_off.upload(off.data(), B + 1));
    if (precision == CMB200_F64)
        ok(p->wchain64.upload(w64.data(), w64.size()));
    else
        ok(p->wchain32.upload(w32.data(), w32.size()));
    ok(cudaDeviceSynchronize());
    if (e != cudaSuccess) {
        set_error("CUDA error %s while creating the plan: %s", cudaGetErrorName(e), cudaGetErrorString(e));
        delete p;
        return CMB200_ERR_CUDA;
    }
    p->hwchain = w64;
    p->hchain_off = off;
    rc = build_segments(p);
    if (rc) {
        delete p;
        return rc;
    }
    *plan = p;
    return CMB200_OK;
}

extern "C" int cmb200_plan_destroy(cmb200_plan *plan) {
    if (!plan) return CMB200_OK;
    cudaSetDevice(plan->device);
    delete plan;
    return CMB200_OK;
}

extern "C" int cmb200_plan_set_star(cmb200_plan *plan, double T_star, double R_star, double dist) {
    CMB_CHECK_ARG(plan, "null plan");
    CMB_CHECK_ARG(T_star > 0 && R_star >= 0 && dist > 0, "bad star parameters");
    plan->T_star = T_star;
    plan->R_star = R_star;
    plan->dist = dist;
    return CMB200_OK;
}

extern "C" int cmb200_plan_set_band_range(cmb200_plan *plan, int32_t band_lo, int32_t band_hi) {
    CMB_CHECK_ARG(plan, "null plan");
    CMB_CHECK_ARG(0 <= band_lo && band_lo <= band_hi && band_hi <= plan->B, "bad band range [%d, %d)", band_lo, band_hi);
    CMB_CUDA(cudaSetDevice(plan->device));
    plan->band_lo = band_lo;
    plan->band_hi = band_hi;
    return build_segments(plan);
}

extern "C" int cmb200_plan_set_segment_range(cmb200_plan *plan, int64_t seg_lo, int64_t seg_hi) {
    CMB_CHECK_ARG(plan, "null plan");
    CMB_CHECK_ARG(0 <= seg_lo && seg_lo <= seg_hi, "bad segment range");
    CMB_CUDA(cudaSetDevice(plan->device));
    plan->seg_lo = seg_lo;
    plan->seg_hi = seg_hi;
    return build_segments(plan);
}

extern "C" int cmb200_plan_segments(const cmb200_plan *plan, int32_t *bins_per_segment, int64_t *n_segments) {
    CMB_CHECK_ARG(plan, "null plan");
    if (bins_per_segment) *bins_per_segment = 32 * kChains;
    if (n_segments) *n_segments = plan->n_seg_total;
    return CMB200_OK;
}

extern "C" int cmb200_plan_info(const cmb200_plan *plan, int32_t *L, int64_t *N, int32_t *B,
                                int32_t *M, int32_t *nT, int32_t *precision, int64_t *n_chains) {
    CMB_CHECK_ARG(plan, "null plan");
    if (L) *L = plan->L;
    if (N) *N = plan->N;
    if (B) *B = plan->B;
    if (M) *M = plan->M;
    if (nT) *nT = plan->nT_pad;
    if (precision) *precision = plan->precision;
    if (n_chains) *n_chains = plan->n_chains;
    return CMB200_OK;
}
