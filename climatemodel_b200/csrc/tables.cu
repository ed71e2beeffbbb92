// Stage 1 of the hot path: absorption-coefficient lookup and optical depth.
//   nearest_index / gather ....... real_gas.py:45-83   (integer index work)
//   optical depth ................. real_gas.py:86-127  (trapezoid prefix sum over levels)
// HBM-bound kernels: one thread per wavenumber bin, consecutive lanes on consecutive bins so
// every table row and every tau row is read / written fully coalesced.
#include "internal.cuh"

namespace cmb {

// argmin_i |v - axis[i]| with NumPy's first-index tie-break (real_gas.py:57-60).
__global__ void nearest_index_kernel(const double *__restrict__ values, int64_t n,
                                     const double *__restrict__ axis, int n_axis,
                                     int32_t *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = values[i];
    double best = fabs(v - axis[0]);
    int bi = 0;
    for (int a = 1; a < n_axis; ++a) {
        double d = fabs(v - axis[a]);
        if (d < best) {
            best = d;
            bi = a;
        }
    }
    out[i] = bi;
}

// As above, but -1 when the value lies outside [axis_min, axis_max] (load_absorption_coef's
// crop, real_gas.py:79-82).
__global__ void nearest_index_crop_kernel(const double *__restrict__ values, int64_t n,
                                          const double *__restrict__ axis, int n_axis,
                                          double lo, double hi, int32_t *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = values[i];
    if (!(v >= lo && v <= hi)) {
        out[i] = -1;
        return;
    }
    double best = fabs(v - axis[0]);
    int bi = 0;
    for (int a = 1; a < n_axis; ++a) {
        double d = fabs(v - axis[a]);
        if (d < best) {
            best = d;
            bi = a;
        }
    }
    out[i] = bi;
}

// The wavenumber axis can be long (1e6 bins against 1e6 table bins): a linear scan per query
// is O(N^2).  Table axes produced by hitran.make_table are ascending (np.arange), so use a
// bisection and then compare the two neighbours with the same |.| < rule; ties resolve to the
// lower index, which is what argmin's first-index rule gives on an ascending axis.
__global__ void nearest_index_sorted_kernel(const double *__restrict__ values, int64_t n,
                                            const double *__restrict__ axis, int n_axis,
                                            int crop, int32_t *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = values[i];
    if (crop && !(v >= axis[0] && v <= axis[n_axis - 1])) {
        out[i] = -1;
        return;
    }
    int lo = 0, hi = n_axis - 1;  // find last index with axis[idx] <= v (or 0)
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (axis[mid] <= v)
            lo = mid;
        else
            hi = mid - 1;
    }
    // scan a small neighbourhood in ascending order so that exact ties keep the first index
    int a0 = lo > 0 ? lo - 1 : 0, a1 = lo + 1 < n_axis ? lo + 1 : n_axis - 1;
    int bi = a0;
    double best = fabs(v - axis[a0]);
    for (int a = a0 + 1; a <= a1; ++a) {
        double d = fabs(v - axis[a]);
        if (d < best) {
            best = d;
            bi = a;
        }
    }
    out[i] = bi;
}

// k_out[lev][n] = k[p_idx[lev]][T_idx[lev]][nu_idx[n]]  (0 where nu_idx < 0)
__global__ void gather_kernel(const double *__restrict__ k, int n_T, int64_t n_nu_tab,
                              const int32_t *__restrict__ p_idx, const int32_t *__restrict__ T_idx,
                              const int32_t *__restrict__ nu_idx, int64_t n_lev, int64_t n_nu,
                              double *__restrict__ out) {
    int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t lev = blockIdx.y;
    if (n >= n_nu) return;
    int32_t j = nu_idx[n];
    const double *row = k + ((int64_t)p_idx[lev] * n_T + T_idx[lev]) * n_nu_tab;
    out[lev * n_nu + n] = j >= 0 ? row[j] : 0.0;
}

// Level table: KL[k][t][n] for one molecule on the plan grid, all table temperatures kept.
template <typename Real>
__global__ void level_table_kernel(const double *__restrict__ k, int n_T, int64_t n_nu_tab,
                                   const int32_t *__restrict__ p_idx,
                                   const int32_t *__restrict__ nu_idx, int L, int nT_pad,
                                   int64_t N, int64_t pitch, Real *__restrict__ KL) {
    int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int lev = blockIdx.y;
    int t = blockIdx.z;
    if (n >= pitch) return;
    Real v = Real(0);
    if (n < N && t < n_T) {
        int32_t j = nu_idx[n];
        if (j >= 0) v = (Real)k[((int64_t)p_idx[lev] * n_T + t) * n_nu_tab + j];
    }
    KL[((int64_t)lev * nT_pad + t) * pitch + n] = v;
}
// Level table in segment space: KS[row][c] = KL[row][seg_bin[c]], row = (molecule, level, table T).
template <typename Real>
__global__ void segment_table_kernel(const Real *__restrict__ KL, int64_t n_rows, int64_t pitch,
                                     const int32_t *__restrict__ seg_bin, int64_t n_chains, int64_t spitch,
                                     Real *__restrict__ KS) {
    const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (c >= n_chains) return;           // (the row's padding beyond the last segment stays zero)
    const int32_t n = seg_bin[c];
    for (int64_t row = blockIdx.y; row < n_rows; row += gridDim.y) KS[row * spitch + c] = KL[row * pitch + n];
}
template __global__ void segment_table_kernel<double>(const double *, int64_t, int64_t, const int32_t *, int64_t, int64_t, double *);
template __global__ void segment_table_kernel<float>(const float *, int64_t, int64_t, const int32_t *, int64_t, int64_t, float *);

template __global__ void level_table_kernel<double>(const double *, int, int64_t, const int32_t *,
                                                    const int32_t *, int, int, int64_t, int64_t,
                                                    double *);
template __global__ void level_table_kernel<float>(const double *, int, int64_t, const int32_t *,
                                                   const int32_t *, int, int, int64_t, int64_t,
                                                   float *);

// optical_depth (real_gas.py:119-126) from already-gathered k rows:
//   integrand = (sum_m k_m q_m) / g ; tau_i = tau_{i-1} + dx_i (y_i + y_{i-1}) / 2
// with the products, sums and the division rounded separately (no FMA contraction) in the
// order NumPy evaluates them, so tau is bit-identical to the reference.
// kg[m][lev][n] : gathered coefficients; q[m][lev]; dx[lev] = diff([0, p...]).
__global__ void optical_depth_kernel(const double *__restrict__ kg, int M, int64_t n_lev,
                                     int64_t n_nu, const double *__restrict__ q,
                                     const double *__restrict__ dx, double g,
                                     double *__restrict__ tau) {
    int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (n >= n_nu) return;
    double y_prev = 0.0, run = 0.0;
    for (int64_t lev = 0; lev < n_lev; ++lev) {
        double acc = 0.0;
        for (int m = 0; m < M; ++m)
            acc = __dadd_rn(acc, __dmul_rn(kg[((int64_t)m * n_lev + lev) * n_nu + n], q[m * n_lev + lev]));
        double y = __ddiv_rn(acc, g);
        double term = __dmul_rn(__dmul_rn(dx[lev], __dadd_rn(y, y_prev)), 0.5);
        run = __dadd_rn(run, term);
        tau[lev * n_nu + n] = run;
        y_prev = y;
    }
}

// Same arithmetic straight from the plan's level table, for n_col columns:
// tidx[m][col][lev] selects the temperature row; qcol = q[m][col|0][lev].
template <typename Real>
__global__ void plan_optical_depth_kernel(const Real *__restrict__ KL, int M, int L, int nT_pad,
                                          int64_t N, int64_t pitch,
                                          const uint8_t *__restrict__ tidx,
                                          const double *__restrict__ q, int64_t q_col_stride,
                                          int64_t n_col, const double *__restrict__ dx, double g,
                                          double *__restrict__ tau) {
    int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t col = blockIdx.y;
    if (n >= N) return;
    double y_prev = 0.0, run = 0.0;
    for (int lev = 0; lev < L; ++lev) {
        double acc = 0.0;
        for (int m = 0; m < M; ++m) {
            int t = tidx[((int64_t)m * n_col + col) * L + lev];
            double kv = (double)KL[(((int64_t)m * L + lev) * nT_pad + t) * pitch + n];
            double qv = q[(int64_t)m * (q_col_stride ? n_col * L : L) + col * q_col_stride + lev];
            acc = __dadd_rn(acc, __dmul_rn(kv, qv));
        }
        double y = __ddiv_rn(acc, g);
        double term = __dmul_rn(__dmul_rn(dx[lev], __dadd_rn(y, y_prev)), 0.5);
        run = __dadd_rn(run, term);
        tau[(col * L + lev) * N + n] = run;
        y_prev = y;
    }
}
template __global__ void plan_optical_depth_kernel<double>(const double *, int, int, int, int64_t,
                                                           int64_t, const uint8_t *, const double *,
                                                           int64_t, int64_t, const double *, double,
                                                           double *);
template __global__ void plan_optical_depth_kernel<float>(const float *, int, int, int, int64_t,
                                                          int64_t, const uint8_t *, const double *,
                                                          int64_t, int64_t, const double *, double,
                                                          double *);

// Temperature-row index per (molecule, column, level): argmin |T - T_tab| first-index.
// T_tau may be null: isothermal at T_g[col] (real_gas.py:285-290).
__global__ void tidx_kernel(const double *__restrict__ T_tau, const double *__restrict__ T_g,
                            int64_t n_col, int L, const double *__restrict__ T_axis, int n_T,
                            uint8_t *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n_col * L) return;
    double v = T_tau ? T_tau[i] : T_g[i / L];
    double best = fabs(v - T_axis[0]);
    int bi = 0;
    for (int a = 1; a < n_T; ++a) {
        double d = fabs(v - T_axis[a]);
        if (d < best) {
            best = d;
            bi = a;
        }
    }
    out[i] = (uint8_t)bi;
}

__global__ void planck_kernel(PlanckConst pc, const double *__restrict__ nu, int64_t n_nu,
                              const double *__restrict__ T, int64_t n_T, double *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n_nu * n_T) return;
    out[i] = planck_wavenumber(pc, nu[i % n_nu], T[i / n_nu]);
}

// ---- host-side drivers ---------------------------------------------------------------------
static bool ascending(const double *a, int64_t n) {
    for (int64_t i = 1; i < n; ++i)
        if (!(a[i] > a[i - 1])) return false;
    return true;
}

int launch_nearest(const double *d_values, int64_t n, const double *d_axis, int n_axis,
                   bool axis_ascending, bool crop, double lo, double hi, int32_t *d_out,
                   cudaStream_t s) {
    if (n == 0) return CMB200_OK;
    int threads = 256;
    unsigned blocks = (unsigned)((n + threads - 1) / threads);
    if (axis_ascending && n_axis > 64) {
        nearest_index_sorted_kernel<<<blocks, threads, 0, s>>>(d_values, n, d_axis, n_axis, crop ? 1 : 0, d_out);
    } else if (crop) {
        nearest_index_crop_kernel<<<blocks, threads, 0, s>>>(d_values, n, d_axis, n_axis, lo, hi, d_out);
    } else {
        nearest_index_kernel<<<blocks, threads, 0, s>>>(d_values, n, d_axis, n_axis, d_out);
    }
    CMB_LAUNCHED();
    return CMB200_OK;
}

int check_table(const cmb200_table *t) {
    CMB_CHECK_ARG(t && t->p && t->T && t->nu && t->k, "table has null pointers");
    CMB_CHECK_ARG(t->n_p > 0 && t->n_T > 0 && t->n_nu > 0, "table has empty axes");
    CMB_CHECK_ARG(t->n_T <= 255, "table has more than 255 temperatures");
    return CMB200_OK;
}

// Uploads one table and computes (p_idx, T_idx, nu_idx) for the given query grids.
struct TableOnDevice {
    DevBuf<double> p, T, nu, k;
    bool nu_ascending = false;
    double nu_lo = 0, nu_hi = 0;
    int upload(const cmb200_table *t, cudaStream_t s) {
        int rc = check_table(t);
        if (rc) return rc;
        CMB_CUDA(p.upload(t->p, t->n_p, s));
        CMB_CUDA(T.upload(t->T, t->n_T, s));
        CMB_CUDA(nu.upload(t->nu, t->n_nu, s));
        CMB_CUDA(k.upload(t->k, (size_t)t->n_p * t->n_T * t->n_nu, s));
        nu_ascending = ascending(t->nu, t->n_nu);
        nu_lo = t->nu[0];
        nu_hi = t->nu[0];
        for (int64_t i = 1; i < t->n_nu; ++i) {
            nu_lo = t->nu[i] < nu_lo ? t->nu[i] : nu_lo;
            nu_hi = t->nu[i] > nu_hi ? t->nu[i] : nu_hi;
        }
        return CMB200_OK;
    }
};

static int gather_on_device(const cmb200_table *t, TableOnDevice &dev, const double *d_p,
                            const double *d_T, int64_t n_lev, const double *d_nu, int64_t n_nu,
                            bool zero_outside, double *d_out, cudaStream_t s) {
    DevBuf<int32_t> ip, iT, inu;
    CMB_CUDA(ip.alloc(n_lev));
    CMB_CUDA(iT.alloc(n_lev));
    CMB_CUDA(inu.alloc(n_nu));
    int rc;
    if ((rc = launch_nearest(d_p, n_lev, dev.p.p, t->n_p, false, false, 0, 0, ip.p, s))) return rc;
    if ((rc = launch_nearest(d_T, n_lev, dev.T.p, t->n_T, false, false, 0, 0, iT.p, s))) return rc;
    if ((rc = launch_nearest(d_nu, n_nu, dev.nu.p, t->n_nu, dev.nu_ascending, zero_outside,
                             dev.nu_lo, dev.nu_hi, inu.p, s)))
        return rc;
    if (n_lev > 0 && n_nu > 0) {
        dim3 grid((unsigned)((n_nu + 255) / 256), (unsigned)n_lev);
        gather_kernel<<<grid, 256, 0, s>>>(dev.k.p, t->n_T, t->n_nu, ip.p, iT.p, inu.p, n_lev, n_nu, d_out);
        CMB_LAUNCHED();
    }
    CMB_CUDA(cudaStreamSynchronize(s));  // index buffers die with this scope
    return CMB200_OK;
}

}  // namespace cmb

using namespace cmb;

extern "C" int cmb200_planck_wavenumber_host(const cmb200_constants *c, const double *nu,
                                             int64_t n_nu, const double *T, int64_t n_T,
                                             double *out) {
    CMB_CHECK_ARG(c && nu && T && out, "null pointer");
    CMB_CHECK_ARG(n_nu >= 0 && n_T >= 0, "negative size");
    if (n_nu == 0 || n_T == 0) return CMB200_OK;
    DevBuf<double> dnu, dT, dout;
    CMB_CUDA(dnu.upload(nu, n_nu));
    CMB_CUDA(dT.upload(T, n_T));
    CMB_CUDA(dout.alloc(n_nu * n_T));
    PlanckConst pc{c->h_planck, c->speed_of_light, c->k_boltzmann};
    int64_t n = n_nu * n_T;
    planck_kernel<<<(unsigned)((n + 255) / 256), 256>>>(pc, dnu.p, n_nu, dT.p, n_T, dout.p);
    CMB_LAUNCHED();
    CMB_CUDA(cudaMemcpy(out, dout.p, n * sizeof(double), cudaMemcpyDeviceToHost));
    return CMB200_OK;
}

extern "C" int cmb200_absorption_coef_host(const cmb200_table *table, const double *p,
                                           const double *T, int64_t n_lev, const double *nu,
                                           int64_t n_nu, int zero_outside, double *k_out) {
    CMB_CHECK_ARG(p && T && nu && k_out, "null pointer");
    CMB_CHECK_ARG(n_lev >= 0 && n_nu >= 0, "negative size");
    TableOnDevice dev;
    int rc = dev.upload(table, 0);
    if (rc) return rc;
    if (n_lev == 0 || n_nu == 0) return CMB200_OK;
    DevBuf<double> dp, dT, dnu, dout;
    CMB_CUDA(dp.upload(p, n_lev));
    CMB_CUDA(dT.upload(T, n_lev));
    CMB_CUDA(dnu.upload(nu, n_nu));
    CMB_CUDA(dout.alloc(n_lev * n_nu));
    rc = gather_on_device(table, dev, dp.p, dT.p, n_lev, dnu.p, n_nu, zero_outside != 0, dout.p, 0);
    if (rc) return rc;
    CMB_CUDA(cudaMemcpy(k_out, dout.p, n_lev * n_nu * sizeof(double), cudaMemcpyDeviceToHost));
    return CMB200_OK;
}

extern "C" int cmb200_optical_depth_host(const cmb200_constants *c, int32_t M,
                                         const cmb200_table *tables, const double *p_sorted,
                                         const double *T, int64_t n_lev, const double *nu,
                                         int64_t n_nu, const double *q, double *tau_out) {
    CMB_CHECK_ARG(c && p_sorted && T && nu && tau_out, "null pointer");
    CMB_CHECK_ARG(M >= 0 && (M == 0 || (tables && q)), "tables / q missing");
    CMB_CHECK_ARG(n_lev >= 0 && n_nu >= 0, "negative size");
    if (n_lev == 0 || n_nu == 0) return CMB200_OK;
    for (int64_t i = 1; i < n_lev; ++i)
        CMB_CHECK_ARG(p_sorted[i] >= p_sorted[i - 1], "p must be sorted ascending");
    DevBuf<double> dp, dT, dnu, dq, ddx, kg, dtau;
    CMB_CUDA(dp.upload(p_sorted, n_lev));
    CMB_CUDA(dT.upload(T, n_lev));
    CMB_CUDA(dnu.upload(nu, n_nu));
    CMB_CUDA(dq.upload(q, (size_t)M * n_lev));
    std::vector<double> dx(n_lev);  // host diff([0, p...]) exactly as np.diff forms it
    for (int64_t i = 0; i < n_lev; ++i) dx[i] = p_sorted[i] - (i ? p_sorted[i - 1] : 0.0);
    CMB_CUDA(ddx.upload(dx.data(), n_lev));
    CMB_CUDA(cudaStreamSynchronize(0));
    CMB_CUDA(kg.alloc((size_t)(M > 0 ? M : 1) * n_lev * n_nu));
    CMB_CUDA(dtau.alloc(n_lev * n_nu));
    for (int m = 0; m < M; ++m) {
        TableOnDevice dev;
        int rc = dev.upload(&tables[m], 0);
        if (rc) return rc;
        rc = gather_on_device(&tables[m], dev, dp.p, dT.p, n_lev, dnu.p, n_nu, true,
                              kg.p + (size_t)m * n_lev * n_nu, 0);
        if (rc) return rc;
    }
    optical_depth_kernel<<<(unsigned)((n_nu + 127) / 128), 128>>>(kg.p, M, n_lev, n_nu, dq.p, ddx.p, c->g, dtau.p);
    CMB_LAUNCHED();
    CMB_CUDA(cudaMemcpy(tau_out, dtau.p, n_lev * n_nu * sizeof(double), cudaMemcpyDeviceToHost));
    return CMB200_OK;
}
