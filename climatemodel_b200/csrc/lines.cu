// Line-by-line absorption coefficient: the producer of the lookup tables the hot path reads
// (get_absorption_coefficient, real_gas_data/hitran.py:223-247 with gamma_extrapolate :135-150,
// s_extrapolate :153-167, lorentzian_profile :170-179, wavenumbers_near_line :182-199;
// SURVEY.md 8f row f3).
//
//   k[ip][i] = sum over lines l with i1_l <= i < i2_l of  S_l(T_ip) (1/pi) g / (g^2 + (nu_i - nu0_l)^2),
//   g = gamma_l (p_ip / p_ref) (T_ref / T_ip)^n_l,   S_l(T) = sw_l (T/T_ref)^n_l exp(-(h c nu0_l / k)(1/T - 1/T_ref)),
// the window [i1, i2) = centre bin +- int(n_line_widths * max_ip(g) / d_nu) clipped to the grid.
//
// Two kernels: line_window_kernel (one thread per line: widest half-width over the pressure
// levels, centre bin, window) and line_sum_kernel (one thread per (level, bin); lines staged
// through shared memory in chunks, chunks whose windows miss the block's bins are skipped; the
// sum runs in line order with separately rounded operations like NumPy's, so it differs from the
// reference only through pow / exp, a few ulp).
#include "internal.cuh"

namespace cmb {

struct LineConst {
    double p_ref, T_ref, hc_over_k;   // hc_over_k = 100 h c / k_B  (wavenumber in cm^-1 -> K)
};

__device__ __forceinline__ double line_gamma(double p, double T, double gamma_ref, double n, const LineConst &c) {
    return __dmul_rn(__dmul_rn(gamma_ref, __ddiv_rn(p, c.p_ref)), pow(__ddiv_rn(c.T_ref, T), n));
}

__global__ void line_window_kernel(LineConst c, int n_p, const double *__restrict__ p, const double *__restrict__ T,
                                   int64_t n_lines, const double *__restrict__ gamma_air,
                                   const double *__restrict__ n_air, const int32_t *__restrict__ centre,
                                   double n_line_widths, double spacing, int64_t n_nu, int32_t *__restrict__ i1,
                                   int32_t *__restrict__ i2) {
    const int64_t l = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (l >= n_lines) return;
    double gmax = -1.0;
    for (int ip = 0; ip < n_p; ++ip) gmax = fmax(gmax, line_gamma(p[ip], T[ip], gamma_air[l], n_air[l], c));
    // int(n_line_widths * line_width / array_spacing): truncation toward zero, clamped to the grid size
    const double w = __ddiv_rn(__dmul_rn(n_line_widths, gmax), spacing);
    const int64_t nw = w >= (double)n_nu ? n_nu : (int64_t)w;
    const int64_t ci = centre[l];
    i1[l] = (int32_t)max((int64_t)0, ci - nw);
    i2[l] = (int32_t)(min(n_nu - 1, ci + nw) + 1);
}

constexpr int kLineChunk = 256;

__global__ void __launch_bounds__(kLineChunk) line_sum_kernel(LineConst c, const double *__restrict__ p,
                                                              const double *__restrict__ T, int64_t n_nu,
                                                              const double *__restrict__ nu, int64_t n_lines,
                                                              const double *__restrict__ line_nu,
                                                              const double *__restrict__ line_sw,
                                                              const double *__restrict__ gamma_air,
                                                              const double *__restrict__ n_air,
                                                              const int32_t *__restrict__ i1,
                                                              const int32_t *__restrict__ i2, double *__restrict__ out) {
    __shared__ double s_nu0[kLineChunk], s_g[kLineChunk], s_S[kLineChunk];
    __shared__ int32_t s_i1[kLineChunk], s_i2[kLineChunk];
    __shared__ int s_any;
    const int ip = blockIdx.y;
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t tile_lo = blockIdx.x * (int64_t)blockDim.x, tile_hi = min(n_nu, tile_lo + blockDim.x);
    const double pp = p[ip], TT = T[ip];
    const double x = i < n_nu ? nu[i] : 0.0;
    double acc = 0.0;
    for (int64_t l0 = 0; l0 < n_lines; l0 += kLineChunk) {
        const int64_t l = l0 + threadIdx.x;
        __syncthreads();
        if (threadIdx.x == 0) s_any = 0;
        __syncthreads();
        int32_t a = 0, b = 0;
        if (l < n_lines) {
            a = i1[l];
            b = i2[l];
        }
        const bool hit = a < tile_hi && b > tile_lo;
        if (hit) s_any = 1;
        s_i1[threadIdx.x] = a;
        s_i2[threadIdx.x] = b;
        __syncthreads();
        if (!s_any) continue;                               // no line of this chunk reaches the block's bins
        if (hit) {
            const double nu0 = line_nu[l], n = n_air[l];
            s_nu0[threadIdx.x] = nu0;
            s_g[threadIdx.x] = line_gamma(pp, TT, gamma_air[l], n, c);
            // sw (T/T_ref)^n exp(-(h freq / k)(1/T - 1/T_ref)),  freq = 100 nu0 c
            const double e = exp(__dmul_rn(-__dmul_rn(c.hc_over_k, nu0), __dsub_rn(__ddiv_rn(1.0, TT), __ddiv_rn(1.0, c.T_ref))));
            s_S[threadIdx.x] = __dmul_rn(__dmul_rn(line_sw[l], pow(__ddiv_rn(TT, c.T_ref), n)), e);
        }
        __syncthreads();
        const int n_here = (int)min((int64_t)kLineChunk, n_lines - l0);
        if (i < n_nu)
            for (int k = 0; k < n_here; ++k) {
                if (i < s_i1[k] || i >= s_i2[k]) continue;
                const double g = s_g[k], d = __dsub_rn(x, s_nu0[k]);
                const double shape = __ddiv_rn(__dmul_rn(1.0 / M_PI, g), __dadd_rn(__dmul_rn(g, g), __dmul_rn(d, d)));
                acc = __dadd_rn(acc, __dmul_rn(s_S[k], shape));
            }
    }
    if (i < n_nu) out[(int64_t)ip * n_nu + i] = acc;
}

}  // namespace cmb

using namespace cmb;

extern "C" int cmb200_line_absorption_host(const cmb200_constants *c, double p_reference, double T_reference,
                                           int64_t n_p, const double *p, const double *T, int64_t n_nu,
                                           const double *nu, int64_t n_lines, const double *line_nu,
                                           const double *line_sw, const double *line_gamma_air,
                                           const double *line_n_air, double n_line_widths, double *k) {
    CMB_CHECK_ARG(c && p && T && nu && k, "null pointer");
    CMB_CHECK_ARG(n_p >= 0 && n_nu >= 0 && n_lines >= 0, "negative size");
    CMB_CHECK_ARG(n_lines == 0 || (line_nu && line_sw && line_gamma_air && line_n_air), "null line data");
    CMB_CHECK_ARG(n_p <= 65535, "at most 65535 pressure levels per call");
    CMB_CHECK_ARG(n_nu < ((int64_t)1 << 31), "wavenumber grid too long");
    CMB_CHECK_ARG(p_reference > 0 && T_reference > 0, "reference pressure and temperature must be positive");
    if (n_p == 0 || n_nu == 0) return CMB200_OK;
    CMB_CHECK_ARG(n_nu >= 2, "need two wavenumbers for the grid spacing");
    DevBuf<double> dp, dT, dnu, dl, dsw, dg, dn, dout;
    DevBuf<int32_t> centre, i1, i2;
    CMB_CUDA(dp.upload(p, n_p));
    CMB_CUDA(dT.upload(T, n_p));
    CMB_CUDA(dnu.upload(nu, n_nu));
    CMB_CUDA(dout.alloc((size_t)n_p * n_nu));
    if (n_lines == 0) {
        CMB_CUDA(cudaMemset(dout.p, 0, (size_t)n_p * n_nu * sizeof(double)));
    } else {
        CMB_CUDA(dl.upload(line_nu, n_lines));
        CMB_CUDA(dsw.upload(line_sw, n_lines));
        CMB_CUDA(dg.upload(line_gamma_air, n_lines));
        CMB_CUDA(dn.upload(line_n_air, n_lines));
        CMB_CUDA(centre.alloc(n_lines));
        CMB_CUDA(i1.alloc(n_lines));
        CMB_CUDA(i2.alloc(n_lines));
        bool ascending = true;
        for (int64_t i = 1; i < n_nu && ascending; ++i) ascending = nu[i] > nu[i - 1];
        int rc = launch_nearest(dl.p, n_lines, dnu.p, (int)n_nu, ascending, false, 0, 0, centre.p, 0);
        if (rc) return rc;
        const LineConst lc{p_reference, T_reference, c->h_planck * (100.0 * c->speed_of_light) / c->k_boltzmann};
        line_window_kernel<<<(unsigned)((n_lines + 127) / 128), 128>>>(lc, (int)n_p, dp.p, dT.p, n_lines, dg.p, dn.p,
                                                                      centre.p, n_line_widths, nu[1] - nu[0], n_nu,
                                                                      i1.p, i2.p);
        CMB_LAUNCHED();
        dim3 grid((unsigned)((n_nu + kLineChunk - 1) / kLineChunk), (unsigned)n_p);
        line_sum_kernel<<<grid, kLineChunk>>>(lc, dp.p, dT.p, n_nu, dnu.p, n_lines, dl.p, dsw.p, dg.p, dn.p, i1.p,
                                              i2.p, dout.p);
        CMB_LAUNCHED();
    }
    CMB_CUDA(cudaMemcpy(k, dout.p, (size_t)n_p * n_nu * sizeof(double), cudaMemcpyDeviceToHost));
    return CMB200_OK;
}
