// Band-averaged transmission function and its q-derivative as standalone calls
// (transmission, real_gas.py:130-154; dtransmission_dq, :157-184).  The flux path does not go
// through these -- it uses the per-bin recurrence -- but notebooks and
// centa_presentation/base.py call them directly.
#include "internal.cuh"

namespace cmb {

// One warp per output entry (a, b); lanes stride over the band's bins; trapezoid in nu formed
// like np.trapz: sum_i d_i (e_{i+1} + e_i) / 2, then divided by the band width.
// Entries with tau[a] > tau[b] overflow to +inf exactly as NumPy's do.
__global__ void transmission_kernel(const double *__restrict__ tau, int64_t n_nu,
                                    const int32_t *__restrict__ rows1, int n1,
                                    const int32_t *__restrict__ rows2, int n2,
                                    const int32_t *__restrict__ bins,
                                    const double *__restrict__ nu_band, int n_bins, double delta,
                                    const double *__restrict__ k_band,
                                    const double *__restrict__ scale, double g,
                                    double *__restrict__ out) {
    const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t e = (int64_t)blockIdx.x * warps + warp;
    if (e >= (int64_t)n1 * n2) return;
    const int ia = rows1[e / n2], ib = rows2[e % n2];
    const double *ta = tau + (int64_t)ia * n_nu, *tb = tau + (int64_t)ib * n_nu;
    const double sc = scale ? scale[e] : 1.0;
    auto f = [&](int i) {
        int n = bins[i];
        double v = exp(ta[n] - tb[n]);
        if (k_band) v = sc * k_band[i] * v / g;   // (p1 - p2) * k * exp(.) / g, :182-183
        return v;
    };
    double acc = 0.0;
    for (int i = lane; i < n_bins - 1; i += 32) acc += (nu_band[i + 1] - nu_band[i]) * (f(i + 1) + f(i)) / 2.0;
    acc = warp_sum(acc);
    if (lane == 0) out[e] = acc / delta;
}

static int transmission_host(const double *tau, int64_t n_lev, int64_t n_nu, const int32_t *rows1,
                             int32_t n1, const int32_t *rows2, int32_t n2, const int32_t *bins,
                             const double *nu_band, int32_t n_bins, double delta,
                             const double *k_band, const double *scale, double g, double *out) {
    CMB_CHECK_ARG(tau && rows1 && rows2 && bins && nu_band && out, "null pointer");
    CMB_CHECK_ARG(n_lev > 0 && n_nu > 0 && n1 >= 0 && n2 >= 0 && n_bins >= 0, "bad size");
    for (int i = 0; i < n1; ++i) CMB_CHECK_ARG(rows1[i] >= 0 && rows1[i] < n_lev, "rows1[%d] out of range", i);
    for (int i = 0; i < n2; ++i) CMB_CHECK_ARG(rows2[i] >= 0 && rows2[i] < n_lev, "rows2[%d] out of range", i);
    for (int i = 0; i < n_bins; ++i) CMB_CHECK_ARG(bins[i] >= 0 && bins[i] < n_nu, "bins[%d] out of range", i);
    if (n1 == 0 || n2 == 0) return CMB200_OK;
    DevBuf<double> dtau, dnu, dk, dsc, dout;
    DevBuf<int32_t> r1, r2, db;
    CMB_CUDA(dtau.upload(tau, (size_t)n_lev * n_nu));
    CMB_CUDA(r1.upload(rows1, n1));
    CMB_CUDA(r2.upload(rows2, n2));
    CMB_CUDA(db.upload(bins, n_bins));
    CMB_CUDA(dnu.upload(nu_band, n_bins));
    if (k_band) CMB_CUDA(dk.upload(k_band, n_bins));
    if (scale) CMB_CUDA(dsc.upload(scale, (size_t)n1 * n2));
    CMB_CUDA(dout.alloc((size_t)n1 * n2));
    const int warps = 4;
    const int64_t entries = (int64_t)n1 * n2;
    transmission_kernel<<<(unsigned)((entries + warps - 1) / warps), warps * 32>>>(
        dtau.p, n_nu, r1.p, n1, r2.p, n2, db.p, dnu.p, n_bins, delta, dk.p, dsc.p, g, dout.p);
    CMB_LAUNCHED();
    CMB_CUDA(cudaMemcpy(out, dout.p, entries * sizeof(double), cudaMemcpyDeviceToHost));
    return CMB200_OK;
}

}  // namespace cmb

using namespace cmb;

extern "C" int cmb200_transmission_host(const double *tau, int64_t n_lev, int64_t n_nu,
                                        const int32_t *rows1, int32_t n1, const int32_t *rows2,
                                        int32_t n2, const int32_t *bins, const double *nu_band,
                                        int32_t n_bins, double delta, double *out) {
    return transmission_host(tau, n_lev, n_nu, rows1, n1, rows2, n2, bins, nu_band, n_bins, delta,
                             nullptr, nullptr, 1.0, out);
}

extern "C" int cmb200_dtransmission_dq_host(const double *tau, int64_t n_lev, int64_t n_nu,
                                            const int32_t *rows1, int32_t n1, const int32_t *rows2,
                                            int32_t n2, const int32_t *bins, const double *nu_band,
                                            int32_t n_bins, double delta, const double *k_band,
                                            const double *scale, double g, double *out) {
    CMB_CHECK_ARG(k_band && scale, "k_band and scale are required");
    return transmission_host(tau, n_lev, n_nu, rows1, n1, rows2, n2, bins, nu_band, n_bins, delta,
                             k_band, scale, g, out);
}
