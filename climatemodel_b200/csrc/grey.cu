// Grey two-stream fluxes, batched over columns (GreyGas.get_lw_flux, grey.py:251-275 and
// get_sw_flux, :277-294; SURVEY.md 8f "grey kernel").  Arrays are laid out like the reference's:
// [level][column], level 0 at the surface, so consecutive threads (columns) read consecutive
// addresses; one thread walks its column from the top boundary condition down to the surface.
// HBM-bound: 16 bytes read and 16-48 written per (level, column).  Products and sums are rounded
// one by one like NumPy's expression; only exp / pow may differ in the last place.
#include "internal.cuh"

namespace cmb {

__global__ void grey_flux_kernel(int nz, int64_t ny, double sigma, const double *__restrict__ T,
                                 const double *__restrict__ dtau, const double *__restrict__ toa_up_lw,
                                 const double *__restrict__ up_sw0, const double *__restrict__ down_sw0,
                                 const double *__restrict__ tau_sw, double *__restrict__ up_lw,
                                 double *__restrict__ down_lw, double *__restrict__ up_sw,
                                 double *__restrict__ down_sw, double *__restrict__ net) {
    const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (j >= ny) return;
    auto sw = [&](int i, double &u, double &d) {
        u = up_sw0 ? up_sw0[j] : 0.0;
        d = down_sw0 ? down_sw0[j] : 0.0;
        if (tau_sw) {                                   // Beer's law on both streams, :292-293
            const double ts = tau_sw[(int64_t)i * ny + j];
            u = __dmul_rn(u, exp(ts));
            d = __dmul_rn(d, exp(-ts));
        }
        if (up_sw) up_sw[(int64_t)i * ny + j] = u;
        if (down_sw) down_sw[(int64_t)i * ny + j] = d;
    };
    double up = toa_up_lw[j], down = 0.0, us, ds;       // top: up balances the absorbed stellar flux, down = 0
    const int64_t top = (int64_t)(nz - 1) * ny + j;
    up_lw[top] = up;
    down_lw[top] = down;
    sw(nz - 1, us, ds);
    if (net) net[top] = __dsub_rn(__dadd_rn(__dsub_rn(up, down), us), ds);
    for (int i = nz - 2; i >= 0; --i) {
        const int64_t at = (int64_t)i * ny + j;
        const double dt = dtau[at], source = __dmul_rn(sigma, pow(T[at], 4.0));
        const double ep = exp(dt), em = exp(-dt);
        up = __dadd_rn(__dmul_rn(up, ep), __dmul_rn(source, __dsub_rn(1.0, ep)));        // :268-270
        down = __dadd_rn(__dmul_rn(down, em), __dmul_rn(source, __dsub_rn(1.0, em)));    // :271-273
        up_lw[at] = up;
        down_lw[at] = down;
        sw(i, us, ds);
        if (net) net[at] = __dsub_rn(__dadd_rn(__dsub_rn(up, down), us), ds);
    }
}

}  // namespace cmb

using namespace cmb;

extern "C" int cmb200_grey_flux_host(int32_t nz, int64_t ny, double sigma, const double *T, const double *dtau,
                                     const double *toa_up_lw, const double *up_sw0, const double *down_sw0,
                                     const double *tau_sw, double *up_lw, double *down_lw, double *up_sw,
                                     double *down_sw, double *net) {
    CMB_CHECK_ARG(T && dtau && toa_up_lw && up_lw && down_lw, "null pointer");
    CMB_CHECK_ARG(nz >= 2 && ny >= 0, "need at least two interfaces and a non-negative column count");
    CMB_CHECK_ARG(!(up_sw || down_sw) || (up_sw0 && down_sw0), "short-wave outputs need their top-of-atmosphere values");
    if (ny == 0) return CMB200_OK;
    const size_t nc = (size_t)(nz - 1) * ny, ni = (size_t)nz * ny;
    DevBuf<double> dT, ddtau, dtoa, du0, dd0, dts, dul, ddl, dus, dds, dnet;
    CMB_CUDA(dT.upload(T, nc));
    CMB_CUDA(ddtau.upload(dtau, nc));
    CMB_CUDA(dtoa.upload(toa_up_lw, ny));
    if (up_sw0) CMB_CUDA(du0.upload(up_sw0, ny));
    if (down_sw0) CMB_CUDA(dd0.upload(down_sw0, ny));
    if (tau_sw) CMB_CUDA(dts.upload(tau_sw, ni));
    CMB_CUDA(dul.alloc(ni));
    CMB_CUDA(ddl.alloc(ni));
    if (up_sw) CMB_CUDA(dus.alloc(ni));
    if (down_sw) CMB_CUDA(dds.alloc(ni));
    if (net) CMB_CUDA(dnet.alloc(ni));
    grey_flux_kernel<<<(unsigned)((ny + 127) / 128), 128>>>(nz, ny, sigma, dT.p, ddtau.p, dtoa.p, du0.p, dd0.p, dts.p,
                                                            dul.p, ddl.p, dus.p, dds.p, dnet.p);
    CMB_LAUNCHED();
    CMB_CUDA(cudaMemcpy(up_lw, dul.p, ni * sizeof(double), cudaMemcpyDeviceToHost));
    CMB_CUDA(cudaMemcpy(down_lw, ddl.p, ni * sizeof(double), cudaMemcpyDeviceToHost));
    if (up_sw) CMB_CUDA(cudaMemcpy(up_sw, dus.p, ni * sizeof(double), cudaMemcpyDeviceToHost));
    if (down_sw) CMB_CUDA(cudaMemcpy(down_sw, dds.p, ni * sizeof(double), cudaMemcpyDeviceToHost));
    if (net) CMB_CUDA(cudaMemcpy(net, dnet.p, ni * sizeof(double), cudaMemcpyDeviceToHost));
    return CMB200_OK;
}
