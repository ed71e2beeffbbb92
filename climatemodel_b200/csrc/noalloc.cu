// Allocation-free variant of cmb200_plan_optical_depth (A3 on the plan grid, real_gas.py:86-127).
//
// No counterpart in the reference: added for callers that cannot
// call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses that forbid it.
//
// cmb200_plan_optical_depth keeps one piece of scratch: tidx[M][n_col][L], the table-temperature row
// every (molecule, column, level) reads (written by tidx_kernel, read by plan_optical_depth_kernel).  The
// regular entry point grows that buffer on the plan with cudaMalloc when a call needs more than the last.
// Here the scratch is a file-scope device array of a fixed, stated size instead: every GPU that loads the
// library holds kNoAllocScratchBytes (1 MiB) until the process exits, and the entry point calls no
// cudaMalloc / cudaFree / allocating library routine -- only cudaSetDevice, two kernel launches per
// molecule count, and (once per device) cudaGetSymbolAddress.
//
// Sharing: every call on a device uses the SAME array, which keeps the last call's contents.  Nothing in
// it has to start at zero -- tidx_kernel overwrites all M * n_col * L entries a call reads before
// plan_optical_depth_kernel reads them, in stream order -- so nothing is cleared.  Two calls whose kernels
// overlap on the GPU (different streams without an event between them, or concurrent graph launches) RACE
// on the array: serialise them.
#include <mutex>

#include "internal.cuh"

namespace cmb {

constexpr size_t kNoAllocScratchBytes = (size_t)1 << 20;      // M * n_col * L <= 1,048,576 (e.g. 2 x 5242 x 100)
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) uint8_t g_noalloc_tidx[kNoAllocScratchBytes];

// In host code the array's name is a host-side stand-in, not a device address: the address a kernel can
// use comes from cudaGetSymbolAddress, per device (each device has its own copy), with that device current.
static uint8_t *g_noalloc_addr[kNoAllocMaxDevices] = {};
static std::mutex g_noalloc_mutex;

static int noalloc_scratch(int device, uint8_t **out) {
    CMB_CHECK_ARG(device >= 0 && device < kNoAllocMaxDevices, "device %d beyond the %d this entry point tracks", device,
                  kNoAllocMaxDevices);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_addr[device]) {
        void *p = nullptr;
        CMB_CUDA(cudaGetSymbolAddress(&p, g_noalloc_tidx));      // the caller made `device` current
        if (!p) {
            set_error("cudaGetSymbolAddress returned a null address for the scratch array on device %d", device);
            return CMB200_ERR_CUDA;
        }
        g_noalloc_addr[device] = (uint8_t *)p;
    }
    *out = g_noalloc_addr[device];
    return CMB200_OK;
}

}  // namespace cmb

using namespace cmb;

extern "C" int cmb200_plan_optical_depth_noalloc(cmb200_plan *plan, int64_t n_col, const double *T_tau,
                                                 const double *T_g, const double *q, int q_per_column,
                                                 double *tau, void *stream) {
    CMB_CHECK_ARG(plan && q && tau, "null pointer");
    CMB_CHECK_ARG(T_tau || T_g, "either T_tau or T_g is required");
    CMB_CHECK_ARG(n_col >= 0, "negative column count");
    if (plan->M == 0) {
        set_error("the plan has no absorption tables (cmb200_plan_set_tables)");
        return CMB200_ERR_STATE;
    }
    if (n_col == 0) return CMB200_OK;
    CMB_CHECK_ARG(n_col <= 65535, "at most 65535 columns per optical-depth call");
    const size_t need = (size_t)plan->M * (size_t)n_col * (size_t)plan->L;
    CMB_CHECK_ARG(need <= kNoAllocScratchBytes,
                  "cmb200_plan_optical_depth_noalloc: molecules x columns x levels = %d x %lld x %d = %zu exceeds the fixed "
                  "scratch of %zu entries; split the columns over several calls or use cmb200_plan_optical_depth",
                  plan->M, (long long)n_col, plan->L, need, kNoAllocScratchBytes);
    CMB_CUDA(cudaSetDevice(plan->device));
    uint8_t *tidx = nullptr;
    int rc = noalloc_scratch(plan->device, &tidx);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t n = n_col * plan->L;
    for (int m = 0; m < plan->M; ++m) {
        tidx_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(T_tau, T_g, n_col, plan->L,
                                                               plan->T_axes.p + (size_t)m * plan->nT_pad, plan->nT_pad,
                                                               tidx + (size_t)m * n);
        CMB_LAUNCHED();
    }
    dim3 grid((unsigned)((plan->N + 127) / 128), (unsigned)n_col);
    if (plan->precision == CMB200_F64)
        plan_optical_depth_kernel<double><<<grid, 128, 0, s>>>(plan->KL64.p, plan->M, plan->L, plan->nT_pad, plan->N,
                                                              plan->pitch, tidx, q, q_per_column ? plan->L : 0, n_col,
                                                              plan->dx.p, plan->c.g, tau);
    else
        plan_optical_depth_kernel<float><<<grid, 128, 0, s>>>(plan->KL32.p, plan->M, plan->L, plan->nT_pad, plan->N,
                                                             plan->pitch, tidx, q, q_per_column ? plan->L : 0, n_col,
                                                             plan->dx.p, plan->c.g, tau);
    CMB_LAUNCHED();
    return CMB200_OK;
}
