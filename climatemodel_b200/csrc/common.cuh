// Shared helpers of the climatemodel_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <atomic>

#include "../../include/climatemodel_b200.h"

namespace cmb {

// ---- error plumbing ---------------------------------------------------------------------
void set_error(const char *fmt, ...);
extern std::atomic<long long> g_launches;

#define CMB_CHECK_ARG(cond, ...)                          \
    do {                                                  \
        if (!(cond)) {                                    \
            cmb::set_error(__VA_ARGS__);                  \
            return CMB200_ERR_ARG;                        \
        }                                                 \
    } while (0)

#define CMB_CUDA(expr)                                                                     \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            cmb::set_error("CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__,   \
                           __LINE__, cudaGetErrorString(_e));                              \
            return CMB200_ERR_CUDA;                                                        \
        }                                                                                  \
    } while (0)

#define CMB_LAUNCHED()                     \
    do {                                   \
        cmb::g_launches.fetch_add(1);      \
        CMB_CUDA(cudaGetLastError());      \
    } while (0)

// RAII device buffer (host-side temporaries of the *_host entry points)
template <typename T>
struct DevBuf {
    T *p = nullptr;
    size_t n = 0;
    DevBuf() {}
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
    cudaError_t alloc(size_t count) {
        release();
        n = count;
        if (count == 0) return cudaSuccess;
        return cudaMalloc((void **)&p, count * sizeof(T));
    }
    cudaError_t upload(const T *h, size_t count, cudaStream_t s = 0) {
        cudaError_t e = alloc(count);
        if (e != cudaSuccess || count == 0) return e;
        return cudaMemcpyAsync(p, h, count * sizeof(T), cudaMemcpyHostToDevice, s);
    }
};

// ---- device helpers ----------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// pi * B_wavenumber(nu, T): the flux-density source of real_gas.py:638-641 / :585-600.
// Operation order follows B_freq (real_gas.py:26-27): u = h f / (k T);
// (2 h f^3 / c^2) / (exp(u) - 1), times dfreq/dnu = 100 c.  exp(u) overflows to +inf for
// short-wave band centres at atmospheric temperatures, giving 0 exactly like NumPy.
struct PlanckConst {
    double h, c, kB;
};
__device__ __forceinline__ double planck_wavenumber(const PlanckConst &pc, double nu, double T) {
    double f = 100.0 * nu * pc.c;
    double u = pc.h * f / (pc.kB * T);
    double num = 2.0 * pc.h * (f * f * f) / (pc.c * pc.c);
    return (100.0 * pc.c) * (num / (exp(u) - 1.0));
}

}  // namespace cmb
