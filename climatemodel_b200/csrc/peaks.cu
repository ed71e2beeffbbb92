// Issue-rate microbenchmarks used as roofline denominators for the pipe-bound regime
// (MEASURED_PEAKS.json only carries HBM copy bandwidth and bf16 tensor throughput; the flux
// kernel is bound by the FP64 -- or, in FP32 mode, the FP32/SFU -- pipe).
#include "internal.cuh"

namespace cmb {

template <typename Real, int ILP>
__global__ void fma_peak_kernel(Real *out, int iters, Real a, Real b) {
    Real x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = (Real)(threadIdx.x + i) * (Real)1e-3;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) x[i] = fma(x[i], a, b);
    }
    Real s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i];
    if (s == (Real)123456.789) out[0] = s;  // keep the loop alive
}

__global__ void copy_kernel(const double4 *__restrict__ in, double4 *__restrict__ out, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = in[i];
}

template <typename Real>
static int time_fma(int device, double *tflops) {
    constexpr int ILP = 8;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    DevBuf<Real> out;
    CMB_CUDA(out.alloc(1));
    const int threads = 512, blocks = sms * 4, iters = 4096;
    cudaEvent_t e0, e1;
    CMB_CUDA(cudaEventCreate(&e0));
    CMB_CUDA(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        CMB_CUDA(cudaEventRecord(e0));
        fma_peak_kernel<Real, ILP><<<blocks, threads>>>(out.p, iters, (Real)0.999, (Real)1e-3);
        CMB_LAUNCHED();
        CMB_CUDA(cudaEventRecord(e1));
        CMB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        CMB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        double fl = 2.0 * ILP * (double)iters * threads * blocks;
        if (rep > 0) best = std::max(best, fl / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tflops = best;
    return CMB200_OK;
}

}  // namespace cmb

using namespace cmb;

extern "C" int cmb200_measure_peaks(int device, double *fp64_tflops, double *fp32_tflops,
                                    double *copy_gbs) {
    CMB_CUDA(cudaSetDevice(device));
    int rc;
    if (fp64_tflops && (rc = time_fma<double>(device, fp64_tflops))) return rc;
    if (fp32_tflops && (rc = time_fma<float>(device, fp32_tflops))) return rc;
    if (copy_gbs) {
        const int64_t n = (int64_t)1 << 25;  // 32 Mi double4 = 1 GiB per buffer (> L2)
        DevBuf<double4> a, b;
        CMB_CUDA(a.alloc(n));
        CMB_CUDA(b.alloc(n));
        CMB_CUDA(cudaMemset(a.p, 0, n * sizeof(double4)));
        cudaEvent_t e0, e1;
        CMB_CUDA(cudaEventCreate(&e0));
        CMB_CUDA(cudaEventCreate(&e1));
        double best = 0.0;
        for (int rep = 0; rep < 5; ++rep) {
            CMB_CUDA(cudaEventRecord(e0));
            copy_kernel<<<148 * 16, 512>>>(a.p, b.p, n);
            CMB_LAUNCHED();
            CMB_CUDA(cudaEventRecord(e1));
            CMB_CUDA(cudaEventSynchronize(e1));
            float ms = 0;
            CMB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            if (rep > 0) best = std::max(best, 2.0 * n * sizeof(double4) / (ms * 1e-3) / 1e9);
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        *copy_gbs = best;
    }
    return CMB200_OK;
}
