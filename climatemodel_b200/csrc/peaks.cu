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

// The same with the multiplier and addend in registers (loaded from memory, so that ptxas cannot fold them
// into constant-bank operands, which cost the FP64 pipe ~10 % of its issue rate on this part).
template <typename Real, int ILP>
__global__ void fma_peak_reg_kernel(Real *out, int iters, const Real *coef) {
    Real x[ILP];
    const Real a = coef[0], b = coef[1];
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = (Real)(threadIdx.x + i) * (Real)1e-3;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) x[i] = fma(x[i], a, b);
    }
    Real s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i];
    if (s == (Real)123456.789) out[0] = s;
}

// L2 -> SM read bandwidth: an L2-resident buffer (64 MB of the 126 MB) read over and over with coalesced
// 16-byte loads, eight independent loads in flight per thread, at full occupancy -- the ceiling of the flux
// kernel's table stream (its rows never leave L2 while columns stream past).
__global__ void __launch_bounds__(512) l2_stream_kernel(const double2 *__restrict__ buf, int64_t n, int passes, double *out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    double acc = 0.0;
    for (int p = 0; p < passes; ++p) {
        // another CTA's lines every pass, so that no pass finds its lines in this SM's L1
        int64_t i = (int64_t)((blockIdx.x + 41 * p) % gridDim.x) * blockDim.x + threadIdx.x;
        for (; i + 7 * stride < n; i += 8 * stride) {
            double2 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = __ldcg(buf + i + k * stride);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc += v[k].x + v[k].y;
        }
    }
    if (acc == 123456.789) out[0] = acc;
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

template <typename Real>
static int time_fma_reg(int device, double *tflops) {
    constexpr int ILP = 8;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    DevBuf<Real> out, coef;
    CMB_CUDA(out.alloc(1));
    const Real h[2] = {(Real)0.999, (Real)1e-3};
    CMB_CUDA(coef.upload(h, 2));
    const int threads = 256, blocks = sms, iters = 1 << 15;       // the flux kernel's occupancy: 2 warps per scheduler
    cudaEvent_t e0, e1;
    CMB_CUDA(cudaEventCreate(&e0));
    CMB_CUDA(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {
        CMB_CUDA(cudaEventRecord(e0));
        fma_peak_reg_kernel<Real, ILP><<<blocks, threads>>>(out.p, iters, coef.p);
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

static int time_l2_stream(int device, double *gbs) {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const int64_t n = (int64_t)4 << 20;                  // 4 Mi double2 = 64 MB
    DevBuf<double2> buf;
    DevBuf<double> out;
    CMB_CUDA(buf.alloc((size_t)n));
    CMB_CUDA(out.alloc(1));
    CMB_CUDA(cudaMemset(buf.p, 0, (size_t)n * sizeof(double2)));
    const int passes = 24;
    cudaEvent_t e0, e1;
    CMB_CUDA(cudaEventCreate(&e0));
    CMB_CUDA(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        CMB_CUDA(cudaEventRecord(e0));
        l2_stream_kernel<<<sms * 4, 512>>>(buf.p, n, passes, out.p);
        CMB_LAUNCHED();
        CMB_CUDA(cudaEventRecord(e1));
        CMB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        CMB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const int64_t per_pass = n / ((int64_t)sms * 4 * 512 * 8) * ((int64_t)sms * 4 * 512 * 8);
        const double bytes = (double)passes * per_pass * sizeof(double2);
        if (rep > 0) best = std::max(best, bytes / (ms * 1e-3) / 1e9);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *gbs = best;
    return CMB200_OK;
}

}  // namespace cmb

using namespace cmb;

extern "C" int cmb200_measure_peaks2(int device, double *fp64_reg_tflops, double *fp32_reg_tflops, double *l2_stream_gbs) {
    CMB_CUDA(cudaSetDevice(device));
    int rc;
    if (fp64_reg_tflops && (rc = time_fma_reg<double>(device, fp64_reg_tflops))) return rc;
    if (fp32_reg_tflops && (rc = time_fma_reg<float>(device, fp32_reg_tflops))) return rc;
    if (l2_stream_gbs && (rc = time_l2_stream(device, l2_stream_gbs))) return rc;
    return CMB200_OK;
}

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
