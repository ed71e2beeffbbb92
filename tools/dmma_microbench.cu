// Does the FP64 tensor path (mma.sync m8n8k4 f64) run beside the FP64 vector pipe on this GPU, and
// at what rate?  Decides whether the flux kernel's warp reduction can be moved onto it.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/dmma_microbench tools/dmma_microbench.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

template <int N_MMA, int N_FMA>
__global__ void mix(double *out, int iters, long long *cycles) {
    double c[4][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};
    double x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = 1.0 + 1e-9 * (threadIdx.x + i);
    const double a = 1.0 + 1e-12 * threadIdx.x, b = 0.5, fa = 0.999999, fb = 1e-7;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < N_MMA; ++i) dmma(c[i % 4][0], c[i % 4][1], a, b);
#pragma unroll
            for (int i = 0; i < N_FMA; ++i) x[i % 8] = fma(x[i % 8], fa, fb);
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) s += c[i][0] + c[i][1];
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int N_MMA, int N_FMA>
void run(int warps_per_smsp, double *out, long long *dcyc) {
    const int iters = 1000;
    for (int rep = 0; rep < 2; ++rep) {
        mix<N_MMA, N_FMA><<<148, warps_per_smsp * 4 * 32>>>(out, iters, dcyc);
        cudaDeviceSynchronize();
    }
    long long c;
    cudaMemcpy(&c, dcyc, sizeof(c), cudaMemcpyDeviceToHost);
    const double per_group = (double)c / (iters * 4);
    printf("  %d warp(s)/scheduler, group of %2d DMMA + %2d DFMA: %.1f cycles per group per warp\n", warps_per_smsp, N_MMA,
           N_FMA, per_group);
}

int main() {
    double *out;
    long long *dcyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(double));
    cudaMalloc(&dcyc, sizeof(long long));
    for (int w = 1; w <= 2; ++w) {
        run<8, 0>(w, out, dcyc);
        run<0, 32>(w, out, dcyc);
        run<8, 32>(w, out, dcyc);
        run<8, 64>(w, out, dcyc);
        run<1, 0>(w, out, dcyc);
    }
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
