// FP64 pipe microbenchmark for the flux kernel's design: dependent-DFMA latency and the issue
// rate a scheduler reaches with W warps of K independent chains each.  Stand-alone:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fp64_microbench tools/fp64_microbench.cu
//   build/fp64_microbench            (on the GPU box)
#include <cstdio>
#include <cuda_runtime.h>

template <int K>
__global__ void chains(double *out, int iters, long long *cycles) {
    double x[K];
#pragma unroll
    for (int i = 0; i < K; ++i) x[i] = 1.0 + 1e-9 * (threadIdx.x + i);
    const double a = 0.999999, b = 1e-7;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int i = 0; i < K; ++i) x[i] = fma(x[i], a, b);
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < K; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int K>
void run(int warps_per_smsp, double *out, long long *dcyc) {
    const int iters = 2000;
    chains<K><<<148, warps_per_smsp * 4 * 32>>>(out, iters, dcyc);
    cudaDeviceSynchronize();
    chains<K><<<148, warps_per_smsp * 4 * 32>>>(out, iters, dcyc);
    cudaDeviceSynchronize();
    long long c;
    cudaMemcpy(&c, dcyc, sizeof(c), cudaMemcpyDeviceToHost);
    const double n = (double)iters * 8 * K;                 // DFMA per warp
    printf("  K=%2d chains, %d warp(s)/scheduler: %.2f cycles per DFMA per warp, %.3f DFMA/clk/scheduler\n", K,
           warps_per_smsp, c / n, n * warps_per_smsp / c);
}

int main() {
    double *out;
    long long *dcyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(double));
    cudaMalloc(&dcyc, sizeof(long long));
    for (int w = 1; w <= 4; w *= 2) {
        printf("warps per scheduler = %d\n", w);
        run<1>(w, out, dcyc);
        run<2>(w, out, dcyc);
        run<4>(w, out, dcyc);
        run<8>(w, out, dcyc);
        run<16>(w, out, dcyc);
    }
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
