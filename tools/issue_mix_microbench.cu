// Issue-mix microbenchmark behind the flux kernel's roofline: how many cycles a scheduler needs
// per DFMA when the warp's instruction stream interleaves other instruction classes with the FP64
// work (8 independent DFMA chains per warp, W warps per scheduler, as in the kernel).  Stand-alone:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/issue_mix tools/issue_mix_microbench.cu
#include <cstdio>
#include <cuda_runtime.h>

__constant__ double kC[4] = {0.999999, 1e-7, 1.000001, -1e-7};

enum Mix { PURE, CONSTOP, IMAD1, ALU1, MOV1, SHFL4, LDS4, DMUL_ONLY, DADD_ONLY, IMAD2, ALU2, ALU1_IMAD1 };

template <int MIX>
__global__ void mix_kernel(double *out, int iters, long long *cycles, double a, double b, int ia, int ib) {
    __shared__ double sh[64 * 8];
    constexpr int K = 8;
    double x[K];
    int y[K], z[K];
#pragma unroll
    for (int i = 0; i < K; ++i) {
        x[i] = 1.0 + 1e-9 * (threadIdx.x + i);
        y[i] = threadIdx.x + i;
        z[i] = threadIdx.x * 3 + i;
    }
    for (int i = threadIdx.x; i < 64 * 8; i += blockDim.x) sh[i] = 1e-9 * i;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < K; ++i) {
                if (MIX == DMUL_ONLY)
                    x[i] = x[i] * a;
                else if (MIX == DADD_ONLY)
                    x[i] = x[i] + b;
                else if (MIX == CONSTOP)
                    x[i] = fma(x[i], kC[r & 1 ? 0 : 2], kC[r & 1 ? 1 : 3]);
                else
                    x[i] = fma(x[i], a, b);
                if (MIX == IMAD1 || MIX == IMAD2 || MIX == ALU1_IMAD1) y[i] = y[i] * ia + ib;
                if (MIX == IMAD2) z[i] = z[i] * ib + ia;
                if (MIX == ALU1 || MIX == ALU2 || MIX == ALU1_IMAD1) asm volatile("xor.b32 %0, %0, %1;" : "+r"(z[i]) : "r"(ia));
                if (MIX == ALU2) asm volatile("and.b32 %0, %0, %1;" : "+r"(y[i]) : "r"(z[i]));
                if (MIX == MOV1) asm volatile("{.reg .pred p; setp.ne.b32 p, %2, 0; selp.b32 %0, %0, %1, p;}" : "+r"(z[i]) : "r"(ib), "r"(lane));
                if (MIX == SHFL4 && (i & 3) == 0) y[i] = __shfl_xor_sync(0xffffffffu, y[i], 1 + (r & 3));
                if (MIX == LDS4 && (i & 3) == 0) x[i] += ((volatile double *)sh)[((i + r) * 32 + lane) & 511];
            }
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < K; ++i) s += x[i] + y[i] + z[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int MIX>
void run(const char *name, int wps, double *out, long long *dcyc) {
    const int iters = 2000;
    for (int rep = 0; rep < 2; ++rep) {
        mix_kernel<MIX><<<148, wps * 4 * 32>>>(out, iters, dcyc, 0.999999, 1e-7, 3, 7);
        cudaDeviceSynchronize();
    }
    long long c;
    cudaMemcpy(&c, dcyc, sizeof(c), cudaMemcpyDeviceToHost);
    const double n = (double)iters * 4 * 8;                 // FP64 instructions per warp
    printf("  %-34s %d warp(s)/scheduler: %.2f cycles per FP64 instruction per scheduler\n", name, wps, c / n / wps);
}

int main() {
    double *out;
    long long *dcyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(double));
    cudaMalloc(&dcyc, sizeof(long long));
    for (int w = 2; w <= 3; ++w) {
        run<PURE>("DFMA only", w, out, dcyc);
        run<CONSTOP>("DFMA, constant-bank operands", w, out, dcyc);
        run<DMUL_ONLY>("DMUL only", w, out, dcyc);
        run<DADD_ONLY>("DADD only", w, out, dcyc);
        run<IMAD1>("DFMA + 1 IMAD", w, out, dcyc);
        run<IMAD2>("DFMA + 2 IMAD", w, out, dcyc);
        run<ALU1>("DFMA + 1 ALU (LOP3)", w, out, dcyc);
        run<ALU2>("DFMA + 2 ALU (LOP3)", w, out, dcyc);
        run<ALU1_IMAD1>("DFMA + IMAD + ALU", w, out, dcyc);
        run<MOV1>("DFMA + ISETP + SEL", w, out, dcyc);
        run<SHFL4>("DFMA + 1/4 SHFL", w, out, dcyc);
        run<LDS4>("DFMA + 1/4 (LDS.64 + DADD)", w, out, dcyc);
    }
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
