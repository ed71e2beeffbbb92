// Issue-mix microbenchmark, round 2: what one extra instruction of a given class costs next to the FP64
// stream of the flux kernel (8 independent DFMA chains per warp, 2 warps per scheduler, one CTA of 256
// threads per SM).  Every extra is written as inline PTX so that the SASS opcode is the one named
// (check with cuobjdump -sass).  Stand-alone:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/issue_mix2_mb tools/issue_mix2_microbench.cu
#include <cstdio>
#include <cuda_runtime.h>

enum Extra { NONE, LOP3, IADD3_CC, SHF, IMAD, IMAD_WIDE, MOV, SEL, VIMNMX, PRMT, LDS64, LDS128B, SHFL, LDG64, FFMA, FMUL, LEA, IMAD_SHL, IADD_PLAIN, DADD_X, F2F };

template <int EXTRA, int PER>   // PER extras per DFMA (1 or 2); EXTRA after every DFMA
__global__ void __launch_bounds__(256, 1) mix_kernel(double *out, int iters, long long *cycles, double a, double b, int ia, int ib,
                                                     const double *g) {
    __shared__ __align__(16) double sh[512];
    constexpr int K = 8;
    double x[K];
    int y[K], z[K];
    float f[K];
#pragma unroll
    for (int i = 0; i < K; ++i) {
        x[i] = 1.0 + 1e-9 * (threadIdx.x + i);
        y[i] = threadIdx.x + i;
        z[i] = threadIdx.x * 3 + i;
        f[i] = 1.0f + i;
    }
    for (int i = threadIdx.x; i < 512; i += blockDim.x) sh[i] = 1e-9 * i;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const unsigned shb = (unsigned)__cvta_generic_to_shared(sh);
    const double *gl = g + threadIdx.x;
    double dacc = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < K; ++i) {
                x[i] = fma(x[i], a, b);
#pragma unroll
                for (int e = 0; e < PER; ++e) {
                    int &w = e ? z[i] : y[i];
                    if (EXTRA == LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(w) : "r"(ia), "r"(ib));
                    const int other = e ? y[(i + r + 1) & 7] : z[(i + r + 1) & 7];     // a different register every time: no strength reduction
                    if (EXTRA == IADD3_CC) asm volatile("add.cc.s32 %0, %0, %1;" : "+r"(w) : "r"(other));
                    if (EXTRA == IADD_PLAIN) asm volatile("add.s32 %0, %0, %1;" : "+r"(w) : "r"(other));
                    if (EXTRA == SHF) asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(w) : "r"(other));
                    if (EXTRA == IMAD) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(w) : "r"(ia), "r"(other));
                    if (EXTRA == IMAD_SHL) asm volatile("{.reg .b32 t; shl.b32 t, %0, 4; xor.b32 %0, t, %1;}" : "+r"(w) : "r"(other));
                    if (EXTRA == LEA) asm volatile("{.reg .b32 t; shl.b32 t, %0, 3; add.s32 %0, t, %1;}" : "+r"(w) : "r"(other));
                    if (EXTRA == IMAD_WIDE) {
                        unsigned lo, hi;
                        asm volatile("{.reg .b64 q; mad.wide.u32 q, %2, 16, %3; mov.b64 {%0, %1}, q;}" : "=r"(lo), "=r"(hi) : "r"(w), "l"((unsigned long long)g));
                        w = lo ^ hi;
                    }
                    if (EXTRA == MOV) asm volatile("{.reg .b32 t; mov.b32 t, %1; xor.b32 %0, %0, t;}" : "+r"(w) : "r"(other));
                    if (EXTRA == SEL) asm volatile("{.reg .pred p; setp.ne.b32 p, %2, 0; selp.b32 %0, %0, %1, p;}" : "+r"(w) : "r"(ib), "r"(lane));
                    if (EXTRA == VIMNMX) asm volatile("min.s32 %0, %0, %1;" : "+r"(w) : "r"(other));
                    if (EXTRA == PRMT) asm volatile("prmt.b32 %0, %0, %1, 0x5210;" : "+r"(w) : "r"(other));
                    if (EXTRA == LDS64) {
                        double v;
                        asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(shb + ((w & 63) * 64 + (lane & 15) * 8) % 4096));
                        dacc += 0;   // keep v alive through an integer use only
                        w ^= __double2loint(v);
                    }
                    if (EXTRA == LDS128B) {
                        int4 v;
                        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(shb + ((i + r) & 31) * 16));
                        w ^= v.x;
                    }
                    if (EXTRA == SHFL) w = __shfl_xor_sync(0xffffffffu, w, 1 + (r & 3));
                    if (EXTRA == LDG64) {
                        double v;
                        asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(gl + ((it * 32 + r * 8 + i) & 1023) * 256));
                        w ^= __double2loint(v);
                    }
                    if (EXTRA == FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"((float)a), "f"((float)b));
                    if (EXTRA == FMUL) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f[i]) : "f"((float)a));
                    if (EXTRA == F2F) {
                        float q;
                        asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(q) : "d"(x[(i + 4) & 7]));
                        f[i] += q;
                    }
                    if (EXTRA == DADD_X) dacc += x[(i + 4) & 7];
                }
            }
        }
    }
    const long long t1 = clock64();
    double s = dacc;
#pragma unroll
    for (int i = 0; i < K; ++i) s += x[i] + y[i] + z[i] + f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int EXTRA, int PER>
void run(const char *name, double *out, long long *dcyc, const double *g) {
    const int iters = 1000;
    for (int rep = 0; rep < 2; ++rep) {
        mix_kernel<EXTRA, PER><<<148, 256>>>(out, iters, dcyc, 0.999999, 1e-7, 3, 7, g);
        cudaDeviceSynchronize();
    }
    long long c;
    cudaMemcpy(&c, dcyc, sizeof(c), cudaMemcpyDeviceToHost);
    const double n = (double)iters * 4 * 8;                 // DFMA per warp
    printf("  DFMA + %d x %-26s %.2f cycles per DFMA per scheduler (2 warps)   %s\n", PER, name, c / n / 2, cudaGetErrorString(cudaGetLastError()));
}

#define BOTH(E, NAME) run<E, 1>(NAME, out, dcyc, g); run<E, 2>(NAME, out, dcyc, g);

int main() {
    double *out, *g;
    long long *dcyc;
    cudaMalloc(&out, 148 * 256 * sizeof(double));
    cudaMalloc(&g, (1024 * 256 + 256) * sizeof(double));
    cudaMemset(g, 0, (1024 * 256 + 256) * sizeof(double));
    cudaMalloc(&dcyc, sizeof(long long));
    run<NONE, 1>("nothing", out, dcyc, g);
    BOTH(LOP3, "LOP3 (lop3.b32)")
    BOTH(IADD3_CC, "IADD3 (add.cc.s32)")
    BOTH(IADD_PLAIN, "add.s32")
    BOTH(SHF, "SHF (shf.l.wrap)")
    BOTH(IMAD_SHL, "shl.b32")
    BOTH(LEA, "shl + add (LEA?)")
    BOTH(IMAD, "IMAD (mad.lo.s32)")
    BOTH(IMAD_WIDE, "IMAD.WIDE (mad.wide.u32)")
    BOTH(MOV, "mov.b32")
    BOTH(SEL, "SEL (selp.b32)")
    BOTH(VIMNMX, "VIMNMX (min.s32)")
    BOTH(PRMT, "PRMT")
    BOTH(LDS64, "LDS.64 conflict-free table")
    BOTH(LDS128B, "LDS.128 broadcast")
    BOTH(SHFL, "SHFL.BFLY")
    BOTH(LDG64, "LDG.E.64 (L2)")
    BOTH(FFMA, "FFMA")
    BOTH(FMUL, "FMUL")
    BOTH(F2F, "F2F.F32.F64 (+FADD)")
    BOTH(DADD_X, "DADD (for scale)")
    return 0;
}
