// Stages 2+3 of the hot path: layer transmissivity, Planck source and the two flux sweeps
// (RealGas.get_flux / flux_integrals, real_gas.py:564-669), net flux (:712-713) and heating
// rate (base.py:166-168), optionally with stage 1 (optical depth, :86-127) fused in front.
//
// Per-bin form (oracle/realgas_oracle.py:get_flux_per_bin): every quantity the
// reference builds from the band-averaged transmission matrix is linear in the per-bin
// products E(i,k) = prod_{l=i}^{k-1} t_l with t_l = exp(-dtau_l), so each (band, bin) "chain"
// is one exp per layer plus two first-order recurrences over the layers (one per direction), and
// the band value is the trapezoid-weighted sum of the chains.
//
// Mapping: one CTA owns one column at a time (or 1/C of it when there are few columns); the
// per-level lookup data of the column sit once in shared memory and each of the CTA's W warps
// takes a contiguous, cost-balanced run of "segments" (32 consecutive wavenumber bins of one
// band, one chain per lane, so table / tau rows are read coalesced).
//
// A warp walks its segments in *passes* over the levels that alternate direction.  Each pass
// carries two chains at once: the FRESH chain of segment s, whose t_k = exp(-dtau_k) is computed
// now and parked in shared memory ts[k][lane], and the REPLAY chain of segment s-1, whose t_k is
// read back from ts (just before the slot is overwritten) to run its sweep in the other
// direction.  So a layer costs one exp, every pass has two independent recurrences to overlap,
// and because both chains of a pass contribute to the same band, direction and interface, their
// values are added *before* the warp reduction -- one reduction per pass instead of two.
//   pass n   (down):  fresh seg n   downward sweep  |  replay seg n-1 downward sweep
//   pass n+1 (up):    fresh seg n+1 upward sweep    |  replay seg n   upward sweep      ...
// A band run ends with a replay-only pass.  Per-(warp, band) partial sums go to a small global
// scratch (L2 resident) and are added in a fixed order, so results do not depend on scheduling.
#pragma once
#include <cstring>
#include "internal.cuh"
#include "tmem.cuh"

namespace cmb {

template <typename Real>
struct FluxArgs {
    int L, B, M, nT_pad, n_slots, band_lo, band_hi;
    int ctas_per_col;          // C: CTAs sharing one column (1 = whole column in one CTA, in-CTA finalize)
    int64_t N, pitch, n_col, n_tasks;
    // plan-static
    const double *geo, *Pint, *hdP;
    const int32_t *band_start, *band_len;
    const double *band_centre, *band_delta, *band_sumw;
    const uint8_t *band_sw;
    const Real *wchain, *seg_w;   // nu-weights per chain; the same per (segment, lane)
    const int32_t *seg_bin;       // wavenumber bin per (segment, lane)
    const int64_t *chain_off;
    const int32_t *seg_band, *seg_c0, *part_seg0, *part_slot0, *band_slot0;
    const double *slot_delta;  // band width of the band each partial-sum slot belongs to
    PlanckConst pc;
    double T_star, star_geom, g, g_over_cp;
    // per column
    const double *T, *Tint, *T_g, *albedo;
    double albedo_scalar;
    const double *tau;
    const int32_t *tau_id;
    const Real *KL;
    const uint8_t *tidx;
    const double *q;
    int64_t q_mol_stride, q_col_stride;
    double *up, *down, *net, *heating;
    // work space (written and re-read inside the kernel: never const / restrict)
    double *scratch;           // [region][n_slots][2][L] partial band sums
    Real *coefbuf;             // [CTA][B][4][Lp] emission coefficients of the long-wave bands
    unsigned long long *task_counter;   // zeroed before the launch
    long long *dbg;            // optional: sweep cycles per warp of CTA 0's first task (CMB200_DEBUG_TIMING)
    int tmem_cols, tmem_cpw;   // TMEM t-store: columns allocated per CTA, columns per warp (V & kVarTmem)
    const double *exp_tab;     // [kXlN] 2^(j/512), correctly rounded (kVarExpLds)
    int dbg_row0;
    int seg_space;             // KL is the segment-space table (column = segment * 32 + lane), pitch its row length
};

// Kernel variants (template parameter V, a bit set):
//   kVarTmem      the t-store lives in tensor memory (one LDTM / STTM per block of kU layers and
//                 lane) instead of shared memory, which frees 26 kB of shared memory per warp;
//   kVarCoefSmem  the long-wave emission coefficients of the band a warp is working on are staged in
//                 its shared memory and read with broadcast 128-bit loads instead of being loaded per
//                 lane from global memory and broadcast with shuffles;
//   kVarExpLds    (FP64) exp(-x) = 2^n 2^(j/512) e^r with the 512-entry table of 2^(j/512) in shared memory,
//                 replicated sixteen times so that lane l reads bank pair l & 15 whatever j it needs (no
//                 conflicts), a one-FMA range reduction and e^r = 1 + r + r^2 (1/2 + r/6): 7 FP64
//                 instructions per layer instead of 15, error <= 1.1e-14 where the parity gate is 1e-10;
constexpr int kVarTmem = 1, kVarCoefSmem = 2;
//   kVarSwDirect  (FP64, fused) short-wave fresh passes run sw_fresh_pass: every block is finished inside its own
//                 iteration (stage 1, exp, both recurrences, t-store traffic), only the block's warp reduction
//                 is carried into the next iteration; implies the shared-memory exp table at shared address
//                 0x10000 (the CTA asks for 128 kB of dynamic shared memory).
//   kVarRedTmem   (with kVarSwDirect) the lane sum of a block's eight values goes through tensor memory instead of
//                 18 shuffles + 28 selects: every lane parks its eight values in 16 exchange columns of its own
//                 TMEM lane (one STTM.32x32b.x16), and two LDTM.16x256b.x2 hand thread (u, c) = (lane & 3, lane >> 2)
//                 the values of layers u and u + 4 from lanes c, c + 8, c + 16, c + 24 -- a transpose done by the
//                 access shapes, no shuffle.  The thread adds its four lanes and accumulates the two sums in ITS OWN
//                 class accumulator pacc[block][u][c] (one LDS.128 / STS.128); the eight classes are added when the
//                 (warp, band) run is written out.  Short-wave replay-only passes take the same route (sw_replay_pass).
//   kVarSmallSt   (with kVarSwDirect) the direct short-wave passes store their blocks to tensor memory with eight
//                 STTM.x2 (the register pair of each value) instead of one STTM.x16.
//   kVarPair      (FP32, with kVarRedTmem) a short-wave pass carries TWO consecutive segments per lane in packed FP32
//                 arithmetic (FFMA2 / FMUL2 / FADD2), see sw_fresh_pair_pass; 8 warps per SM.
constexpr int kVarExpLds = 1024, kVarSwDirect = 4096, kVarRedTmem = 8192, kVarSmallSt = 16384, kVarPair = 32768;
// Class accumulators of kVarRedTmem: [2 directions][Lp / kU blocks][32 lanes] double2 (x: layer u, y: layer u + 4 of the
// block, u = lane & 3; class c = lane >> 2): every thread owns slot `lane` of every block, so a warp's LDS.128 /
// STS.128 covers 512 contiguous bytes in lane order (no bank conflicts).
__host__ __device__ inline size_t red_warp_bytes(int Lp) { return (size_t)2 * Lp * 8 * sizeof(double); }
struct RedAux {
    uint32_t scratch;               // tensor-memory address of this warp's 16 exchange columns (first lane of its quarter)
    unsigned up, dn;                // shared-window address of this thread's slot of block 0 of the class accumulators, per direction
};
// kVarSwDirect: shared-window address of the exp table (64 kB, 64 kB aligned so that its entry address is one LOP3).
// With kVarRedTmem the class accumulators need the room below it, so it moves up and the staged coefficient rows
// of the warps go above it.
constexpr unsigned xl_window(int V) { return (V & 8192) ? 0x20000u : 0x10000u; }
constexpr unsigned kXlBytes = 0x10000u;
constexpr int kXlN = 512, kXlLog = 9, kXlCopies = 16;       // table entries, log2, replicas
// Per-warp handles of the shared-memory exp table (kVarExpLds): already offset by lane & 15.
struct ExpAux {
    const double *lds;
    unsigned saddr;       // kVarSwDirect: xl_window(V) | (lane & 15) * 8
};
__device__ __forceinline__ double xl_scale(double tt, int n);
// Entry j of the table as the kernels keep it: 2^(j/512), correctly rounded, with j * 2^11 subtracted from the high
// word (integer arithmetic on the bit pattern; xl_scale adds n * 2^11 back before the value is used as a double).
inline double xl_table_entry(double two_to_j_over_512, int j) {
    unsigned long long bits;
    memcpy(&bits, &two_to_j_over_512, 8);
    bits -= (unsigned long long)j << (32 + 11);
    double r;
    memcpy(&r, &bits, 8);
    return r;
}
// Threads per CTA the kernel is compiled for: 8 warps, or 12 (168 registers) for the FP32 direct passes with one
// segment per pass, whose t-stores are half as wide (three warps per lane quarter of tensor memory instead of two).
template <typename Real>
constexpr int variant_threads(int V) {
    return (sizeof(Real) == 4 && (V & 8192) && !(V & 32768)) ? 384 : 256;
}
// Tensor-memory words per layer and lane of the t-store.
template <typename Real>
constexpr int variant_tstore_words(int V) {
    return (sizeof(Real) == 4 && (V & 32768)) ? 2 : (int)(sizeof(Real) / 4);
}

constexpr int kU = 8;          // layers per block of a pass (instruction-level parallelism)

// Per-level, warp-uniform lookup data of one molecule kept in shared memory: q/g and the offset
// of the table row KL[m][k][tidx] (fused path) or of the tau row k (resident-tau path).  A block
// of kU layers k0..k0+kU-1 touches the kU+1 interfaces k0..k0+kU; a downward pass carries the
// first and loads k0+1..k0+kU, an upward pass carries the last and loads k0..k0+kU-1.  Interface
// k sits at index k + 3 in off_dn / qg and at k + 4 in off_up, so both offset groups start on a
// 16-byte boundary; everything is padded beyond L-1 with q/g = 0 and a valid offset so that the
// sweeps need no bounds checks.
constexpr int kLevShift = 3;
template <typename Real>
struct LevelView {
    const unsigned *off_dn, *off_up;   // [MR][stride]
    const Real *qg;                    // [MR][stride]
    int stride;
    const Real *hdp;                   // [Lp] half layer thickness, 0 in the padding
};

// kU consecutive, 16-byte aligned shared-memory values -> registers with 128-bit loads.
template <typename T>
__device__ __forceinline__ void lds_block(const T *p, T (&o)[kU]) {
    constexpr int n16 = kU * sizeof(T) / 16;
    union {
        uint4 v[n16];
        T t[kU];
    } buf;
#pragma unroll
    for (int i = 0; i < n16; ++i) buf.v[i] = reinterpret_cast<const uint4 *>(p)[i];
#pragma unroll
    for (int u = 0; u < kU; ++u) o[u] = buf.t[u];
}
// The same for a run that starts one element *before* a 16-byte boundary (p + 1 is aligned).
__device__ __forceinline__ void lds_block_m1(const double *p, double (&o)[kU]) {
    o[0] = p[0];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double2 v = *reinterpret_cast<const double2 *>(p + 1 + 2 * i);
        o[1 + 2 * i] = v.x;
        o[2 + 2 * i] = v.y;
    }
    o[7] = p[7];
}
__device__ __forceinline__ void lds_block_m1(const float *p, float (&o)[kU]) {
    o[0] = p[0];
    const float4 v = *reinterpret_cast<const float4 *>(p + 1);
    o[1] = v.x, o[2] = v.y, o[3] = v.z, o[4] = v.w;
    const float2 w = *reinterpret_cast<const float2 *>(p + 5);
    o[5] = w.x, o[6] = w.y;
    o[7] = p[7];
}

__host__ __device__ inline int padded_layers(int L) { return ((L - 1) + kU - 1) / kU * kU; }

// Shared-memory layout of one CTA (every sub-array 16-byte aligned): the column's level tables
// once, then per warp the band accumulators and the t-store.
template <typename Real, int MM, int V = 0>
struct CtaSmem {
    size_t off_offdn, off_offup, off_qg, off_hdp, shared_total, warp_bacc, warp_coef, warp_pacc, coef_bytes, warp_ts, warp_total;
    int lev_stride;
    // ts_rows: layers of the t-store a warp keeps in shared memory (-1: all of them, or none with kVarTmem)
    size_t off_xtab, off_bsrc;      // off_bsrc: [2][B] boundary sources of the column's bands (kVarRedTmem)
    __host__ __device__ CtaSmem(int L, int ts_rows = -1, int B = 0) {
        auto up16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
        const int Lp = padded_layers(L), MR = MM > 0 ? MM : 1;
        if (ts_rows < 0) ts_rows = (V & kVarTmem) ? 0 : Lp;
        lev_stride = Lp + kU;
        size_t b = 0;
        off_offdn = b;
        b = up16(b + (size_t)MR * lev_stride * sizeof(unsigned));
        off_offup = b;
        b = up16(b + (size_t)MR * lev_stride * sizeof(unsigned));
        off_qg = b;
        b = up16(b + (size_t)MR * lev_stride * sizeof(Real));
        off_hdp = b;
        b = up16(b + (size_t)Lp * sizeof(Real) * ((V & 32768) ? 2 : 1));      // kVarPair: + the same times -log2 e
        off_xtab = b;
        if ((V & kVarExpLds) && !(V & kVarSwDirect)) b = up16(b + (size_t)kXlN * kXlCopies * sizeof(double));
        off_bsrc = b;
        if (V & kVarRedTmem) b = up16(b + (size_t)2 * B * sizeof(double));
        shared_total = b;
        // per warp: band accumulators (bacc_up, bacc_dn [Lp + kU] each), the staged coefficient rows [4][Lp],
        // the class accumulators (kVarRedTmem), the t-store [Lp][32] (without kVarTmem)
        warp_bacc = 0;
        warp_coef = up16((size_t)2 * (Lp + kU) * sizeof(double));
        coef_bytes = (V & kVarCoefSmem) ? (size_t)4 * Lp * sizeof(Real) : 0;
        if (V & kVarRedTmem) {      // class accumulators here; FP64: the coefficient rows go above the exp table
            warp_pacc = sizeof(Real) == 8 ? warp_coef : up16(warp_coef + coef_bytes);
            warp_ts = up16(warp_pacc + red_warp_bytes(Lp));
        } else {
            warp_pacc = 0;
            warp_ts = up16(warp_coef + coef_bytes);
        }
        warp_total = up16(warp_ts + (size_t)ts_rows * 32 * sizeof(Real));
    }
    __host__ __device__ size_t total(int warps) const { return shared_total + warp_total * warps; }
};

// Where a warp keeps its t-store: a shared-memory row pointer (already offset by the lane) or a
// tensor-memory address (lane quarter and first column of the warp).
template <typename Real>
struct TStore {
    Real *lane;        // shared-memory row of layer k: lane + k * 32
    uint32_t taddr;    // tensor-memory column of layer k: taddr + k * (sizeof(Real) / 4)
};

// exp(-x) for the layer transmissivity.  FP64: n = rint(-x log2 e), r = -x - n ln2 (two-term
// Cody-Waite), degree-11 near-minimax polynomial on |r| <= ln2/2 (max rel. error 1.6e-17 before
// rounding), scaled by 2^n through the exponent field; no special cases beyond the argument clamp.
// Coefficients sit in constant memory so that each DFMA takes its constant as an operand.
static __constant__ double kExpC[16] = {      // (static: the kernels are instantiated in several translation units)
    2.51100492048186583e-08, 2.76326547225277896e-07, 2.75572408872298695e-06, 2.48014854415613131e-05,
    1.98412698900764028e-04, 1.38888889523528631e-03, 8.33333333331958900e-03, 4.16666666664879531e-02,
    1.66666666666666796e-01, 5.00000000000001887e-01, 1.0, 1.4426950408889634074,
    6755399441055744.0 /* 1.5 * 2^52 */, -6.93147180369123816490e-01, -1.90821492927058770002e-10, 0.0};

// Optical thicknesses beyond 708 (exp(-708) = 3e-308, nothing at the 1e-10 parity level) are
// clamped through the high word -- one integer min -- so that 2^n below stays a normal number
// and the range reduction never sees huge arguments.  (Layer thicknesses are >= 0 on this path.)
__device__ __forceinline__ double clamp_thickness(double x) {
    return __hiloint2double(min(__double2hiint(x), 0x40862000), __double2loint(x));
}
__device__ __forceinline__ double exp_neg(double x) {
    const double y = -clamp_thickness(x);
    double tn = fma(y, kExpC[11], kExpC[12]);
    const int n = __double2loint(tn);
    tn -= kExpC[12];
    double r = fma(tn, kExpC[13], y);
    r = fma(tn, kExpC[14], r);
    double p = kExpC[0];
#pragma unroll
    for (int i = 1; i <= 10; ++i) p = fma(p, r, kExpC[i]);
    p = fma(p, r, kExpC[10]);
    return __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
}
// The same for kU values at once, written step-by-step across the values so that the kU
// independent dependency chains are adjacent in program order and the FP64 pipe stays busy.
__device__ __forceinline__ void exp_neg_block(const double (&xin)[kU], double (&out)[kU]) {
    double x[kU], tn[kU], r[kU], p[kU];
    int n[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
        x[u] = clamp_thickness(xin[u]);
        tn[u] = fma(-x[u], kExpC[11], kExpC[12]);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
        n[u] = __double2loint(tn[u]);
        tn[u] -= kExpC[12];
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) r[u] = fma(tn[u], kExpC[13], -x[u]);
#pragma unroll
    for (int u = 0; u < kU; ++u) {
        r[u] = fma(tn[u], kExpC[14], r[u]);
        p[u] = kExpC[0];
    }
#pragma unroll
    for (int i = 1; i <= 10; ++i)
#pragma unroll
        for (int u = 0; u < kU; ++u) p[u] = fma(p[u], r[u], kExpC[i]);
#pragma unroll
    for (int u = 0; u < kU; ++u) {
        p[u] = fma(p[u], r[u], kExpC[10]);
        out[u] = __hiloint2double(__double2hiint(p[u]) + (n[u] << 20), __double2loint(p[u]));
    }
}
__device__ __forceinline__ void exp_neg_block(const float (&x)[kU], float (&out)[kU]) {
#pragma unroll
    for (int u = 0; u < kU; ++u) out[u] = __expf(-x[u]);
}
__device__ __forceinline__ float exp_neg(float x) { return __expf(-x); }

// pi B(nu, T) for the column prologue, which evaluates it ~4000 times per column (every long-wave band at every
// interface and cell centre): K = h f / k_B and N = pi 100 c 2 h f^3 / c^2 come per band (f = 100 c nu), rT = 1 / T per
// level, so what is left is exp(K rT) with the kernel's own polynomial, a reciprocal by two Newton steps on
// rcp.approx and two multiplies -- ~30 instructions where B_wavenumber's expression with libdevice exp and two IEEE
// divisions takes ~120 (real_gas.py:14-42; agreement ~1e-15, the stand-alone Planck entry points keep that expression).
__device__ __forceinline__ double rcp_fast(double d) {          // d normal and positive
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    r = fma(r, fma(-d, r, 1.0), r);
    return fma(r, fma(-d, r, 1.0), r);
}
__device__ __forceinline__ double planck_fast(double K, double N, double rT) {
    const double u = fmin(K * rT, 708.0);                       // beyond: B < 1e-290, nothing at any tolerance here
    return N * rcp_fast(exp_neg(-u) - 1.0);
}
__device__ __forceinline__ void planck_band_constants(const PlanckConst &pc, double nu, double &K, double &N) {
    const double f = 100.0 * nu * pc.c;
    K = pc.h * f / pc.kB;
    N = M_PI * (100.0 * pc.c) * (2.0 * pc.h * (f * f * f) / (pc.c * pc.c));
}

__device__ __forceinline__ double shfl_f64(double v, int src) {
    const int lo = __shfl_sync(0xffffffffu, __double2loint(v), src), hi = __shfl_sync(0xffffffffu, __double2hiint(v), src);
    return __hiloint2double(hi, lo);
}

// Sum eight per-lane values over the warp at once: three exchange steps that halve the number
// of values a lane carries while doubling the lanes each partial sum covers, then two plain
// butterfly steps -- 9 additions and 9 shuffles for 8 sums instead of 40 + 40.  Afterwards
// every lane holds the total of value ((lane>>4)&1)*4 + ((lane>>3)&1)*2 + ((lane>>2)&1).
template <typename Real>
__device__ __forceinline__ Real reduce8(const Real (&v)[8], int lane) {
    const unsigned full = 0xffffffffu;
    const bool h4 = lane & 16, h3 = lane & 8, h2 = lane & 4;
    Real a4[4], a2[2];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const Real keep = h4 ? v[i + 4] : v[i], send = h4 ? v[i] : v[i + 4];
        a4[i] = keep + __shfl_xor_sync(full, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const Real keep = h3 ? a4[i + 2] : a4[i], send = h3 ? a4[i] : a4[i + 2];
        a2[i] = keep + __shfl_xor_sync(full, send, 8);
    }
    const Real keep = h2 ? a2[1] : a2[0], send = h2 ? a2[0] : a2[1];
    Real c = keep + __shfl_xor_sync(full, send, 4);
    c += __shfl_xor_sync(full, c, 2);
    c += __shfl_xor_sync(full, c, 1);
    return c;
}

// Table values (or tau) a fresh pass starts from, fetched while the pass before it is still running:
// the loaded interfaces of its first block and the raw values behind the carried interface.
template <typename Real, int MM>
struct PassFeed {
    using Src = typename std::conditional<(MM > 0), Real, double>::type;
    Src raw[kU][MM > 0 ? MM : 1];
    Src icraw[MM > 0 ? MM : 1];
};
template <typename Real, int MM>
__device__ __forceinline__ void prefetch_pass(const bool up, const int Lp, const void *src_lane,
                                              const LevelView<Real> &lv, PassFeed<Real, MM> &fd) {
    using Src = typename PassFeed<Real, MM>::Src;
    constexpr int MR = MM > 0 ? MM : 1;
    constexpr unsigned kOffUnit = MM > 0 ? 16u : 8u;
    const char *src_bytes = (const char *)src_lane;
    const unsigned *offs = up ? lv.off_up : lv.off_dn;
    const int k0 = up ? Lp - kU : 0, kc = (up ? Lp : 0) + kLevShift;
#pragma unroll
    for (int m = 0; m < MR; ++m) {
        unsigned off[kU];
        lds_block(offs + m * lv.stride + k0 + 4, off);
#pragma unroll
        for (int u = 0; u < kU; ++u) fd.raw[u][m] = *(const Src *)(src_bytes + (size_t)off[u] * kOffUnit);
        fd.icraw[m] = *(const Src *)(src_bytes + (size_t)lv.off_dn[m * lv.stride + kc] * kOffUnit);
    }
}

// One pass over the levels for the (up to) two chains this lane holds.
//   UP = false: TOA -> surface, value of layer k belongs to interface k+1 (caller passes bacc_dn + 1)
//   UP = true:  surface -> TOA, value of layer k belongs to interface k   (caller passes bacc_up)
//   FRESH:      segment whose t_k are computed in this pass and stored to ts (weight wF, rows src_lane)
//   replay:     segment whose t_k are read back from ts (weight wR; wR = 0 on the first pass of a band)
//   LW:         long-wave band (emission recurrences; coefL / coefG = this direction's coefficient rows)
// b0 is the boundary source of the direction (stellar at the top, ground at the bottom); the
// nu-weight is folded into the initial state so that the values handed to the warp reduction are
// already weighted.
//
// The pass is software-pipelined by hand over blocks of kU layers, because a lane's work is a set
// of long dependency chains and only two warps per scheduler fit next to the t-store: while the kU
// independent exp evaluations of block b are in flight (stage A), the serial recurrences of block
// b-1 (stage B, both chains) and the warp reduction of their summed values (stage C) are issued
// in between, statement by statement -- ptxas keeps this order.
// Two consecutive, 8/16-byte aligned shared-memory values with one load.
__device__ __forceinline__ void lds_pair(const double *p, double (&o)[2]) {
    const double2 v = *reinterpret_cast<const double2 *>(p);
    o[0] = v.x, o[1] = v.y;
}
__device__ __forceinline__ void lds_pair(const float *p, float (&o)[2]) {
    const float2 v = *reinterpret_cast<const float2 *>(p);
    o[0] = v.x, o[1] = v.y;
}

template <typename Real, int MM, bool LW, bool UP, bool FRESH, int V>
__device__ __forceinline__ void sweep_pass(const int lane, const int nl, const int Lp, const Real wF, const Real wR,
                                           const Real b0, const void *src_lane, const void *next_src,
                                           PassFeed<Real, MM> &fd, const LevelView<Real> lv, const TStore<Real> tst,
                                           double *bacc, const Real *coefL, const Real *coefG, const ExpAux xaux) {
    constexpr bool FUSED = MM > 0;
    constexpr int MR = MM > 0 ? MM : 1;
    constexpr bool F64 = sizeof(Real) == 8;
    constexpr bool TM = (V & kVarTmem) != 0, CS = (V & kVarCoefSmem) != 0;
    constexpr bool XL = F64 && (V & kVarExpLds) != 0;
    constexpr int WPR = sizeof(Real) / 4, NW = kU * WPR;   // 32-bit words per value / per block and lane
    using Src = typename std::conditional<FUSED, Real, double>::type;
    const char *src_bytes = (const char *)src_lane;
    constexpr unsigned kOffUnit = FUSED ? 16u : 8u;   // table rows are 64-byte aligned, tau rows only 8
    const unsigned full = 0xffffffffu;
    const bool h4 = lane & 16, h3 = lane & 8, h2 = lane & 4;
    const int ridx = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
    const bool rlane = (lane & 3) == 0;
    const int kfirst = UP ? Lp - kU : 0, kstep = UP ? -kU : kU;
    Real *const tsl = tst.lane;                            // layer k sits at tsl[k * 32]
    // Tensor memory: one block (kU layers) of the t-store <-> NW words per lane.
    auto ts_fetch = [&](int k0, uint32_t(&w)[NW]) { tmem::ld<NW>(tst.taddr + (uint32_t)(k0 * WPR), w); };
    auto ts_park = [&](int k0, const Real(&t)[kU]) {
        uint32_t tw[NW];
        tmem::pack<kU>(t, tw);
        tmem::st<NW>(tst.taddr + (uint32_t)(k0 * WPR), tw);
    };

    // ---- warp reduction of v[0..7], split into five steps for the interleave ----------------------
    Real v[kU], a4[4], a2[2], c1;
    auto C1 = [&]() {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const Real keep = h4 ? v[i + 4] : v[i], send = h4 ? v[i] : v[i + 4];
            a4[i] = keep + __shfl_xor_sync(full, send, 16);
        }
    };
    auto C2 = [&]() {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const Real keep = h3 ? a4[i + 2] : a4[i], send = h3 ? a4[i] : a4[i + 2];
            a2[i] = keep + __shfl_xor_sync(full, send, 8);
        }
    };
    auto C3 = [&]() {
        const Real keep = h2 ? a2[1] : a2[0], send = h2 ? a2[0] : a2[1];
        c1 = keep + __shfl_xor_sync(full, send, 4);
    };
    auto C4 = [&]() { c1 += __shfl_xor_sync(full, c1, 2); };
    auto C5 = [&]() { c1 += __shfl_xor_sync(full, c1, 1); };
    // band accumulator of the reduced block: fetched early so that the read latency is off the tail
    double bold = 0.0;
    auto Cload = [&](int kc) {
        if (rlane && kc + ridx < nl) bold = bacc[kc + ridx];
    };
    auto Cstore = [&](int kc) {
        if (rlane && kc + ridx < nl) bacc[kc + ridx] = bold + (double)c1;
    };

    // ---- recurrence step of layer kb+u for both chains (stage B) --------------------------------------
    Real EF = wF * b0, HF = Real(0), ER = wR * b0, HR = Real(0);
    Real tp[kU], tR[kU];
    // Long-wave emission coefficients.  Default: they live in global memory (L2): one coalesced load
    // per block, a block ahead, gives lane 8 r + i the coefficient of row r (coefL, coefG = coefL + Lp)
    // and layer k0 + i; the recurrence broadcasts them with shuffles.  kVarCoefSmem: coefL / coefG
    // point into the warp's shared memory and two layers (one pair) are read per broadcast load, two
    // pairs ahead, into the rolling buffers cA / cB ([row][layer parity]).
    const Real *crow = coefL + ((lane >> 3) & 1) * (coefG - coefL) + (lane & (kU - 1));
    Real ccur = Real(0), cnext = Real(0);
    Real cA[2][2] = {{Real(0), Real(0)}, {Real(0), Real(0)}}, cB[2][2] = {{Real(0), Real(0)}, {Real(0), Real(0)}};
    auto ldpair = [&](int kb, int pi, Real(&c)[2][2]) {
        if constexpr (LW && CS) {
            const int k = kb + 2 * (UP ? 3 - pi : pi);
            lds_pair(coefL + k, c[0]);
            lds_pair(coefG + k, c[1]);
        }
    };
    auto ord = [](int o) { return UP ? kU - 1 - o : o; };   // order in which a block's layers are swept
    auto Bstep = [&](int kb, int o, Real(&out)[kU]) {
        const int u = ord(o);
        Real cl = Real(0), cg = Real(0), valF = Real(0), valR;
        if (LW) {
            if constexpr (CS) {
                const Real(&c)[2][2] = ((o >> 1) & 1) ? cB : cA;
                cl = c[0][u & 1];
                cg = c[1][u & 1];
            } else {
                cl = __shfl_sync(full, ccur, u);        // lanes 0-7 hold the block's "last/first" row, 8-15 its "gen" row
                cg = __shfl_sync(full, ccur, kU + u);
            }
        }
        if constexpr (FRESH) {
            EF *= tp[u];
            if (LW) {
                const Real om = fma(-wF, tp[u], wF);          // w (1 - t)
                const Real S = tp[u] * HF;
                valF = EF + S + cl * om;
                HF = fma(cg, om, S);
            } else {
                valF = EF;
            }
        }
        ER *= tR[u];
        if (LW) {
            const Real om = fma(-wR, tR[u], wR);
            const Real S = tR[u] * HR;
            valR = ER + S + cl * om;
            HR = fma(cg, om, S);
        } else {
            valR = ER;
        }
        out[u] = FRESH ? valF + valR : valR;
    };
    // Steps 2 pi, 2 pi + 1 of block kb; afterwards the pair buffer they used is refilled with pair
    // pi + 2 of this block or pair pi - 2 of the next block (knext; < 0: there is none).
    auto Bpair = [&](int kb, int pi, Real(&out)[kU], int knext) {
        Bstep(kb, 2 * pi, out);
        Bstep(kb, 2 * pi + 1, out);
        if constexpr (LW && CS) {
            if (pi < 2)
                ldpair(kb, pi + 2, pi & 1 ? cB : cA);
            else if (knext >= 0)
                ldpair(knext, pi - 2, pi & 1 ? cB : cA);
        }
    };

    if constexpr (FRESH) {
        // raw[u][m]: table value (or tau) of the block's loaded interfaces, fetched one block ahead so
        // that the L2 / HBM latency hides behind the previous block's work.
        const unsigned *offs = UP ? lv.off_up : lv.off_dn;   // either way the group of block k0 starts at k0 + 4
        auto &raw = fd.raw;                                  // first block already in flight (prefetch_pass)
        auto ld = [&](unsigned off) { return *(const Src *)(src_bytes + (size_t)off * kOffUnit); };
        auto load_block = [&](int k0) {
#pragma unroll
            for (int m = 0; m < MR; ++m) {
                unsigned off[kU];
                lds_block(offs + m * lv.stride + k0 + 4, off);
#pragma unroll
                for (int u = 0; u < kU; ++u) raw[u][m] = ld(off[u]);
            }
        };
        Src Ic;                                               // the carried interface of the next block
        if (FUSED) {
            Real acc = Real(0);
#pragma unroll
            for (int m = 0; m < MR; ++m) acc += lv.qg[m * lv.stride + (UP ? Lp : 0) + kLevShift] * (Real)fd.icraw[m];
            Ic = (Src)acc;
        } else {
            Ic = fd.icraw[0];
        }

        // t-store traffic of a block.  Shared memory: spread over the iteration -- once the recurrences
        // are done with layer kb+u the fresh t (still in tp) is parked in its row, and once they are done
        // with the whole block the replay chain's t of block k0 is fetched from the rows that block k0's
        // fresh t will overwrite one iteration later.  Tensor memory: the replay block is fetched with one
        // LDTM at the same point, and at the end of the iteration (after the wait on that load) the fresh
        // t of block k0 goes to the same columns with one STTM.
        uint32_t tRw[NW];
#pragma unroll
        for (int i = 0; i < NW; ++i) tRw[i] = 0u;
        auto park = [&](int kb, int o) {
            if constexpr (!TM) tsl[(kb + ord(o)) * 32] = tp[ord(o)];
        };
        auto fetch_replay = [&](int k0) {
            if constexpr (TM) {
                ts_fetch(k0, tRw);
            } else {
#pragma unroll
                for (int u = 0; u < kU; ++u) tR[u] = tsl[(k0 + u) * 32];
            }
        };
        auto iter = [&](auto doA_t, auto doB_t, const int k0) {
            constexpr bool DO_A = decltype(doA_t)::value, DO_B = decltype(doB_t)::value;
            const int kb = k0 - kstep;
            const int knext = DO_A ? k0 : -1;     // the block stage B works on in the next iteration
            Real x[kU], t[kU];
            if (DO_B) Cload(kb);
            if (LW && DO_A && !CS) cnext = crow[k0];
            // stage A set-up: layer integrand at the block's interfaces, next block's loads, dtau
            if (DO_A) {
                Src In[kU];
                if (FUSED) {
                    Real qg[kU], acc[kU];
#pragma unroll
                    for (int m = 0; m < MR; ++m) {
                        if (UP)
                            lds_block_m1(lv.qg + m * lv.stride + k0 + kLevShift, qg);
                        else
                            lds_block(lv.qg + m * lv.stride + k0 + 1 + kLevShift, qg);
#pragma unroll
                        for (int u = 0; u < kU; ++u) acc[u] = m ? fma(qg[u], (Real)raw[u][m], acc[u]) : qg[u] * (Real)raw[u][0];
                    }
#pragma unroll
                    for (int u = 0; u < kU; ++u) In[u] = (Src)acc[u];
                } else {
#pragma unroll
                    for (int u = 0; u < kU; ++u) In[u] = raw[u][0];
                }
                if (UP ? k0 >= kU : k0 + kU < Lp) load_block(k0 + kstep);
                Real hd[kU];
                if (FUSED) lds_block(lv.hdp + k0, hd);
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    const Src lo = UP ? In[u] : (u ? In[u - 1] : Ic);
                    const Src hi = UP ? (u < kU - 1 ? In[u + 1] : Ic) : In[u];
                    x[u] = FUSED ? (Real)(hd[u] * (Real)(lo + hi)) : (Real)(hi - lo);
                }
                Ic = UP ? In[0] : In[kU - 1];
            }
            if constexpr (XL) {
                // exp(-x) with the shared-memory 2^(j/512) table: n' = rint(-x 512/ln2) = 512 n + j,
                // r = -x - n' ln2/512 (one FMA with the rounded constant: error <= n' 2^-63, below 1e-13
                // for every x <= 708), t = 2^n T_j (1 + r + r^2 (1/2 + r/6)); step by step across the layers
                constexpr double kK = 0x1.71547652b82fep+9, kMagic = 6755399441055744.0, kC = -0x1.62e42fefa39efp-10;
                double tn[kU], r[kU], r2[kU], p[kU], Tj[kU];
                int nn[kU];
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) {
                        x[u] = clamp_thickness(x[u]);
                        tn[u] = fma(-x[u], kK, kMagic);
                    }
                }
                if (DO_B) { Bpair(kb, 0, v, knext); park(kb, 0); park(kb, 1); }
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) {
                        nn[u] = __double2loint(tn[u]);
                        tn[u] -= kMagic;
                        Tj[u] = xl_scale(xaux.lds[(nn[u] & (kXlN - 1)) * kXlCopies], nn[u]);      // 2^(j/512) 2^n
                    }
                }
                if (DO_B) { Bpair(kb, 1, v, knext); park(kb, 2); park(kb, 3); }
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) r[u] = fma(tn[u], kC, -x[u]);
                }
                if (DO_B) { Bpair(kb, 2, v, knext); park(kb, 4); park(kb, 5); }
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) {
                        r2[u] = r[u] * r[u];
                        p[u] = fma(r[u], 1.0 / 6.0, 0.5);
                    }
                }
                if (DO_B) { Bpair(kb, 3, v, knext); park(kb, 6); park(kb, 7); C1(); }
                if (DO_A) fetch_replay(k0);
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) p[u] = fma(r2[u], p[u], r[u]);
                }
                if (DO_B) { C2(); C3(); }
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) t[u] = fma(Tj[u], p[u], Tj[u]);
                }
                if (DO_B) { C4(); C5(); }
            } else if constexpr (F64) {
                double tn[kU], r[kU], p[kU];
                int n[kU];
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) {
                        x[u] = clamp_thickness(x[u]);
                        tn[u] = fma(-x[u], kExpC[11], kExpC[12]);
                    }
                }
                if (DO_B) { Bpair(kb, 0, v, knext); park(kb, 0); park(kb, 1); }
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) {
                        n[u] = __double2loint(tn[u]);
                        tn[u] -= kExpC[12];
                    }
                }
                if (DO_B) { Bpair(kb, 1, v, knext); park(kb, 2); park(kb, 3); }
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) r[u] = fma(tn[u], kExpC[13], -x[u]);
                }
                if (DO_B) { Bpair(kb, 2, v, knext); park(kb, 4); park(kb, 5); }
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) {
                        r[u] = fma(tn[u], kExpC[14], r[u]);
                        p[u] = kExpC[0];
                    }
                }
                if (DO_B) { Bpair(kb, 3, v, knext); park(kb, 6); park(kb, 7); }
                if (DO_A) fetch_replay(k0);
#pragma unroll
                for (int i = 1; i <= 10; ++i) {
                    if (DO_A) {
#pragma unroll
                        for (int u = 0; u < kU; ++u) p[u] = fma(p[u], r[u], kExpC[i]);
                    }
                    if (DO_B) {
                        if (i == 1) C1();
                        if (i == 3) C2();
                        if (i == 5) C3();
                        if (i == 7) C4();
                        if (i == 9) C5();
                    }
                }
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) {
                        p[u] = fma(p[u], r[u], kExpC[10]);
                        t[u] = __hiloint2double(__double2hiint(p[u]) + (n[u] << 20), __double2loint(p[u]));
                    }
                }
            } else {
                if (DO_A) {
#pragma unroll
                    for (int u = 0; u < kU; ++u) t[u] = exp_neg(x[u]);
                }
                if (DO_B) {
#pragma unroll
                    for (int pi = 0; pi < kU / 2; ++pi) {
                        Bpair(kb, pi, v, knext);
                        park(kb, 2 * pi);
                        park(kb, 2 * pi + 1);
                    }
                }
                if (DO_A) fetch_replay(k0);
                if (DO_B) { C1(); C2(); C3(); C4(); C5(); }
            }
            if (DO_B) Cstore(kb);
            if (DO_A) {
                if constexpr (TM) {
                    tmem::wait_ld<NW>(tRw);
                    tmem::unpack<kU>(tRw, tR);
                    ts_park(k0, t);
                }
#pragma unroll
                for (int u = 0; u < kU; ++u) tp[u] = t[u];
                ccur = cnext;
                if (!DO_B) {          // first iteration of the pass: both coefficient pair buffers are empty
                    ldpair(k0, 0, cA);
                    ldpair(k0, 1, cB);
                }
            }
        };
        iter(std::true_type{}, std::false_type{}, kfirst);
        if (UP) {
            for (int k0 = Lp - 2 * kU; k0 >= 0; k0 -= kU) iter(std::true_type{}, std::true_type{}, k0);
            if (next_src) prefetch_pass<Real, MM>(!UP, Lp, next_src, lv, fd);   // raw is free from here on
            iter(std::false_type{}, std::true_type{}, -kU);
        } else {
            for (int k0 = kU; k0 < Lp; k0 += kU) iter(std::true_type{}, std::true_type{}, k0);
            if (next_src) prefetch_pass<Real, MM>(!UP, Lp, next_src, lv, fd);   // raw is free from here on
            iter(std::false_type{}, std::true_type{}, Lp);
        }
        if constexpr (TM) tmem::wait_st();   // the next pass reads these columns back
    } else {
        // ---- replay only: recurrence of block b interleaved with the reduction of the block before --
        Real tn8[kU], vp[kU];
        uint32_t tnw[NW];
#pragma unroll
        for (int i = 0; i < NW; ++i) tnw[i] = 0u;
        auto iter = [&](auto doB_t, auto doC_t, const int k0) {
            constexpr bool DO_B = decltype(doB_t)::value, DO_C = decltype(doC_t)::value;
            const bool more = DO_B && (UP ? k0 >= kU : k0 + kU < Lp);
            const int knext = more ? k0 + kstep : -1;
            if (DO_C) Cload(k0 - kstep);
            if (more) {
                if constexpr (TM) {
                    ts_fetch(k0 + kstep, tnw);
                } else {
#pragma unroll
                    for (int u = 0; u < kU; ++u) tn8[u] = tsl[(k0 + kstep + u) * 32];
                }
                if (LW && !CS) cnext = crow[k0 + kstep];
            }
            if (DO_B) Bstep(k0, 0, vp);
            if (DO_C) C1();
            if (DO_B) {
                Bstep(k0, 1, vp);
                if (LW && CS) ldpair(k0, 2, cA);
                Bstep(k0, 2, vp);
            }
            if (DO_C) C2();
            if (DO_B) {
                Bstep(k0, 3, vp);
                if (LW && CS) ldpair(k0, 3, cB);
                Bstep(k0, 4, vp);
            }
            if (DO_C) C3();
            if (DO_B) {
                Bstep(k0, 5, vp);
                if (LW && CS && knext >= 0) ldpair(knext, 0, cA);
            }
            if (DO_C) C4();
            if (DO_B) {
                Bstep(k0, 6, vp);
                Bstep(k0, 7, vp);
                if (LW && CS && knext >= 0) ldpair(knext, 1, cB);
            }
            if (DO_C) {
                C5();
                Cstore(k0 - kstep);
            }
            if (DO_B) {
                if constexpr (TM) {
                    if (more) tmem::wait_ld<NW>(tnw);
                    tmem::unpack<kU>(tnw, tR);
                } else {
#pragma unroll
                    for (int u = 0; u < kU; ++u) tR[u] = tn8[u];
                }
#pragma unroll
                for (int u = 0; u < kU; ++u) v[u] = vp[u];
                ccur = cnext;
            }
        };
        if (next_src) prefetch_pass<Real, MM>(!UP, Lp, next_src, lv, fd);
        if constexpr (TM) {
            ts_fetch(kfirst, tnw);
            tmem::wait_ld<NW>(tnw);
            tmem::unpack<kU>(tnw, tR);
        } else {
#pragma unroll
            for (int u = 0; u < kU; ++u) tR[u] = tsl[(kfirst + u) * 32];
        }
        if (LW && !CS) ccur = crow[kfirst];
        ldpair(kfirst, 0, cA);
        ldpair(kfirst, 1, cB);
        iter(std::true_type{}, std::false_type{}, kfirst);
        if (UP) {
            for (int k0 = Lp - 2 * kU; k0 >= 0; k0 -= kU) iter(std::true_type{}, std::true_type{}, k0);
            iter(std::false_type{}, std::true_type{}, -kU);
        } else {
            for (int k0 = kU; k0 < Lp; k0 += kU) iter(std::true_type{}, std::true_type{}, k0);
            iter(std::false_type{}, std::true_type{}, Lp);
        }
    }
}

// kVarRedTmem: the lane sum of a block's eight values through tensor memory (see the variant list).
//   red_park   every lane writes its eight values to the warp's 16 exchange columns;
//   red_fetch  (after the stores have landed) thread (u, c) = (lane & 3, lane >> 2) reads values u and u + 4 of
//              lanes c, c + 8 (q0) and c + 16, c + 24 (q1); valid after the next tcgen05.wait::ld;
//   red_accumulate  adds them, lanes in ascending order pairwise, to the thread's class accumulator.
// A block of kU values -> consecutive columns of the lane (two per FP64 value, one per FP32 value): one STTM.x16 /
// .x8, or (SMALL) one STTM per value whose operand is the register (pair) the value was produced in -- ptxas
// assembles wider operand groups with moves (two per FP64 value), a single value never needs one.
template <bool SMALL>
__device__ __forceinline__ void tm_store_block(uint32_t addr, const double (&v)[kU]) {
    if constexpr (SMALL) {
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            uint32_t q[2] = {(uint32_t)__double2loint(v[u]), (uint32_t)__double2hiint(v[u])};
            tmem::st_small<2>(addr + (uint32_t)(2 * u), q);
        }
    } else {
        uint32_t w[2 * kU];
        tmem::pack<kU>(v, w);
        tmem::st<2 * kU>(addr, w);
    }
}
template <bool SMALL>
__device__ __forceinline__ void tm_store_block(uint32_t addr, const float (&v)[kU]) {
    if constexpr (SMALL) {
#pragma unroll
        for (int u = 0; u < kU; ++u) tmem::st_one(addr + (uint32_t)u, __float_as_uint(v[u]));
    } else {
        uint32_t w[kU];
        tmem::pack<kU>(v, w);
        tmem::st<kU>(addr, w);
    }
}
template <bool SMALL, typename Real>
__device__ __forceinline__ void red_park(const Real (&v)[kU], uint32_t scratch) {
    tm_store_block<SMALL>(scratch, v);
}
// The transposed values of a parked block: FP64 8 words from lanes c, c + 8 and 8 from c + 16, c + 24 (values u, u + 4,
// u = lane & 3); FP32 4 + 4 words (values 2 u, 2 u + 1).
template <typename Real>
struct RedQ {
    static constexpr int N = sizeof(Real) == 8 ? 8 : 4;
    uint32_t q0[N], q1[N];
};
__device__ __forceinline__ void red_fetch_issued(uint32_t addr, RedQ<double> &r) {     // the stores have landed already
    tmem::ld_16x256b_x2(addr, r.q0);
    tmem::ld_16x256b_x2(addr + (16u << 16), r.q1);
}
__device__ __forceinline__ void red_fetch_issued(uint32_t addr, RedQ<float> &r) {
    tmem::ld_16x256b_x1(addr, r.q0);
    tmem::ld_16x256b_x1(addr + (16u << 16), r.q1);
}
template <typename Real>
__device__ __forceinline__ void red_fetch(uint32_t scratch, RedQ<Real> &r) {
    tmem::wait_st();
    red_fetch_issued(scratch, r);
}
// (one 128-bit access each way, written as PTX: left to itself nvcc splits the pair into two 64-bit accesses
// whose 16-byte lane stride costs two extra wavefronts)
__device__ __forceinline__ double2 red_slot_load(unsigned slot) {
    double2 a;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(a.x), "=d"(a.y) : "r"(slot));
    return a;
}
__device__ __forceinline__ void red_slot_store(unsigned slot, double2 a) {
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(slot), "d"(a.x), "d"(a.y) : "memory");
}
// Four lanes added pairwise in ascending order, into the class accumulator: .x = value u (FP64) / 2 u (FP32) of the
// block, .y = value u + 4 / 2 u + 1.  FP32 chains are added in FP32 over the four lanes, in FP64 from there on.
__device__ __forceinline__ void red_accumulate(const RedQ<double> &r, double2 acc, unsigned slot) {
    auto d = [](const uint32_t(&q)[8], int i) { return __hiloint2double((int)q[2 * i + 1], (int)q[2 * i]); };
    acc.x += (d(r.q0, 0) + d(r.q0, 1)) + (d(r.q1, 0) + d(r.q1, 1));
    acc.y += (d(r.q0, 2) + d(r.q0, 3)) + (d(r.q1, 2) + d(r.q1, 3));
    red_slot_store(slot, acc);
}
__device__ __forceinline__ void red_accumulate(const RedQ<float> &r, double2 acc, unsigned slot) {
    auto f = [](uint32_t w) { return __uint_as_float(w); };
    acc.x += (double)((f(r.q0[0]) + f(r.q0[2])) + (f(r.q1[0]) + f(r.q1[2])));
    acc.y += (double)((f(r.q0[1]) + f(r.q0[3])) + (f(r.q1[1]) + f(r.q1[3])));
    red_slot_store(slot, acc);
}
template <typename Real>
__device__ __forceinline__ void red_wait(RedQ<Real> &r) {
    tmem::wait_ld_pair(r.q0, r.q1);
}
template <typename Real, int NW>
__device__ __forceinline__ void red_wait(uint32_t (&tw)[NW], RedQ<Real> &r) {
    tmem::wait_ld_three(tw, r.q0, r.q1);
}

// Replay-only pass of a short-wave band run (the last segment's chain in the direction it has not been swept
// in yet), in two phases.  Sweep: ER *= t read back from the t-store, and the value of every layer goes straight
// back into the t-store columns its t came from (nobody needs that t again; the next run's first pass replays
// with weight 0 and only wants finite numbers there).  Lane sums: the whole t-store is then read back with the
// transposing loads, block after block with nothing serial between them, into the class accumulators.
template <typename Real, bool UP, bool SMALL>
__device__ __forceinline__ void sw_replay_pass(const int Lp, const Real wR, const Real b0, const TStore<Real> tst,
                                               const RedAux ra) {
    constexpr int WPR = sizeof(Real) / 4, NW = kU * WPR;
    const int kfirst = UP ? Lp - kU : 0, kstep = UP ? -kU : kU, nb = Lp / kU;
    Real ER = wR * b0;
    uint32_t tw[NW];
    tmem::ld<NW>(tst.taddr + (uint32_t)(kfirst * WPR), tw);
    for (int it = 0, k0 = kfirst; it < nb; ++it, k0 += kstep) {
        Real tR[kU], v[kU];
        tmem::wait_ld<NW>(tw);
        tmem::unpack<kU>(tw, tR);
        if (it + 1 < nb) tmem::ld<NW>(tst.taddr + (uint32_t)((k0 + kstep) * WPR), tw);
#pragma unroll
        for (int o = 0; o < kU; ++o) {
            const int u = UP ? kU - 1 - o : o;
            ER *= tR[u];
            v[u] = ER;
        }
        tm_store_block<SMALL>(tst.taddr + (uint32_t)(k0 * WPR), v);
    }
    tmem::wait_st();
    const unsigned base = UP ? ra.up : ra.dn;
    RedQ<Real> q;
    red_fetch_issued(tst.taddr, q);
    for (int b = 0; b < nb; ++b) {
        const double2 acc = red_slot_load(base + b * 512);
        red_wait(q);
        const RedQ<Real> p = q;
        if (b + 1 < nb) red_fetch_issued(tst.taddr + (uint32_t)((b + 1) * NW), q);
        red_accumulate(p, acc, base + b * 512);
    }
}

// ---- short-wave fresh pass, direct form (kVarSwDirect) -----------------------------------------------
// Integer helpers written as PTX so that ptxas keeps them on the ALU pipe in the shape intended.
__device__ __forceinline__ double xl_lookup(int n, unsigned saddr_lane) {
    // entry j = n & 511 of the replicated table: byte offset j * 128 inside the 64 kB window the table fills
    unsigned s, a;
    asm("shf.l.wrap.b32 %0, %1, %1, 7;" : "=r"(s) : "r"(n));
    asm("lop3.b32 %0, %1, 0xff80, %2, 0xEA;" : "=r"(a) : "r"(s), "r"(saddr_lane));      // (s & 0xff80) | saddr_lane
    double T;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(T) : "r"(a));
    return T;
}
__device__ __forceinline__ double xl_scale(double tt, int n) {
    // 2^(j/512) 2^(n >> 9) from table entry j = n & 511.  The exponent field has to gain (n >> 9) << 20 =
    // n * 2^11 - j * 2^11; the entries are stored with j * 2^11 already taken off their high word
    // (xl_table_entry), so one multiply-add with the whole n does it -- no masking of n's low bits.
    int hi;
    asm("mad.lo.s32 %0, %1, 2048, %2;" : "=r"(hi) : "r"(n), "r"(__double2hiint(tt)));
    return __hiloint2double(hi, __double2loint(tt));
}

// One downward (UP = false) or upward pass of a short-wave segment pair: fresh chain of this segment (rows
// src_lane, weight wF) and replay chain of the previous one (weight wR, its t read back from the t-store in
// tensor memory).  Same conventions as sweep_pass (level tables, carried interface, t-store addressing, band
// accumulators), but a block of kU layers is finished inside its iteration:
//   table values of the block (loaded one block ahead) -> layer integrands -> thicknesses -> exp ->
//   both recurrences -> park t;  then the next block's loads go out into the same registers.
// Only the lane sum of the block's eight values is carried into the next iteration: five shuffle stages
// interleaved with the next block's exp, or (kVarRedTmem) the transpose through tensor memory whose loads are
// issued next to the replay chain's and complete behind the exp.
// The table factor 2^(j/512) gets the 2^n of the result added to its exponent field BEFORE the last FMA, so
// that FMA writes t straight into the register pair the t-store block and the recurrences read.
// Level data by byte offset from the CTA's dynamic shared array: the address space is then known at every load
// (LDS with the array's link-time base as immediate); through the generic pointers of LevelView the compiler
// rebuilds the shared window base (S2UR of the CTA id + shifts) in front of every group of loads.
extern __shared__ __align__(16) unsigned char smem_raw[];
__device__ __forceinline__ void lds_real8(unsigned off, double (&o)[kU]) {
#pragma unroll
    for (int i = 0; i < kU / 2; ++i) {
        const double2 q = *reinterpret_cast<const double2 *>(smem_raw + off + 16u * i);
        o[2 * i] = q.x, o[2 * i + 1] = q.y;
    }
}
__device__ __forceinline__ void lds_real8(unsigned off, float (&o)[kU]) {
#pragma unroll
    for (int i = 0; i < kU / 4; ++i) {
        const float4 q = *reinterpret_cast<const float4 *>(smem_raw + off + 16u * i);
        o[4 * i] = q.x, o[4 * i + 1] = q.y, o[4 * i + 2] = q.z, o[4 * i + 3] = q.w;
    }
}
// The same for a run that starts one element before a 16-byte boundary.
__device__ __forceinline__ void lds_real8_m1(unsigned off, double (&o)[kU]) {
    o[0] = *reinterpret_cast<const double *>(smem_raw + off);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double2 q = *reinterpret_cast<const double2 *>(smem_raw + off + 8u + 16u * i);
        o[1 + 2 * i] = q.x, o[2 + 2 * i] = q.y;
    }
    o[7] = *reinterpret_cast<const double *>(smem_raw + off + 56u);
}
__device__ __forceinline__ void lds_real8_m1(unsigned off, float (&o)[kU]) {
    o[0] = *reinterpret_cast<const float *>(smem_raw + off);
    const float4 q = *reinterpret_cast<const float4 *>(smem_raw + off + 4u);
    o[1] = q.x, o[2] = q.y, o[3] = q.z, o[4] = q.w;
    const float2 w = *reinterpret_cast<const float2 *>(smem_raw + off + 20u);
    o[5] = w.x, o[6] = w.y;
    o[7] = *reinterpret_cast<const float *>(smem_raw + off + 28u);
}
__device__ __forceinline__ void lds_u32x8(unsigned off, unsigned (&o)[kU]) {
#pragma unroll
    for (int i = 0; i < kU / 4; ++i) {
        const uint4 q = *reinterpret_cast<const uint4 *>(smem_raw + off + 16u * i);
        o[4 * i] = q.x, o[4 * i + 1] = q.y, o[4 * i + 2] = q.z, o[4 * i + 3] = q.w;
    }
}

// exp(-x) of a block for the direct pass.  FP64: the shared-memory table routine (see the variant list), eight
// evaluations step by step; with RED the five shuffle stages of the previous block's lane sum are issued in between.
// FP32: ex2.approx, as everywhere in FP32 mode.
// (Every `if constexpr` around a `#pragma unroll` loop is braced: unbraced, the discarded statement of a false condition
// was the pragma, not the loop, and with nvcc 12.9 the pass came back with its arithmetic folded to NaN.)
template <int V, bool RED>
__device__ __forceinline__ void direct_exp(double (&x)[kU], double (&t)[kU], const ExpAux xaux, const int lane,
                                           const double (&v)[kU], double &c1) {
    constexpr double kK = 0x1.71547652b82fep+9, kMagic = 6755399441055744.0, kC = -0x1.62e42fefa39efp-10;
    const unsigned full = 0xffffffffu;
    const bool h4 = lane & 16, h3 = lane & 8, h2 = lane & 4;
    double a4[4], a2[2];
    double tn[kU], r[kU], r2[kU], p[kU], Tj[kU];
    int nn[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
        int hi = __double2hiint(x[u]);
        asm("min.s32 %0, %0, 0x40862000;" : "+r"(hi));            // thickness clamp (clamp_thickness), in place
        x[u] = __hiloint2double(hi, __double2loint(x[u]));
        tn[u] = fma(-x[u], kK, kMagic);
    }
    if constexpr (RED) {                                             // reduction step 1 (xor 16)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const double keep = h4 ? v[i + 4] : v[i], send = h4 ? v[i] : v[i + 4];
            a4[i] = keep + __shfl_xor_sync(full, send, 16);
        }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
        nn[u] = __double2loint(tn[u]);
        tn[u] -= kMagic;
        Tj[u] = xl_scale(xl_lookup(nn[u], xaux.saddr), nn[u]);       // 2^(j/512) 2^n
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) r[u] = fma(tn[u], kC, -x[u]);
    if constexpr (RED) {                                             // step 2 (xor 8)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const double keep = h3 ? a4[i + 2] : a4[i], send = h3 ? a4[i] : a4[i + 2];
            a2[i] = keep + __shfl_xor_sync(full, send, 8);
        }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
        r2[u] = r[u] * r[u];
        p[u] = fma(r[u], 1.0 / 6.0, 0.5);
    }
    if constexpr (RED) {                                             // step 3 (xor 4)
        const double keep = h2 ? a2[1] : a2[0], send = h2 ? a2[0] : a2[1];
        c1 = keep + __shfl_xor_sync(full, send, 4);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) p[u] = fma(r2[u], p[u], r[u]);
    if constexpr (RED) c1 += __shfl_xor_sync(full, c1, 2);           // step 4
#pragma unroll
    for (int u = 0; u < kU; ++u) t[u] = fma(Tj[u], p[u], Tj[u]);
    if constexpr (RED) c1 += __shfl_xor_sync(full, c1, 1);           // step 5
}
template <int V, bool RED>
__device__ __forceinline__ void direct_exp(float (&x)[kU], float (&t)[kU], const ExpAux, const int lane,
                                           const float (&v)[kU], float &c1) {
    // exp(-x) = 2^(-x log2 e): one multiply and one MUFU.EX2.  (__expf wraps ex2.approx in a range test and two more
    // multiplies so that denormal RESULTS survive; a transmissivity below 1e-38 is zero at any tolerance of this mode.)
#pragma unroll
    for (int u = 0; u < kU; ++u) asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t[u]) : "f"(x[u] * -1.4426950408889634f));
    if constexpr (RED) c1 = reduce8<float>(v, lane);
}

template <typename Real, int MM, bool UP, int V>
__device__ __forceinline__ void sw_fresh_pass(const int lane, const int nl, const int Lp, const Real wF, const Real wR,
                                              const Real b0, const void *src_lane, const void *next_src,
                                              PassFeed<Real, MM> &fd, const LevelView<Real> lv,
                                              const TStore<Real> tst, double *bacc, const ExpAux xaux, const RedAux ra) {
    static_assert(MM > 0, "table-fused path");
    constexpr bool RT = (V & kVarRedTmem) != 0, SMALL = (V & kVarSmallSt) != 0;
    constexpr int WPR = sizeof(Real) / 4, NW = kU * WPR;
    constexpr unsigned kRB = sizeof(Real);
    const int ridx = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
    const bool rlane = (lane & 3) == 0;
    const int kfirst = UP ? Lp - kU : 0, kstep = UP ? -kU : kU, nb = Lp / kU;
    const char *src_bytes = (const char *)src_lane;
    // byte offsets of the level tables in the shared array: row offsets of this direction, q/g, half thicknesses
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned s_offs = (unsigned)__cvta_generic_to_shared(UP ? lv.off_up : lv.off_dn) - s0;
    const unsigned s_qg = (unsigned)__cvta_generic_to_shared(lv.qg) - s0, s_hdp = (unsigned)__cvta_generic_to_shared(lv.hdp) - s0;

    Real EF = wF * b0, ER = wR * b0;
    Real Ic = Real(0);
#pragma unroll
    for (int m = 0; m < MM; ++m) Ic = fma(lv.qg[m * lv.stride + (UP ? Lp : 0) + kLevShift], fd.icraw[m], Ic);

    Real v[kU], c1 = Real(0);
#pragma unroll
    for (int u = 0; u < kU; ++u) v[u] = Real(0);
    int kprev = -kU * 4;        // block whose values v holds (none yet: its accumulate is predicated off)
    // kVarRedTmem: class accumulator the pending lane sums go to; the first accumulate adds the zeros parked here
    const unsigned cbase = UP ? ra.up : ra.dn;
    unsigned slot = cbase + kfirst * 64;
    if constexpr (RT) red_park<SMALL>(v, ra.scratch);

    // (Tried and dropped: a second buffer of table values, i.e. loads issued two blocks ahead, the loop unrolled by
    // two -- 9 % slower.)
    for (int it = 0, k0 = kfirst; it < nb; ++it, k0 += kstep) {
        auto &raw = fd.raw;
        // replay chain's t of this block
        uint32_t tRw[NW];
        tmem::ld<NW>(tst.taddr + (uint32_t)(k0 * WPR), tRw);
        // band accumulator of the block being reduced (read early, written at the end)
        double bold = 0.0;
        const bool acc_on = !RT && rlane && kprev >= 0 && kprev + ridx < nl;
        if (acc_on) bold = bacc[kprev + ridx];

        // ---- stage 1: integrand at the block's loaded interfaces, thickness of its layers ------------
        Real In[kU], x[kU];
#pragma unroll
        for (int m = 0; m < MM; ++m) {
            Real qg[kU];
            if (UP)
                lds_real8_m1(s_qg + (unsigned)(m * lv.stride + k0 + kLevShift) * kRB, qg);
            else
                lds_real8(s_qg + (unsigned)(m * lv.stride + k0 + 1 + kLevShift) * kRB, qg);
#pragma unroll
            for (int u = 0; u < kU; ++u) In[u] = m ? fma(qg[u], raw[u][m], In[u]) : qg[u] * raw[u][0];
        }
        {
            Real hd[kU];
            lds_real8(s_hdp + (unsigned)k0 * kRB, hd);
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const Real lo = UP ? In[u] : (u ? In[u - 1] : Ic);
                const Real hi = UP ? (u < kU - 1 ? In[u + 1] : Ic) : In[u];
                x[u] = hd[u] * (lo + hi);
            }
            Ic = UP ? In[0] : In[kU - 1];
        }
        // ---- the next block's table values (or the next pass's first block) into the same registers ----
        if (it + 1 < nb) {
#pragma unroll
            for (int m = 0; m < MM; ++m) {
                unsigned off[kU];
                lds_u32x8(s_offs + (unsigned)(m * lv.stride + k0 + kstep + 4) * 4u, off);
#pragma unroll
                for (int u = 0; u < kU; ++u) raw[u][m] = *(const Real *)(src_bytes + (size_t)off[u] * 16u);
            }
        } else if (next_src) {
            prefetch_pass<Real, MM>(!UP, Lp, next_src, lv, fd);
        }
        // ---- kVarRedTmem: the previous block's values, transposed, are on their way behind the exp ------
        RedQ<Real> rq;
        double2 cacc = make_double2(0.0, 0.0);
        if constexpr (RT) {
            red_fetch(ra.scratch, rq);
            cacc = red_slot_load(slot);
        }

        // ---- exp of the block's eight layers (with the previous block's shuffle lane sum in between) ----
        Real t[kU];
        constexpr bool RED = !RT;
        direct_exp<V, RED>(x, t, xaux, lane, v, c1);

        // ---- both recurrences in sweep order; the block's values for the next iteration's lane sum ------
        Real tR[kU];
        if constexpr (RT)
            red_wait(tRw, rq);
        else
            tmem::wait_ld<NW>(tRw);
        tmem::unpack<kU>(tRw, tR);
        if constexpr (RT) red_accumulate(rq, cacc, slot);
#pragma unroll
        for (int o = 0; o < kU; ++o) {
            const int u = UP ? kU - 1 - o : o;
            EF *= t[u];
            ER *= tR[u];
            v[u] = EF + ER;
        }
        tm_store_block<SMALL>(tst.taddr + (uint32_t)(k0 * WPR), t);
        if (acc_on) bacc[kprev + ridx] = bold + (double)c1;
        kprev = k0;
        if constexpr (RT) {
            red_park<SMALL>(v, ra.scratch);
            slot = cbase + k0 * 64;
        }
    }
    // ---- lane sum of the last block -------------------------------------------------------------------
    if constexpr (RT) {
        RedQ<Real> rq;
        red_fetch(ra.scratch, rq);                  // its wait::st also covers the t-store: the next pass reads it back
        const double2 cacc = red_slot_load(slot);
        red_wait(rq);
        red_accumulate(rq, cacc, slot);
    } else {
        const Real c = reduce8<Real>(v, lane);
        if (rlane && kprev + ridx < nl) bacc[kprev + ridx] += (double)c;
        tmem::wait_st();
    }
}

// ---- FP32, two segments per pass (kVarPair) -------------------------------------------------------------
// The FP32 passes are bound by instruction issue, not by any pipe.  Blackwell's packed FP32 instructions (FFMA2 /
// FMUL2 / FADD2: two FP32 operations on a register pair, a 32-bit operand is broadcast for free) halve the issue
// slots of the arithmetic, so a lane carries the same chain of TWO consecutive segments of a band (.x = segment s,
// .y = segment s + 1: in segment space its table values sit 128 bytes apart, one address for both loads).  Per
// layer and lane everything is then two words wide -- the t-store, the parked values, the transposing loads are
// those of the FP64 kernel -- and a warp needs FP64's share of tensor memory: 8 warps per SM instead of 12.
using f2 = float2;
__device__ __forceinline__ f2 f2_set(float a) { return make_float2(a, a); }
__device__ __forceinline__ f2 f2_mul(f2 a, f2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ f2 f2_add(f2 a, f2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ f2 f2_fma(f2 a, f2 b, f2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ f2 f2_ld(const char *p) {      // segments s and s + 1, this lane
    return make_float2(*(const float *)p, *(const float *)(p + 128));
}
template <int MM>
struct PairFeed {
    f2 raw[kU][MM > 0 ? MM : 1];
    f2 icraw[MM > 0 ? MM : 1];
};
template <int MM>
__device__ __forceinline__ void prefetch_pair_pass(const bool up, const int Lp, const void *src_lane,
                                                   const LevelView<float> &lv, PairFeed<MM> &fd) {
    const char *src_bytes = (const char *)src_lane;
    const unsigned *offs = up ? lv.off_up : lv.off_dn;
    const int k0 = up ? Lp - kU : 0, kc = (up ? Lp : 0) + kLevShift;
#pragma unroll
    for (int m = 0; m < MM; ++m) {
        unsigned off[kU];
        lds_block(offs + m * lv.stride + k0 + 4, off);
#pragma unroll
        for (int u = 0; u < kU; ++u) fd.raw[u][m] = f2_ld(src_bytes + (size_t)off[u] * 16u);
        fd.icraw[m] = f2_ld(src_bytes + (size_t)lv.off_dn[m * lv.stride + kc] * 16u);
    }
}
template <bool SMALL>
__device__ __forceinline__ void tm_store_block(uint32_t addr, const f2 (&v)[kU]) {
    if constexpr (SMALL) {
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            uint32_t q[2] = {__float_as_uint(v[u].x), __float_as_uint(v[u].y)};
            tmem::st_small<2>(addr + (uint32_t)(2 * u), q);
        }
    } else {
        uint32_t w[2 * kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) w[2 * u] = __float_as_uint(v[u].x), w[2 * u + 1] = __float_as_uint(v[u].y);
        tmem::st<2 * kU>(addr, w);
    }
}
__device__ __forceinline__ void unpack_pairs(const uint32_t (&w)[2 * kU], f2 (&v)[kU]) {
#pragma unroll
    for (int u = 0; u < kU; ++u) v[u] = make_float2(__uint_as_float(w[2 * u]), __uint_as_float(w[2 * u + 1]));
}
// The transposed words are those of an FP64 block (RedQ<double>): value u / u + 4 of lanes c, c + 8 | c + 16, c + 24,
// each value the pair of the lane's two segments.  Four lanes and both segments in FP32, FP64 from there on.
__device__ __forceinline__ void red_accumulate_pairs(const RedQ<double> &r, double2 acc, unsigned slot) {
    auto f = [](const uint32_t(&q)[8], int i) { return make_float2(__uint_as_float(q[2 * i]), __uint_as_float(q[2 * i + 1])); };
    const f2 lo = f2_add(f2_add(f(r.q0, 0), f(r.q0, 1)), f2_add(f(r.q1, 0), f(r.q1, 1)));
    const f2 hi = f2_add(f2_add(f(r.q0, 2), f(r.q0, 3)), f2_add(f(r.q1, 2), f(r.q1, 3)));
    acc.x += (double)(lo.x + lo.y);
    acc.y += (double)(hi.x + hi.y);
    red_slot_store(slot, acc);
}

// sw_fresh_pass for a segment pair (always with the tensor-memory lane sum).
template <int MM, bool UP, int V>
__device__ __forceinline__ void sw_fresh_pair_pass(const int Lp, const f2 wF, const f2 wR, const float b0,
                                                   const void *src_lane, const void *next_src, PairFeed<MM> &fd,
                                                   const LevelView<float> lv, const uint32_t taddr, const RedAux ra) {
    static_assert(MM > 0, "table-fused path");
    constexpr bool SMALL = (V & kVarSmallSt) != 0;
    constexpr int NW = 2 * kU;
    const int kfirst = UP ? Lp - kU : 0, kstep = UP ? -kU : kU, nb = Lp / kU;
    const char *src_bytes = (const char *)src_lane;
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(smem_raw);
    const unsigned s_offs = (unsigned)__cvta_generic_to_shared(UP ? lv.off_up : lv.off_dn) - s0;
    const unsigned s_qg = (unsigned)__cvta_generic_to_shared(lv.qg) - s0, s_hdp = (unsigned)__cvta_generic_to_shared(lv.hdp) - s0;

    f2 EF = f2_mul(wF, f2_set(b0)), ER = f2_mul(wR, f2_set(b0));
    f2 Ic = f2_set(0.f);
#pragma unroll
    for (int m = 0; m < MM; ++m) Ic = f2_fma(f2_set(lv.qg[m * lv.stride + (UP ? Lp : 0) + kLevShift]), fd.icraw[m], Ic);

    f2 v[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) v[u] = f2_set(0.f);
    const unsigned cbase = UP ? ra.up : ra.dn;
    unsigned slot = cbase + kfirst * 64;
    tm_store_block<SMALL>(ra.scratch, v);

    for (int it = 0, k0 = kfirst; it < nb; ++it, k0 += kstep) {
        auto &raw = fd.raw;
        uint32_t tRw[NW];
        tmem::ld<NW>(taddr + (uint32_t)(k0 * 2), tRw);
        // ---- integrand at the block's loaded interfaces, thickness of its layers --------------------------
        f2 In[kU], x[kU];
#pragma unroll
        for (int m = 0; m < MM; ++m) {
            float qg[kU];
            if (UP)
                lds_real8_m1(s_qg + (unsigned)(m * lv.stride + k0 + kLevShift) * 4u, qg);
            else
                lds_real8(s_qg + (unsigned)(m * lv.stride + k0 + 1 + kLevShift) * 4u, qg);
#pragma unroll
            for (int u = 0; u < kU; ++u) In[u] = m ? f2_fma(f2_set(qg[u]), raw[u][m], In[u]) : f2_mul(f2_set(qg[u]), raw[u][0]);
        }
        {
            float hd[kU];
            lds_real8(s_hdp + (unsigned)(Lp + k0) * 4u, hd);        // half layer thickness times -log2 e (second half of the array)
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const f2 lo = UP ? In[u] : (u ? In[u - 1] : Ic);
                const f2 hi = UP ? (u < kU - 1 ? In[u + 1] : Ic) : In[u];
                x[u] = f2_mul(f2_set(hd[u]), f2_add(lo, hi));       // -thickness * log2 e
            }
            Ic = UP ? In[0] : In[kU - 1];
        }
        // ---- the next block's table values (or the next pass's first block) into the same registers ----------
        if (it + 1 < nb) {
#pragma unroll
            for (int m = 0; m < MM; ++m) {
                unsigned off[kU];
                lds_u32x8(s_offs + (unsigned)(m * lv.stride + k0 + kstep + 4) * 4u, off);
#pragma unroll
                for (int u = 0; u < kU; ++u) raw[u][m] = f2_ld(src_bytes + (size_t)off[u] * 16u);
            }
        } else if (next_src) {
            prefetch_pair_pass<MM>(!UP, Lp, next_src, lv, fd);
        }
        // ---- the previous block's values, transposed, are on their way behind the exp -------------------------
        RedQ<double> rq;
        red_fetch(ra.scratch, rq);
        const double2 cacc = red_slot_load(slot);
        f2 t[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t[u].x) : "f"(x[u].x));
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t[u].y) : "f"(x[u].y));
        }
        // ---- both recurrences in sweep order; the block's values for the next iteration's lane sum -------------
        f2 tR[kU];
        red_wait(tRw, rq);
        unpack_pairs(tRw, tR);
        red_accumulate_pairs(rq, cacc, slot);
#pragma unroll
        for (int o = 0; o < kU; ++o) {
            const int u = UP ? kU - 1 - o : o;
            EF = f2_mul(EF, t[u]);
            ER = f2_mul(ER, tR[u]);
            v[u] = f2_add(EF, ER);
        }
        tm_store_block<SMALL>(taddr + (uint32_t)(k0 * 2), t);
        tm_store_block<SMALL>(ra.scratch, v);
        slot = cbase + k0 * 64;
    }
    RedQ<double> rq;
    red_fetch(ra.scratch, rq);                      // its wait::st also covers the t-store: the next pass reads it back
    const double2 cacc = red_slot_load(slot);
    red_wait(rq);
    red_accumulate_pairs(rq, cacc, slot);
}

// sw_replay_pass for a segment pair.
template <bool UP, bool SMALL>
__device__ __forceinline__ void sw_replay_pair_pass(const int Lp, const f2 wR, const float b0, const uint32_t taddr,
                                                    const RedAux ra) {
    constexpr int NW = 2 * kU;
    const int kfirst = UP ? Lp - kU : 0, kstep = UP ? -kU : kU, nb = Lp / kU;
    f2 ER = f2_mul(wR, f2_set(b0));
    uint32_t tw[NW];
    tmem::ld<NW>(taddr + (uint32_t)(kfirst * 2), tw);
    for (int it = 0, k0 = kfirst; it < nb; ++it, k0 += kstep) {
        f2 tR[kU], v[kU];
        tmem::wait_ld<NW>(tw);
        unpack_pairs(tw, tR);
        if (it + 1 < nb) tmem::ld<NW>(taddr + (uint32_t)((k0 + kstep) * 2), tw);
#pragma unroll
        for (int o = 0; o < kU; ++o) {
            const int u = UP ? kU - 1 - o : o;
            ER = f2_mul(ER, tR[u]);
            v[u] = ER;
        }
        tm_store_block<SMALL>(taddr + (uint32_t)(k0 * 2), v);
    }
    tmem::wait_st();
    const unsigned base = UP ? ra.up : ra.dn;
    RedQ<double> q;
    red_fetch_issued(taddr, q);
    for (int b = 0; b < nb; ++b) {
        const double2 acc = red_slot_load(base + b * 512);
        red_wait(q);
        const RedQ<double> p = q;
        if (b + 1 < nb) red_fetch_issued(taddr + (uint32_t)((b + 1) * NW), q);
        red_accumulate_pairs(p, acc, base + b * 512);
    }
}

template <typename Real, int MM, bool FRESH, int V>
__device__ __forceinline__ void dispatch_pass(const bool lw, const bool up, const int lane, const int nl, const int Lp,
                                              const Real wF, const Real wR, const Real bs, const Real bg,
                                              const void *src_lane, const void *next_src, PassFeed<Real, MM> &fd,
                                              const LevelView<Real> &lv, const TStore<Real> tst, double *bacc_up, double *bacc_dn,
                                              const Real *coef, const ExpAux xaux, const RedAux ra) {
    // coefficient rows of a band: [0] dn_last, [1] dn_gen, [2] up_first, [3] up_gen.  The value of layer k
    // of a downward pass belongs to interface k + 1 (the plain accumulators are indexed by interface).
    double *const dn = bacc_dn + 1;
    if (lw) {
        if (up)
            sweep_pass<Real, MM, true, true, FRESH, V>(lane, nl, Lp, wF, wR, bg, src_lane, next_src, fd, lv, tst, bacc_up, coef + 2 * Lp, coef + 3 * Lp, xaux);
        else
            sweep_pass<Real, MM, true, false, FRESH, V>(lane, nl, Lp, wF, wR, bs, src_lane, next_src, fd, lv, tst, dn, coef, coef + Lp, xaux);
    } else {
        if constexpr (FRESH && MM > 0 && (V & kVarSwDirect) != 0) {
            if (up)
                sw_fresh_pass<Real, MM, true, V>(lane, nl, Lp, wF, wR, bg, src_lane, next_src, fd, lv, tst, bacc_up, xaux, ra);
            else
                sw_fresh_pass<Real, MM, false, V>(lane, nl, Lp, wF, wR, bs, src_lane, next_src, fd, lv, tst, dn, xaux, ra);
        } else if constexpr (!FRESH && MM > 0 && (V & kVarSwDirect) != 0 && (V & kVarRedTmem) != 0) {
            if (next_src) prefetch_pass<Real, MM>(!up, Lp, next_src, lv, fd);
            if (up)
                sw_replay_pass<Real, true, (V & kVarSmallSt) != 0>(Lp, wR, bg, tst, ra);
            else
                sw_replay_pass<Real, false, (V & kVarSmallSt) != 0>(Lp, wR, bs, tst, ra);
        } else if (up)
            sweep_pass<Real, MM, false, true, FRESH, V>(lane, nl, Lp, wF, wR, bg, src_lane, next_src, fd, lv, tst, bacc_up, coef, coef, xaux);
        else
            sweep_pass<Real, MM, false, false, FRESH, V>(lane, nl, Lp, wF, wR, bs, src_lane, next_src, fd, lv, tst, dn, coef, coef, xaux);
    }
}

// MM = number of molecules of the table-fused path (compile-time so the prefetched table values
// live in registers); MM = 0 selects the resident-tau path.  Persistent CTAs: task = (column, c)
// with c < ctas_per_col; warp w of the CTA sweeps part c * W + w of the column.
template <typename Real, int MM, int V = 0>
__global__ void __launch_bounds__(variant_threads<Real>(V), 1) flux_kernel(const FluxArgs<Real> a) {
    constexpr bool FUSED = MM > 0;
    constexpr int MR = MM > 0 ? MM : 1;
    constexpr bool TM = (V & kVarTmem) != 0, CS = (V & kVarCoefSmem) != 0;
    constexpr bool PAIR = sizeof(Real) == 4 && MM > 0 && (V & kVarPair) != 0;
    static_assert(!(V & kVarPair) || (PAIR && (V & kVarRedTmem) && (V & kVarSwDirect)), "kVarPair: FP32 table-fused direct passes");
    constexpr int WPR = variant_tstore_words<Real>(V), NW = kU * WPR;
    const int tid = threadIdx.x, nthr = blockDim.x, W = nthr >> 5, warp = tid >> 5, lane = tid & 31;
    const int L = a.L, nl = a.L - 1, Lp = padded_layers(a.L), LS = Lp + kU, C = a.ctas_per_col;

    const CtaSmem<Real, MM, V> lay(L, -1, a.B);
    unsigned *lev_offdn = (unsigned *)(smem_raw + lay.off_offdn);
    unsigned *lev_offup = (unsigned *)(smem_raw + lay.off_offup);
    Real *lev_qg = (Real *)(smem_raw + lay.off_qg);
    Real *hdp = (Real *)(smem_raw + lay.off_hdp);
    const int lstride = lay.lev_stride;
    unsigned char *wbase = smem_raw + lay.shared_total + (size_t)warp * lay.warp_total;
    double *bacc_up = (double *)(wbase + lay.warp_bacc);
    double *bacc_dn = bacc_up + LS;
    const int n_bacc = 2 * LS;                                    // accumulator doubles per warp
    constexpr bool RT = (V & kVarRedTmem) != 0;
    constexpr unsigned kXlWindow = xl_window(V);
    const unsigned smem_window0 = (unsigned)__cvta_generic_to_shared(smem_raw);      // shared-window address of smem_raw
    // staged coefficient rows (kVarCoefSmem): in the warp's region, or (kVarRedTmem) above the exp table
    Real *cw = (RT && sizeof(Real) == 8) ? (Real *)(smem_raw + (kXlWindow + kXlBytes - smem_window0) + (size_t)warp * lay.coef_bytes)
                                         : (Real *)(wbase + lay.warp_coef);
    Real *ts = (Real *)(wbase + lay.warp_ts);
    double *sacc = (double *)(smem_raw + lay.shared_total + lay.warp_bacc);   // warp 0's accumulators, free in the epilogue
    Real *coef_cta = a.coefbuf + (size_t)blockIdx.x * a.B * 4 * Lp;
    const LevelView<Real> lv{lev_offdn, lev_offup, lev_qg, lstride, hdp};

    for (int k = tid; k < Lp; k += nthr) {
        hdp[k] = k < nl ? (Real)a.hdP[k] : Real(0);
        if constexpr (PAIR) hdp[Lp + k] = hdp[k] * Real(-1.4426950408889634);
    }
    for (int i = lane; i < n_bacc; i += 32) bacc_up[i] = 0.0;
    // The t-store: the first replay of a run reads it with weight 0, so it has to hold finite numbers.
    __shared__ uint32_t s_tmem_base;
    TStore<Real> tst{ts + lane, 0u};
    // the replicated 2^(j/512) table (kVarExpLds)
    double *xtab = (double *)(smem_raw + lay.off_xtab);
    if constexpr ((V & kVarSwDirect) != 0 && sizeof(Real) == 8)      // the table sits at shared-window address kXlWindow (launch_flux sizes the window)
        xtab = (double *)(smem_raw + (kXlWindow - smem_window0));
    if constexpr ((V & (kVarExpLds | kVarSwDirect)) != 0 && sizeof(Real) == 8)
        for (int i = tid; i < kXlN * kXlCopies; i += nthr) xtab[i] = a.exp_tab[i / kXlCopies];
    const ExpAux xaux{xtab + (lane & (kXlCopies - 1)), kXlWindow | ((unsigned)(lane & (kXlCopies - 1)) << 3)};
    // kVarRedTmem: this warp's class accumulators [2][Lp / kU][4][8] double2 and its 16 exchange columns in tensor memory
    RedAux ra{0u, 0u, 0u};
    double2 *pacc = nullptr;
    const int n_pacc = 2 * Lp * 4;                  // double2 per warp, both directions
    if constexpr (RT) {
        pacc = (double2 *)(wbase + lay.warp_pacc);
        ra.up = (unsigned)__cvta_generic_to_shared(pacc + lane);
        ra.dn = ra.up + (unsigned)Lp * 64u;
        for (int i = lane; i < n_pacc; i += 32) pacc[i] = make_double2(0.0, 0.0);
    }
    if constexpr (TM) {
        if (warp == 0) tmem::acquire(&s_tmem_base, (uint32_t)a.tmem_cols);
        tmem::fence_before_sync();
        __syncthreads();
        tmem::fence_after_sync();
        // (shuffled from lane 0 so that ptxas knows the addresses are warp-uniform: every LDTM / STTM then takes its
        // address from one uniform register + immediate instead of a fresh R2UR each)
        tst.taddr = __shfl_sync(0xffffffffu, tmem::address(s_tmem_base, warp, (warp >> 2) * a.tmem_cpw), 0);
        if constexpr (RT)
            ra.scratch = __shfl_sync(0xffffffffu, tmem::address(s_tmem_base, warp, (W >> 2) * a.tmem_cpw + (warp >> 2) * 2 * kU), 0);
        uint32_t zero[NW];
#pragma unroll
        for (int i = 0; i < NW; ++i) zero[i] = 0u;
        for (int k0 = 0; k0 < Lp; k0 += kU) tmem::st<NW>(tst.taddr + (uint32_t)(k0 * WPR), zero);
        tmem::wait_st();
    } else {
        for (int i = lane; i < Lp * 32; i += 32) ts[i] = Real(0);
    }

    // Static band data the column prologue walks, in registers for up to 64 bands: which bands are long-wave (a bit
    // mask) and the band centres (lane l holds bands l and l + 32, broadcast with shuffles).  Read through global
    // memory they cost an L2 round trip each and one after the other -- the CTA's shared memory leaves almost no L1.
    const bool meta_regs = a.B <= 64;
    unsigned long long lwmask = 0ULL;
    double bK0 = 1.0, bK1 = 1.0, bN0 = 0.0, bN1 = 0.0;          // Planck constants of bands lane, lane + 32
    if (meta_regs) {
        const bool l0 = lane < a.B && !a.band_sw[lane < a.B ? lane : 0], l1 = lane + 32 < a.B && !a.band_sw[lane + 32 < a.B ? lane + 32 : 0];
        lwmask = (unsigned long long)__ballot_sync(0xffffffffu, l0) | ((unsigned long long)__ballot_sync(0xffffffffu, l1) << 32);
        if (lane < a.B) planck_band_constants(a.pc, a.band_centre[lane], bK0, bN0);
        if (lane + 32 < a.B) planck_band_constants(a.pc, a.band_centre[lane + 32], bK1, bN1);
    }
    auto band_is_sw = [&](int j) { return meta_regs ? !((lwmask >> j) & 1ULL) : (bool)__ldg(a.band_sw + j); };
    auto band_KN = [&](int j, double &K, double &N) {             // whole warp, uniform j
        if (meta_regs) {
            K = shfl_f64(j < 32 ? bK0 : bK1, j & 31);
            N = shfl_f64(j < 32 ? bN0 : bN1, j & 31);
        } else {
            planck_band_constants(a.pc, __ldg(a.band_centre + j), K, N);
        }
    };

    // Tasks are handed out through a global counter: SMs differ a little in speed (L2 distance,
    // table rows hit), and a static split would leave the fast ones idle at the end.
    __shared__ unsigned long long s_task;
    for (int n_done = 0;; ++n_done) {
        __syncthreads();   // the previous task is done with the tables, the scratch, sacc and s_task
        if (tid == 0) s_task = atomicAdd(a.task_counter, 1ULL);
        __syncthreads();
        const int64_t task = (int64_t)s_task;
        if (task >= a.n_tasks) break;
        const int64_t col = task / C;
        const int cpart = (int)(task - col * C);
        const int64_t region = C == 1 ? (int64_t)blockIdx.x : col;
        const long long t_task0 = a.dbg ? clock64() : 0;

        // The inputs of the column this CTA is likely to get next (tasks are handed out in order, so about gridDim.x
        // further on) are pulled into L2 now: the prologue is a chain of dependent global loads, and from HBM each
        // link costs three times what it costs from L2.  A wrong guess only wastes a few 128-byte lines.
        if (C == 1) {
            const int64_t nc = col + gridDim.x;
            if (nc < a.n_col) {
                auto pull = [&](const void *base, int bytes) {
                    for (int off = tid * 128; off < bytes; off += nthr * 128)
                        asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)base + off));
                };
                pull(a.T + nc * nl, nl * 8);
                pull(a.Tint + nc * L, L * 8);
                pull(a.T_g + nc, 8);
                if (a.albedo) pull(a.albedo + nc, 8);
                if (FUSED) {
                    for (int m = 0; m < MR; ++m) {
                        pull(a.tidx + ((int64_t)m * a.n_col + nc) * L, L);
                        if (a.q_col_stride) pull(a.q + m * a.q_mol_stride + nc * a.q_col_stride, L * 8);
                    }
                }
            }
        }

        // ---- the column's level tables ---------------------------------------------------------------
        for (int idx = tid; idx < MR * lstride; idx += nthr) {
            const int m = idx / lstride, pos = idx - m * lstride;
            const int k = pos - kLevShift, kc = a.dbg_row0 ? 0 : min(max(k, 0), L - 1), kcu = a.dbg_row0 ? 0 : min(max(k - 1, 0), L - 1);
            if (FUSED) {
                const uint8_t *tcol = a.tidx + ((int64_t)m * a.n_col + col) * L;
                lev_offdn[idx] = (unsigned)(((((int64_t)m * L + kc) * a.nT_pad + tcol[kc]) * a.pitch * sizeof(Real)) >> 4);
                lev_offup[idx] = (unsigned)(((((int64_t)m * L + kcu) * a.nT_pad + tcol[kcu]) * a.pitch * sizeof(Real)) >> 4);
                lev_qg[idx] = (k >= 0 && k < L) ? (Real)(a.q[m * a.q_mol_stride + col * a.q_col_stride + k] / a.g) : Real(0);
            } else {
                lev_offdn[idx] = (unsigned)((int64_t)kc * a.N);
                lev_offup[idx] = (unsigned)((int64_t)kcu * a.N);
                lev_qg[idx] = Real(0);
            }
        }

        const long long t_lev = a.dbg ? clock64() : 0;
        // ---- emission coefficients of the long-wave bands this CTA touches (real_gas.py:634-638) -----
        const int sA = a.part_seg0[cpart * W], sB = a.part_seg0[(cpart + 1) * W];
        const double Tg = a.T_g[col];
        const double albedo = a.albedo ? a.albedo[col] : a.albedo_scalar;
        const double solar = a.star_geom * (1.0 - albedo) / 4.0;  // (R*/d)^2 (1-albedo)/4, :640-642
        double *bsrc = (double *)(smem_raw + lay.off_bsrc);       // kVarRedTmem: boundary sources of every band, once per column
        if constexpr (RT)
            for (int j = tid; j < a.B; j += nthr) {
                double K, N;
                if (meta_regs) {                                  // j = tid here: this lane's band
                    K = j < 32 ? bK0 : bK1, N = j < 32 ? bN0 : bN1;
                } else {
                    planck_band_constants(a.pc, __ldg(a.band_centre + j), K, N);
                }
                bsrc[j] = planck_fast(K, N, 1.0 / Tg);                                    // :638
                bsrc[a.B + j] = planck_fast(K, N, 1.0 / a.T_star) * solar;                // :640-642
            }
        if (sA < sB) {
            // Lane = one layer of a 31-layer chunk (lane 31 only supplies B at the next interface); the
            // warps that share a chunk take every other long-wave band.  Temperatures and the pressure
            // weights of the layer are fetched once, and the bands are walked two at a time so that the
            // Planck evaluations (exp + divide, latency bound) overlap.
            const int jA = a.seg_band[sA], jB = a.seg_band[sB - 1];
            const int nch = (nl + 30) / 31, chs = min(W, nch), groups = W / chs, grp = warp / chs;
            for (int ch = warp % chs; grp < groups && ch < nch; ch += chs) {
                const int k = ch * 31 + lane;
                const double Ti = k < L ? (k == L - 1 ? Tg : a.Tint[col * L + k]) : 1.0;      // :636
                const double Tc = k < nl ? a.T[col * nl + k] : 1.0;
                const double rTi = rcp_fast(Ti), rTc = rcp_fast(Tc);
                const bool writer = lane < 31 && k < nl;
                double gw[10];
#pragma unroll
                for (int r = 0; r < 10; ++r) gw[r] = writer ? __ldg(a.geo + r * nl + k) : 0.0;
                auto one_band = [&](int j, double bi, double bc) {
                    const double bi1 = __shfl_down_sync(0xffffffffu, bi, 1);
                    if (writer) {
                        Real *cf = coef_cta + (size_t)j * 4 * Lp + k;
                        cf[0] = (Real)(gw[0] * bi + gw[1] * bc + gw[2] * bi1);        // dn_last
                        cf[Lp] = (Real)(gw[3] * bi + gw[4] * bc);                     // dn_gen
                        cf[2 * Lp] = (Real)(gw[5] * bi + gw[6] * bc + gw[7] * bi1);   // up_first
                        cf[3 * Lp] = (Real)(gw[8] * bc + gw[9] * bi1);                // up_gen
                    }
                };
                int pend = -1;                                      // a long-wave band waiting for a partner
                for (int j = jA + grp; j <= jB; j += groups) {
                    if (band_is_sw(j)) continue;
                    if (pend < 0) {
                        pend = j;
                        continue;
                    }
                    double K0, N0, K1, N1;
                    band_KN(pend, K0, N0);
                    band_KN(j, K1, N1);
                    const double bi0 = planck_fast(K0, N0, rTi), bi1 = planck_fast(K1, N1, rTi);
                    const double bc0 = planck_fast(K0, N0, rTc), bc1 = planck_fast(K1, N1, rTc);
                    one_band(pend, bi0, bc0);
                    one_band(j, bi1, bc1);
                    pend = -1;
                }
                if (pend >= 0) {
                    double K0, N0;
                    band_KN(pend, K0, N0);
                    one_band(pend, planck_fast(K0, N0, rTi), planck_fast(K0, N0, rTc));
                }
            }
        }
        const long long t_coef = a.dbg ? clock64() : 0;
        __syncthreads();

        // ---- this warp's run of segments -------------------------------------------------------------
        const double *tau_col = FUSED ? nullptr : a.tau + (int64_t)(a.tau_id ? a.tau_id[col] : 0) * L * a.N;
        const int part = cpart * W + warp;
        int s = a.part_seg0[part];
        const int s1 = a.part_seg0[part + 1];
        int slot = a.part_slot0[part];
        bool up = false;
        const long long t_sweep0 = a.dbg ? clock64() : 0;
        // Parameters of a segment: one chain (wavenumber bin of its band) per lane.  They, and the
        // first table rows of the pass, are fetched one pass ahead so that no pass starts on a miss.
        auto seg_params = [&](int sidx, Real &w, const void *&src) {
            w = __ldg(a.seg_w + (int64_t)sidx * 32 + lane);
            if (FUSED && a.seg_space) {
                src = (const void *)(a.KL + ((int64_t)sidx * 32 + lane));
            } else {
                const int64_t n = __ldg(a.seg_bin + (int64_t)sidx * 32 + lane);
                src = FUSED ? (const void *)(a.KL + n) : (const void *)(tau_col + n);
            }
        };
        PassFeed<Real, MM> fd;
        PairFeed<MM> fd2;                   // kVarPair: the feed of a short-wave pair pass
        Real wN = Real(0), wN1 = Real(0);   // wN1 (kVarPair): weight of the segment after the one wN belongs to
        const void *srcN = nullptr;
        auto seg_params_pair = [&](int sidx) {
            seg_params(sidx, wN, srcN);
            wN1 = sidx + 1 < s1 ? __ldg(a.seg_w + (int64_t)(sidx + 1) * 32 + lane) : Real(0);
        };
        if (s < s1) {
            if constexpr (PAIR) {
                seg_params_pair(s);
                if (band_is_sw(__ldg(a.seg_band + s)))
                    prefetch_pair_pass<MM>(up, Lp, srcN, lv, fd2);
                else
                    prefetch_pass<Real, MM>(up, Lp, srcN, lv, fd);
            } else {
                seg_params(s, wN, srcN);
                prefetch_pass<Real, MM>(up, Lp, srcN, lv, fd);
            }
        }
        // Parameters of a (warp, band) run -- band, first chain, band length, sum of the nu-weights -- come from global
        // memory through two dependent loads; those of the NEXT run are requested when a run starts and arrive
        // behind its passes.
        int rj = 0, rc0 = 0, rlen = 0;
        double rsumw = 0.0;
        auto load_run = [&](int sidx, int &j_, int &c0_, int &len_, double &sumw_) {
            j_ = __ldg(a.seg_band + sidx);
            c0_ = __ldg(a.seg_c0 + sidx);
            len_ = __ldg(a.band_len + j_);
            sumw_ = __ldg(a.band_sumw + j_);
        };
        if (s < s1) load_run(s, rj, rc0, rlen, rsumw);
        while (s < s1) {
            const int j = rj;
            // the run ends with the band (its remaining chains, 32 per segment) or with this warp's share
            const int e = min(s1, s + ((rlen - rc0 + 31) >> 5));
            const bool first_chain = rc0 == 0;
            const double sumw = rsumw;
            if (e < s1) load_run(e, rj, rc0, rlen, rsumw);
            const bool lw = !band_is_sw(j);
            Real bg, bs;
            if constexpr (RT) {
                bg = (Real)bsrc[j];
                bs = (Real)bsrc[a.B + j];
            } else {
                const double c = a.band_centre[j];
                bg = (Real)(M_PI * planck_wavenumber(a.pc, c, Tg));                   // :638
                bs = (Real)(M_PI * planck_wavenumber(a.pc, c, a.T_star) * solar);     // :640-642
            }
            const Real *coef = coef_cta + (size_t)j * 4 * Lp;
            if constexpr (CS) {          // stage the band's four coefficient rows in this warp's shared memory
                if (lw) {
                    for (int i = lane; i < 4 * Lp; i += 32) cw[i] = coef[i];
                    __syncwarp();
                }
                coef = cw;
            }
            // boundary terms: Trans(i,i) = sum of the nu-weights
            const double bnd_up = first_chain ? (double)bg * sumw : 0.0;
            const double bnd_dn = first_chain ? (double)bs * sumw : 0.0;
            if (first_chain && lane == 0) {
                bacc_up[L - 1] += bnd_up;
                bacc_dn[0] += bnd_dn;
            }
            if constexpr (PAIR) {
                // The first pass of the NEXT run is fed at this run's end, by its kind (rj is the next run's band by now).
                auto feed_next_run = [&](bool up_next) {
                    if (e >= s1) return;
                    if (band_is_sw(rj))
                        prefetch_pair_pass<MM>(up_next, Lp, srcN, lv, fd2);
                    else
                        prefetch_pass<Real, MM>(up_next, Lp, srcN, lv, fd);
                };
                if (lw) {
                    Real wR = Real(0);
                    for (; s < e; ++s) {
                        const Real wF = wN;
                        const void *src_lane = srcN;
                        if (s + 1 < s1) seg_params_pair(s + 1);
                        const void *nxt = s + 1 < e ? srcN : nullptr;
                        if (up)
                            sweep_pass<Real, MM, true, true, true, V>(lane, nl, Lp, wF, wR, bg, src_lane, nxt, fd, lv, tst, bacc_up, coef + 2 * Lp, coef + 3 * Lp, xaux);
                        else
                            sweep_pass<Real, MM, true, false, true, V>(lane, nl, Lp, wF, wR, bs, src_lane, nxt, fd, lv, tst, bacc_dn + 1, coef, coef + Lp, xaux);
                        wR = wF;
                        up = !up;
                    }
                    feed_next_run(!up);
                    if (up)
                        sweep_pass<Real, MM, true, true, false, V>(lane, nl, Lp, Real(0), wR, bg, nullptr, nullptr, fd, lv, tst, bacc_up, coef + 2 * Lp, coef + 3 * Lp, xaux);
                    else
                        sweep_pass<Real, MM, true, false, false, V>(lane, nl, Lp, Real(0), wR, bs, nullptr, nullptr, fd, lv, tst, bacc_dn + 1, coef, coef + Lp, xaux);
                    up = !up;
                } else {
                    f2 wR = f2_set(0.f);
                    while (s < e) {
                        const f2 wF = make_float2(wN, s + 1 < e ? wN1 : 0.f);
                        const void *src_lane = srcN;
                        const int sn = min(s + 2, e);
                        if (sn < s1) seg_params_pair(sn);
                        const void *nxt = sn < e ? srcN : nullptr;
                        if (up)
                            sw_fresh_pair_pass<MM, true, V>(Lp, wF, wR, bg, src_lane, nxt, fd2, lv, tst.taddr, ra);
                        else
                            sw_fresh_pair_pass<MM, false, V>(Lp, wF, wR, bs, src_lane, nxt, fd2, lv, tst.taddr, ra);
                        wR = wF;
                        up = !up;
                        s = sn;
                    }
                    feed_next_run(!up);
                    if (up)
                        sw_replay_pair_pass<true, (V & kVarSmallSt) != 0>(Lp, wR, bg, tst.taddr, ra);
                    else
                        sw_replay_pair_pass<false, (V & kVarSmallSt) != 0>(Lp, wR, bs, tst.taddr, ra);
                    up = !up;
                }
            } else {
            Real wR = Real(0);
            for (; s < e; ++s) {
                const Real wF = wN;
                const void *src_lane = srcN;
                if (s + 1 < s1) seg_params(s + 1, wN, srcN);
                dispatch_pass<Real, MM, true, V>(lw, up, lane, nl, Lp, wF, wR, bs, bg, src_lane, s + 1 < e ? srcN : nullptr, fd, lv,
                                                 tst, bacc_up, bacc_dn, coef, xaux, ra);
                wR = wF;
                up = !up;
            }
            dispatch_pass<Real, MM, false, V>(lw, up, lane, nl, Lp, Real(0), wR, bs, bg, nullptr, s < s1 ? srcN : nullptr, fd, lv,
                                              tst, bacc_up, bacc_dn, coef, xaux, ra);
            up = !up;
            }
            // ---- partial band sums of this (warp, band) run -> scratch slot ---------------------------
            __syncwarp();
            double *o = a.scratch + ((region * a.n_slots + slot) * 2) * (int64_t)L;
            if constexpr (RT) {
                // class accumulators of the short-wave passes: indexed by layer; an upward layer k is interface k
                // (padding layers >= nl hold junk), a downward layer k interface k + 1.  Slot of layer k, class c:
                // FP64 (k / 8) * 32 + (k & 3) + 4 c, .x for k % 8 < 4, .y otherwise; FP32 (k / 8) * 32 + (k % 8) / 2 + 4 c,
                // .x for even k; classes added in order.
                // (kVarPair parks two words per value, so its classes are laid out like FP64's)
                constexpr bool WIDE = sizeof(Real) == 8 || PAIR;
                auto pslot = [](int k) { return (k >> 3) * 32 + (WIDE ? (k & 3) : ((k & 7) >> 1)); };
                auto phigh = [](int k) { return WIDE ? (k & 4) != 0 : (k & 1) != 0; };
                for (int i = lane; i < L; i += 32) {
                    double su = bacc_up[i], sd = bacc_dn[i];
                    if (!lw) {
                        if (i < nl) {
                            const double2 *pu = pacc + pslot(i);
#pragma unroll
                            for (int c = 0; c < 8; ++c) su += phigh(i) ? pu[4 * c].y : pu[4 * c].x;
                        }
                        if (i >= 1) {
                            const int k = i - 1;
                            const double2 *pd = pacc + Lp * 4 + pslot(k);
#pragma unroll
                            for (int c = 0; c < 8; ++c) sd += phigh(k) ? pd[4 * c].y : pd[4 * c].x;
                        }
                    }
                    o[i] = su;
                    o[L + i] = sd;
                }
                __syncwarp();
                if (!lw)
                    for (int i = lane; i < n_pacc; i += 32) pacc[i] = make_double2(0.0, 0.0);
            } else {
                for (int i = lane; i < L; i += 32) {
                    o[i] = bacc_up[i];
                    o[L + i] = bacc_dn[i];
                }
            }
            ++slot;
            __syncwarp();
            for (int i = lane; i < n_bacc / 2; i += 32) reinterpret_cast<double2 *>(bacc_up)[i] = make_double2(0.0, 0.0);
            __syncwarp();
        }
        const bool dbg_here = a.dbg && n_done == (int)(a.n_tasks / gridDim.x / 2) && blockIdx.x == 0 && lane == 0;
        if (dbg_here) {
            a.dbg[warp] = clock64() - t_sweep0;
            a.dbg[32 + warp] = t_sweep0 - t_task0;
            if (warp < 4) a.dbg[60 + warp] = warp == 0 ? t_lev - t_task0 : warp == 1 ? t_coef - t_lev : warp == 2 ? t_coef - t_task0 : t_lev - t_task0;
        }
        if (C != 1) continue;      // several CTAs per column: flux_band_sums_kernel + flux_net_kernel finish

        // ---- column epilogue: band sums, net flux (:712-713), heating rate (base.py:166-168) ----------
        __syncthreads();
        const double *sc = a.scratch + region * a.n_slots * 2 * (int64_t)L;
        for (int idx = tid; idx < 2 * L; idx += nthr) {
            const int half = idx >= L, i = idx - half * L;
            double *out = half ? a.down : a.up;
            double acc = 0.0;
            if (!out) {     // only net / heating wanted: independent loads, slot by slot
                const double *v = sc + half * (int64_t)L + i;
                // sixteen slots in flight at a time (the scratch is L2 resident: a batch costs one L2 round trip);
                // the sum itself stays in slot order
                for (int q = 0; q < a.n_slots; q += 16) {
                    double x[16], d[16];
#pragma unroll
                    for (int r = 0; r < 16; ++r) {
                        const int qq = min(q + r, a.n_slots - 1);
                        x[r] = v[qq * 2 * (int64_t)L];
                        d[r] = __ldg(a.slot_delta + qq);
                    }
#pragma unroll
                    for (int r = 0; r < 16; ++r)
                        if (q + r < a.n_slots) acc = fma(x[r], d[r], acc);
                }
            } else
            for (int j = a.band_lo; j < a.band_hi; ++j) {
                double sum = 0.0;
                const int q0 = a.band_slot0[j], q1 = a.band_slot0[j + 1];
                for (int q = q0; q < q1; ++q) sum += sc[(q * 2 + half) * (int64_t)L + i];
                if (out) out[(col * L + i) * a.B + j] = sum;
                acc += sum * a.band_delta[j];
            }
            sacc[idx] = acc;
        }
        __syncthreads();
        for (int i = tid; i < L; i += nthr) {
            const double net = sacc[i] - sacc[L + i];
            if (a.net) a.net[col * L + i] = net;
            if (a.heating && i < nl) {
                const double n1 = sacc[i + 1] - sacc[L + i + 1];
                a.heating[col * nl + i] = a.g_over_cp * (n1 - net) / (a.Pint[i + 1] - a.Pint[i]);
            }
        }
        __syncthreads();
        for (int idx = tid; idx < 2 * L; idx += nthr) sacc[idx] = 0.0;
        if (dbg_here) a.dbg[48 + warp] = clock64() - t_task0;
    }
    if constexpr (TM) {      // every warp left the task loop behind the same barrier: nobody uses TMEM any more
        tmem::fence_before_sync();
        __syncthreads();
        if (warp == 0) tmem::release(s_tmem_base, (uint32_t)a.tmem_cols);
    }
}

// Split mode (few columns), two small kernels.  flux_band_sums_kernel: one CTA per (column, band),
// one thread per (direction, interface); adds the band's per-part partial sweeps in slot order
// (loads issued eight at a time, the adds stay sequential) and leaves the band sum in the band's
// first slot.  flux_net_kernel: one CTA per column, band sums -> net flux (:712-713) -> heating
// rate (base.py:166-168), as in the fused epilogue.
static __global__ void flux_band_sums_kernel(int L, int B, int band_lo, int n_slots,
                                      const int32_t *__restrict__ band_slot0, double *scratch,
                                      double *up, double *down) {
    const int64_t col = blockIdx.x;
    const int j = band_lo + blockIdx.y;
    const int q0 = band_slot0[j], q1 = band_slot0[j + 1];
    for (int idx = threadIdx.x; idx < 2 * L; idx += blockDim.x) {
        const int half = idx >= L, i = idx - half * L;
        double *v = scratch + col * n_slots * 2 * (int64_t)L + idx;
        double sum = 0.0;
        int q = q0;
        for (; q + 8 <= q1; q += 8) {
            double x[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) x[r] = v[(q + r) * 2 * (int64_t)L];
#pragma unroll
            for (int r = 0; r < 8; ++r) sum += x[r];
        }
        for (; q < q1; ++q) sum += v[q * 2 * (int64_t)L];
        if (q0 < q1) v[q0 * 2 * (int64_t)L] = sum;
        double *out = half ? down : up;
        if (out) out[(col * L + i) * B + j] = sum;
    }
}

static __global__ void flux_net_kernel(int L, int band_lo, int band_hi, int n_slots,
                                const int32_t *__restrict__ band_slot0,
                                const double *__restrict__ band_delta, const double *__restrict__ scratch,
                                const double *__restrict__ Pint, double g_over_cp, double *net, double *heating) {
    extern __shared__ double snet[];
    const int64_t col = blockIdx.x;
    const double *sc = scratch + col * n_slots * 2 * (int64_t)L;
    for (int i = threadIdx.x; i < L; i += blockDim.x) {
        double nu = 0.0, nd = 0.0;
        for (int j = band_lo; j < band_hi; ++j) {
            const int q0 = band_slot0[j];
            if (q0 == band_slot0[j + 1]) continue;          // band outside this plan's segment range
            nu += sc[(q0 * 2) * (int64_t)L + i] * band_delta[j];
            nd += sc[(q0 * 2 + 1) * (int64_t)L + i] * band_delta[j];
        }
        snet[i] = nu - nd;
        if (net) net[col * L + i] = nu - nd;
    }
    __syncthreads();
    if (heating)
        for (int k = threadIdx.x; k < L - 1; k += blockDim.x)
            heating[col * (L - 1) + k] = g_over_cp * (snet[k + 1] - snet[k]) / (Pint[k + 1] - Pint[k]);
}

// T_interface = A T (the not-a-knot spline of real_gas.py:634 as a matrix, At = A transposed:
// [L-1][L]), T_interface[L-1] = T_g (:636).  One warp per (column, 32 interfaces); the sum runs
// over the layers in order, loads batched by the unrolling.
static __global__ void interface_temperature_kernel(int L, int64_t n_col, const double *__restrict__ At,
                                             const double *__restrict__ T,
                                             const double *__restrict__ T_g,
                                             double *__restrict__ Tint) {
    const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunks = (L + 31) >> 5;
    const int64_t item = (int64_t)blockIdx.x * warps + warp;
    const int64_t col = item / chunks;
    if (col >= n_col) return;
    const int i = (int)(item - col * chunks) * 32 + lane;
    if (i >= L) return;
    const int nl = L - 1;
    const double *Tc = T + col * nl;
    double acc = 0.0;
#pragma unroll 8
    for (int k = 0; k < nl; ++k) acc += At[k * L + i] * Tc[k];
    Tint[col * L + i] = (i == L - 1) ? T_g[col] : acc;
}

}  // namespace cmb
