// Stages 2+3 of the hot path: layer transmissivity, Planck source and the two flux sweeps
// (RealGas.get_flux / flux_integrals, real_gas.py:564-669), net flux (:712-713) and heating
// rate (base.py:166-168), optionally with stage 1 (optical depth, :86-127) fused in front.
//
// Per-bin form (oracle/realgas_oracle.py:get_flux_per_bin): every quantity the
// reference builds from the band-averaged transmission matrix is linear in the per-bin
// products E(i,k) = prod_{l=i}^{k-1} t_l with t_l = exp(-dtau_l), so each (band, bin) "chain"
// is one exp per layer plus two first-order recurrences over the layers, and the band value is
// the trapezoid-weighted sum of the chains.
//
// Mapping: one warp owns one (column, part) task; each lane holds one chain at a time
// (consecutive lanes = consecutive wavenumber bins, so table / tau rows are read coalesced).
// The downward sweep (TOA -> surface) computes t_k and parks it in shared memory
// ts[k][lane]; the upward sweep (surface -> TOA) replays it, so each layer costs one
// exp, not two.  Per interface the chains are summed over the warp with shuffles and added to
// the band accumulators; bands are folded into net flux and heating rate in the epilogue.
#pragma once
#include "internal.cuh"

namespace cmb {

template <typename Real>
struct FluxArgs {
    int L, B, M, nT_pad, parts, n_slots;
    int64_t N, pitch, n_col;
    // plan-static
    const double *geo, *Pint, *hdP;
    const int32_t *band_start, *band_len;
    const double *band_centre, *band_delta, *band_sumw;
    const uint8_t *band_sw;
    const Real *wchain;
    const int64_t *chain_off;
    const int32_t *seg_band, *seg_c0, *part_seg0, *part_slot0;
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
    double *up, *down, *net, *heating, *scratch;
};

constexpr int kU = 8;   // layers per block of the sweeps (instruction-level parallelism)

// Per-level, warp-uniform lookup data of one molecule kept in shared memory: q/g and the offset
// of the table row KL[m][k][tidx] (fused path) or of the tau row k (resident-tau path).  Level k
// sits at index k + kLevShift so that the kU levels k0+1 .. k0+kU of a block start on a 16-byte
// boundary and are read with vector loads; levels are padded beyond L-1 with q/g = 0 and a valid
// offset so that the sweeps need no bounds checks.
constexpr int kLevShift = 3;
template <typename Real>
struct LevelView {
    const unsigned *off;   // [MR][stride]
    const Real *qg;        // [MR][stride]
    int stride;
    const Real *hdp;       // [Lp]   half layer thickness, 0 in the padding
    const Real *coef;      // [4][Lp] emission coefficients of the current long-wave band
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

__host__ __device__ inline int padded_layers(int L) { return ((L - 1) + kU - 1) / kU * kU; }

// Per-warp shared-memory layout (every sub-array 16-byte aligned).
template <typename Real, int MM>
struct WarpSmem {
    size_t off_nacc, off_off, off_qg, off_hdp, off_coef, off_ts, total;
    int lev_stride;
    __host__ __device__ WarpSmem(int L) {
        auto up16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
        const int Lp = padded_layers(L), MR = MM > 0 ? MM : 1;
        size_t b = up16((size_t)2 * (Lp + kU) * sizeof(double));          // bacc_up, bacc_dn
        off_nacc = b;
        b = up16(b + (size_t)2 * L * sizeof(double));                     // nacc_up, nacc_dn
        lev_stride = Lp + 2 * kU;                                         // level k at k + kLevShift
        off_off = b;
        b = up16(b + (size_t)MR * lev_stride * sizeof(unsigned));
        off_qg = b;
        b = up16(b + (size_t)MR * lev_stride * sizeof(Real));
        off_hdp = b;
        b = up16(b + (size_t)Lp * sizeof(Real));
        off_coef = b;
        b = up16(b + (size_t)4 * Lp * sizeof(Real));
        off_ts = b;
        size_t ts_bytes = (size_t)Lp * 32 * sizeof(Real);
        size_t planck_bytes = (size_t)2 * L * sizeof(double);             // band-prologue scratch
        b = up16(b + (ts_bytes > planck_bytes ? ts_bytes : planck_bytes));
        total = b;
    }
};
template <typename Real, int MM>
__host__ __device__ inline size_t flux_warp_smem(int L) {
    return WarpSmem<Real, MM>(L).total;
}

// exp(-x) for the layer transmissivity.  FP64: n = rint(-x log2 e), r = -x - n ln2 (two-term
// Cody-Waite), degree-11 near-minimax polynomial on |r| <= ln2/2 (max rel. error 1.6e-17 before
// rounding), scaled by 2^n through the exponent field; no special cases beyond the argument clamp.
// Coefficients sit in constant memory so that each DFMA takes its constant as an operand.
__constant__ double kExpC[16] = {
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

// One direction pair of sweeps for the chain held by this lane.  LW selects the long-wave
// variant (emission recurrences); MM as in flux_kernel.
//   down, TOA -> surface:  t_k = exp(-dtau_k);  E <- E t_k;  value at interface k+1
//   up,   surface -> TOA:  replay t_k from shared memory;    value at interface k
// The nu-weight w (and the boundary source) is folded into the initial state so that the values
// handed to the warp reduction are already weighted.
//
// Both sweeps are software-pipelined by hand over blocks of kU layers, because a lane's work is
// one long dependency chain and only ~1.5 warps per scheduler fit next to the t-store: while the
// kU independent exp evaluations of block b are in flight (stage A), the serial recurrence of
// block b-1 (stage B) and its warp reduction (stage C) are issued in between, statement by
// statement -- ptxas keeps this order, so the FP64 pipe always has independent work.
template <typename Real, int MM, bool LW>
__device__ __forceinline__ void sweep_chain(const int lane, const int nl, const int Lp, const Real w, const Real bs,
                                            const Real bg, const void *src_lane, const LevelView<Real> lv, Real *ts,
                                            double *bacc_up, double *bacc_dn) {
    constexpr bool FUSED = MM > 0;
    constexpr int MR = MM > 0 ? MM : 1;
    constexpr bool F64 = sizeof(Real) == 8;
    using Src = typename std::conditional<FUSED, Real, double>::type;
    const char *src_bytes = (const char *)src_lane;
    constexpr unsigned kOffUnit = FUSED ? 16u : 8u;   // table rows are 64-byte aligned, tau rows only 8
    const unsigned full = 0xffffffffu;
    const bool h4 = lane & 16, h3 = lane & 8, h2 = lane & 4;
    const int ridx = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
    const bool rlane = (lane & 3) == 0;
    const Real *hdp = lv.hdp, *coef = lv.coef;

    // raw[u][m]: table value (or tau) of interface k0+1+u, loaded one block ahead so the L2/HBM
    // latency hides behind the previous block's work.
    Src raw[kU][MR];
    auto load_block = [&](int kfirst) {      // kfirst = k0 + 1, k0 a multiple of kU
#pragma unroll
        for (int m = 0; m < MR; ++m) {
            unsigned off[kU];
            lds_block(lv.off + m * lv.stride + kfirst + kLevShift, off);
#pragma unroll
            for (int u = 0; u < kU; ++u) raw[u][m] = *(const Src *)(src_bytes + (size_t)off[u] * kOffUnit);
        }
    };
    Src Iprev;
    if (FUSED) {
        Real acc = Real(0);
#pragma unroll
        for (int m = 0; m < MR; ++m)
            acc += lv.qg[m * lv.stride + kLevShift] * (Real)(*(const Src *)(src_bytes + (size_t)lv.off[m * lv.stride + kLevShift] * kOffUnit));
        Iprev = (Src)acc;
    } else {
        Iprev = *(const Src *)(src_bytes + (size_t)lv.off[kLevShift] * kOffUnit);
    }
    load_block(1);

    // ---- shared pieces of the pipeline stages -----------------------------------------------------
    Real E, H, tp[kU], v[kU], a4[4], a2[2], c1;      // recurrence state, block b-1's t and values
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

    // ---- downward sweep ---------------------------------------------------------------------------------
    auto Bdn = [&](int kb, int u) {           // recurrence step of layer kb+u (block b-1)
        E *= tp[u];
        if (LW) {
            const Real om = fma(-w, tp[u], w);              // w (1 - t)
            const Real S = tp[u] * H;
            v[u] = E + S + coef[kb + u] * om;
            H = fma(coef[Lp + kb + u], om, S);
        } else {
            v[u] = E;
        }
    };
    auto down_iter = [&](auto doA_t, auto doB_t, const int k0) {
        constexpr bool DO_A = decltype(doA_t)::value, DO_B = decltype(doB_t)::value;
        const int kb = k0 - kU;
        Real x[kU], t[kU];
        // stage A set-up: layer integrand of the block's interfaces, next block's loads, dtau
        if (DO_A) {
            Src In[kU];
            if (FUSED) {
                Real qg[kU], acc[kU];
                lds_block(lv.qg + k0 + 1 + kLevShift, qg);
#pragma unroll
                for (int u = 0; u < kU; ++u) acc[u] = qg[u] * (Real)raw[u][0];
#pragma unroll
                for (int m = 1; m < MR; ++m) {
                    lds_block(lv.qg + m * lv.stride + k0 + 1 + kLevShift, qg);
#pragma unroll
                    for (int u = 0; u < kU; ++u) acc[u] = fma(qg[u], (Real)raw[u][m], acc[u]);
                }
#pragma unroll
                for (int u = 0; u < kU; ++u) In[u] = (Src)acc[u];
            } else {
#pragma unroll
                for (int u = 0; u < kU; ++u) In[u] = raw[u][0];
            }
            if (k0 + kU < Lp) load_block(k0 + 1 + kU);
            Real hd[kU];
            if (FUSED) lds_block(hdp + k0, hd);
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const Src lo = u ? In[u - 1] : Iprev;
                x[u] = FUSED ? (Real)(hd[u] * (Real)(lo + In[u])) : (Real)(In[u] - lo);
            }
            Iprev = In[kU - 1];
        }
        if constexpr (F64) {
            double tn[kU], r[kU], p[kU];
            int n[kU];
            if (DO_A) {
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    x[u] = clamp_thickness(x[u]);
                    tn[u] = fma(-x[u], kExpC[11], kExpC[12]);
                }
            }
            if (DO_B) { Bdn(kb, 0); Bdn(kb, 1); }
            if (DO_A) {
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    n[u] = __double2loint(tn[u]);
                    tn[u] -= kExpC[12];
                }
            }
            if (DO_B) { Bdn(kb, 2); Bdn(kb, 3); }
            if (DO_A) {
#pragma unroll
                for (int u = 0; u < kU; ++u) r[u] = fma(tn[u], kExpC[13], -x[u]);
            }
            if (DO_B) { Bdn(kb, 4); Bdn(kb, 5); }
            if (DO_A) {
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    r[u] = fma(tn[u], kExpC[14], r[u]);
                    p[u] = kExpC[0];
                }
            }
            if (DO_B) { Bdn(kb, 6); Bdn(kb, 7); }
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
                for (int u = 0; u < kU; ++u) Bdn(kb, u);
                C1(); C2(); C3(); C4(); C5();
            }
        }
        if (DO_B) {
            if (rlane && kb + ridx < nl) bacc_dn[kb + ridx + 1] += (double)c1;
        }
        if (DO_A) {
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                ts[(k0 + u) * 32 + lane] = t[u];
                tp[u] = t[u];
            }
        }
    };
    E = w * bs;
    H = Real(0);
    down_iter(std::true_type{}, std::false_type{}, 0);
    for (int k0 = kU; k0 < Lp; k0 += kU) down_iter(std::true_type{}, std::true_type{}, k0);
    down_iter(std::false_type{}, std::true_type{}, Lp);

    // ---- upward sweep: recurrence of block b interleaved with the reduction of block b+1 ------------
    Real S = Real(0), tn8[kU], vp[kU];
    auto Bup = [&](int kb, int u) {
        E *= tp[u];
        if (LW) {
            const Real om = fma(-w, tp[u], w);
            const Real y = tp[u] * S;
            vp[u] = E + y + coef[2 * Lp + kb + u] * om;
            S = fma(coef[3 * Lp + kb + u], om, y);
        } else {
            vp[u] = E;
        }
    };
    auto up_iter = [&](auto doB_t, auto doC_t, const int k0) {
        constexpr bool DO_B = decltype(doB_t)::value, DO_C = decltype(doC_t)::value;
        // prefetch t of the block after this one (k0 - kU) while this block's chain runs
        if (DO_B && k0 >= kU) {
#pragma unroll
            for (int u = 0; u < kU; ++u) tn8[u] = ts[(k0 - kU + u) * 32 + lane];
        }
        if (DO_B) Bup(k0, 7);
        if (DO_C) C1();
        if (DO_B) { Bup(k0, 6); Bup(k0, 5); }
        if (DO_C) C2();
        if (DO_B) { Bup(k0, 4); Bup(k0, 3); }
        if (DO_C) C3();
        if (DO_B) Bup(k0, 2);
        if (DO_C) C4();
        if (DO_B) { Bup(k0, 1); Bup(k0, 0); }
        if (DO_C) {
            C5();
            const int kc = k0 + kU;
            if (rlane && kc + ridx < nl) bacc_up[kc + ridx] += (double)c1;
        }
        if (DO_B) {
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                v[u] = vp[u];
                tp[u] = tn8[u];
            }
        }
    };
    E = w * bg;
#pragma unroll
    for (int u = 0; u < kU; ++u) tp[u] = ts[(Lp - kU + u) * 32 + lane];
    up_iter(std::true_type{}, std::false_type{}, Lp - kU);
    for (int k0 = Lp - 2 * kU; k0 >= 0; k0 -= kU) up_iter(std::true_type{}, std::true_type{}, k0);
    up_iter(std::false_type{}, std::true_type{}, -kU);
}

// MM = number of molecules of the table-fused path (compile-time so the prefetched table values
// live in registers); MM = 0 selects the resident-tau path.
template <typename Real, int MM>
__global__ void __launch_bounds__(256, 1) flux_kernel(const FluxArgs<Real> a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr bool FUSED = MM > 0;
    constexpr int MR = MM > 0 ? MM : 1;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int L = a.L, nl = a.L - 1, Lp = padded_layers(a.L), LS = Lp + kU;
    const int64_t task = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp;
    if (task >= a.n_col * a.parts) return;
    const int64_t col = task / a.parts;
    const int part = (int)(task % a.parts);

    const WarpSmem<Real, MM> lay(L);
    unsigned char *base = smem_raw + (size_t)warp * lay.total;
    double *bacc_up = (double *)base;
    double *bacc_dn = bacc_up + LS;
    double *nacc_up = (double *)(base + lay.off_nacc);
    double *nacc_dn = nacc_up + L;
    unsigned *lev_off = (unsigned *)(base + lay.off_off);
    Real *lev_qg = (Real *)(base + lay.off_qg);
    const int lstride = lay.lev_stride;
    Real *hdp = (Real *)(base + lay.off_hdp);
    Real *coef = (Real *)(base + lay.off_coef);
    Real *ts = (Real *)(base + lay.off_ts);

    const double Tg = a.T_g[col];
    const double albedo = a.albedo ? a.albedo[col] : a.albedo_scalar;
    const double solar = a.star_geom * (1.0 - albedo) / 4.0;  // (R*/d)^2 (1-albedo)/4, :640-642

    for (int i = lane; i < 2 * LS; i += 32) bacc_up[i] = 0.0;
    for (int i = lane; i < 2 * L; i += 32) nacc_up[i] = 0.0;
    for (int i = lane; i < 4 * Lp; i += 32) coef[i] = Real(0);
    for (int k = lane; k < Lp; k += 32) hdp[k] = k < nl ? (Real)a.hdP[k] : Real(0);
    for (int idx = lane; idx < MR * lstride; idx += 32) {
        const int m = idx / lstride, k = idx - m * lstride - kLevShift, kc = min(max(k, 0), L - 1);
        if (FUSED) {
            const int t = a.tidx[((int64_t)m * a.n_col + col) * L + kc];
            lev_off[idx] = (unsigned)(((((int64_t)m * L + kc) * a.nT_pad + t) * a.pitch * sizeof(Real)) >> 4);
            lev_qg[idx] = (k >= 0 && k < L) ? (Real)(a.q[m * a.q_mol_stride + col * a.q_col_stride + k] / a.g) : Real(0);
        } else {
            lev_off[idx] = (unsigned)((int64_t)kc * a.N);
            lev_qg[idx] = Real(0);
        }
    }
    const LevelView<Real> lv{lev_off, lev_qg, lstride, hdp, coef};
    const double *tau_col = FUSED ? nullptr : a.tau + (int64_t)(a.tau_id ? a.tau_id[col] : 0) * L * a.N;
    __syncwarp();

    const int s0 = a.part_seg0[part], s1 = a.part_seg0[part + 1];
    int slot = a.parts > 1 ? a.part_slot0[part] - 1 : 0;
    int cur = -1;
    bool lw = false;
    Real bg = 0, bs = 0;

    auto flush = [&](int j) {
        __syncwarp();
        if (a.parts == 1) {
            const double dj = a.band_delta[j];
            for (int i = lane; i < L; i += 32) {
                double u = bacc_up[i], d = bacc_dn[i];
                if (a.up) a.up[(col * L + i) * a.B + j] = u;
                if (a.down) a.down[(col * L + i) * a.B + j] = d;
                nacc_up[i] += u * dj;
                nacc_dn[i] += d * dj;
            }
        } else {
            ++slot;
            double *o = a.scratch + ((col * a.n_slots + slot) * 2) * (int64_t)L;
            for (int i = lane; i < L; i += 32) {
                o[i] = bacc_up[i];
                o[L + i] = bacc_dn[i];
            }
        }
        __syncwarp();
        for (int i = lane; i < 2 * LS; i += 32) bacc_up[i] = 0.0;
        __syncwarp();
    };

    for (int s = s0; s < s1; ++s) {
        const int j = a.seg_band[s];
        const int c0 = a.seg_c0[s];
        if (j != cur) {
            if (cur >= 0) flush(cur);
            cur = j;
            // ---- band prologue: Planck sources and the pressure-quadrature coefficients ----
            const double c = a.band_centre[j];
            lw = !a.band_sw[j];
            bg = (Real)(M_PI * planck_wavenumber(a.pc, c, Tg));                   // :638
            bs = (Real)(M_PI * planck_wavenumber(a.pc, c, a.T_star) * solar);     // :640-642
            if (lw) {
                double *sbi = (double *)ts, *sbc = sbi + L;
                for (int k = lane; k < L; k += 32) {
                    double Tk = (k == L - 1) ? Tg : a.Tint[col * L + k];          // :636
                    sbi[k] = M_PI * planck_wavenumber(a.pc, c, Tk);
                    if (k < nl) sbc[k] = M_PI * planck_wavenumber(a.pc, c, a.T[col * nl + k]);
                }
                __syncwarp();
                for (int k = lane; k < nl; k += 32) {
                    const double bi0 = sbi[k], bi1 = sbi[k + 1], bc = sbc[k];
                    const double *g0 = a.geo + k;
                    coef[k] = (Real)(g0[0] * bi0 + g0[nl] * bc + g0[2 * nl] * bi1);                 // dn_last
                    coef[Lp + k] = (Real)(g0[3 * nl] * bi0 + g0[4 * nl] * bc);                      // dn_gen
                    coef[2 * Lp + k] = (Real)(g0[5 * nl] * bi0 + g0[6 * nl] * bc + g0[7 * nl] * bi1);   // up_first
                    coef[3 * Lp + k] = (Real)(g0[8 * nl] * bc + g0[9 * nl] * bi1);                  // up_gen
                }
                __syncwarp();
            }
        }
        // ---- this segment's chain (one wavenumber bin of band j per lane) ----------------------
        const int ci = c0 + lane;
        const bool act = ci < a.band_len[j];
        const Real w = act ? a.wchain[a.chain_off[j] + ci] : Real(0);
        const int64_t n = a.band_start[j] + (act ? ci : 0);
        if (c0 == 0 && lane == 0) {   // boundary terms: Trans(i,i) = sum of the nu-weights
            bacc_up[L - 1] += (double)bg * a.band_sumw[j];
            bacc_dn[0] += (double)bs * a.band_sumw[j];
        }
        const void *src_lane = FUSED ? (const void *)(a.KL + n) : (const void *)(tau_col + n);
        if (lw)
            sweep_chain<Real, MM, true>(lane, nl, Lp, w, bs, bg, src_lane, lv, ts, bacc_up, bacc_dn);
        else
            sweep_chain<Real, MM, false>(lane, nl, Lp, w, bs, bg, src_lane, lv, ts, bacc_up, bacc_dn);
    }
    if (cur >= 0) flush(cur);

    // ---- column epilogue: net flux (:712-713) and heating rate (base.py:166-168) -----------------
    if (a.parts == 1) {
        if (a.net)
            for (int i = lane; i < L; i += 32) a.net[col * L + i] = nacc_up[i] - nacc_dn[i];
        if (a.heating)
            for (int k = lane; k < nl; k += 32) {
                double n1 = nacc_up[k + 1] - nacc_dn[k + 1], n0 = nacc_up[k] - nacc_dn[k];
                a.heating[col * nl + k] = a.g_over_cp * (n1 - n0) / (a.Pint[k + 1] - a.Pint[k]);
            }
    }
}

// Split mode (few columns): sums the per-(part, band) partial sweeps in slot order, then the
// same band -> net -> heating epilogue.  One warp per column, lanes over interfaces.
__global__ void flux_finalize_kernel(int L, int B, int band_lo, int band_hi, int n_slots,
                                     int64_t n_col, const int32_t *__restrict__ band_slot0,
                                     const double *__restrict__ band_delta,
                                     const double *__restrict__ scratch,
                                     const double *__restrict__ Pint, double g_over_cp,
                                     double *up, double *down, double *net, double *heating) {
    extern __shared__ double snet[];
    const int64_t col = blockIdx.x;
    for (int i = threadIdx.x; i < L; i += blockDim.x) {
        double nu = 0.0, nd = 0.0;
        for (int j = band_lo; j < band_hi; ++j) {
            double u = 0.0, d = 0.0;
            for (int s = band_slot0[j]; s < band_slot0[j + 1]; ++s) {
                const double *o = scratch + ((col * n_slots + s) * 2) * (int64_t)L;
                u += o[i];
                d += o[L + i];
            }
            if (up) up[(col * L + i) * B + j] = u;
            if (down) down[(col * L + i) * B + j] = d;
            nu += u * band_delta[j];
            nd += d * band_delta[j];
        }
        snet[i] = nu - nd;
        if (net) net[col * L + i] = nu - nd;
    }
    __syncthreads();
    if (heating)
        for (int k = threadIdx.x; k < L - 1; k += blockDim.x)
            heating[col * (L - 1) + k] = g_over_cp * (snet[k + 1] - snet[k]) / (Pint[k + 1] - Pint[k]);
}

// Interface temperatures (real_gas.py:634-636): T_int = A T with A the not-a-knot cubic-spline
// operator of the (shared) pressure grid, stored transposed At[k][i]; surface value := T_g.
__global__ void interface_temperature_kernel(int L, int64_t n_col, const double *__restrict__ At,
                                             const double *__restrict__ T,
                                             const double *__restrict__ T_g,
                                             double *__restrict__ Tint) {
    const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t col = (int64_t)blockIdx.x * warps + warp;
    if (col >= n_col) return;
    const int nl = L - 1;
    const double *Tc = T + col * nl;
    for (int i0 = 0; i0 < L; i0 += 32) {
        int i = i0 + lane;
        double acc = 0.0;
        if (i < L)
            for (int k = 0; k < nl; ++k) acc += At[k * L + i] * Tc[k];
        if (i < L) Tint[col * L + i] = (i == L - 1) ? T_g[col] : acc;
    }
}

}  // namespace cmb
