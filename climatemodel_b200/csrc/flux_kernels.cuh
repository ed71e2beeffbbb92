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

// Per-warp shared-memory layout (every sub-array 16-byte aligned).
template <typename Real, bool FUSED>
struct WarpSmem {
    size_t off_koff, off_qg, off_coef, off_ts, total;
    __host__ __device__ WarpSmem(int L, int M) {
        auto up16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
        size_t b = up16((size_t)4 * L * sizeof(double));                 // bacc_up/dn, nacc_up/dn
        off_koff = b;
        if (FUSED) b = up16(b + (size_t)M * L * sizeof(long long));
        off_qg = b;
        if (FUSED) b = up16(b + (size_t)M * L * sizeof(Real));
        off_coef = b;
        b = up16(b + (size_t)4 * (L - 1) * sizeof(Real));
        off_ts = b;
        size_t ts_bytes = (size_t)(L - 1) * kChains * 32 * sizeof(Real);
        size_t planck_bytes = (size_t)2 * L * sizeof(double);            // band-prologue scratch
        b = up16(b + (ts_bytes > planck_bytes ? ts_bytes : planck_bytes));
        total = b;
    }
};
template <typename Real, bool FUSED>
__host__ __device__ inline size_t flux_warp_smem(int L, int M) {
    return WarpSmem<Real, FUSED>(L, M).total;
}

constexpr int kU = 8;   // layers per block of the sweeps (instruction-level parallelism)

// exp(-x) for the layer transmissivity.  FP64: n = rint(-x log2 e), r = -x - n ln2 (two-term
// Cody-Waite), degree-11 near-minimax polynomial on |r| <= ln2/2 (max rel. error 1.6e-17 before
// rounding), scaled by 2^n through the exponent field.  Arguments beyond exp(-708) clamp there
// (3e-308, nothing at the 1e-10 parity level); no other special cases are needed on this path.
__device__ __forceinline__ double exp_neg(double x) {
    double y = fmin(fmax(-x, -708.0), 709.0);
    const double shifter = 6755399441055744.0;                 // 1.5 * 2^52
    double tn = fma(y, 1.4426950408889634074, shifter);
    const int n = __double2loint(tn);
    tn -= shifter;
    double r = fma(tn, -6.93147180369123816490e-01, y);
    r = fma(tn, -1.90821492927058770002e-10, r);
    double p = 2.51100492048186583e-08;
    p = fma(p, r, 2.76326547225277896e-07);
    p = fma(p, r, 2.75572408872298695e-06);
    p = fma(p, r, 2.48014854415613131e-05);
    p = fma(p, r, 1.98412698900764028e-04);
    p = fma(p, r, 1.38888889523528631e-03);
    p = fma(p, r, 8.33333333331958900e-03);
    p = fma(p, r, 4.16666666664879531e-02);
    p = fma(p, r, 1.66666666666666796e-01);
    p = fma(p, r, 5.00000000000001887e-01);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    return __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
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

// MM = number of molecules of the table-fused path (compile-time so the prefetched table values
// live in registers); MM = 0 selects the resident-tau path.
template <typename Real, int MM>
__global__ void __launch_bounds__(256, 1) flux_kernel(const FluxArgs<Real> a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr bool FUSED = MM > 0;
    constexpr int MR = MM > 0 ? MM : 1;
    using Src = typename std::conditional<FUSED, Real, double>::type;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int L = a.L, nl = a.L - 1, M = MM;
    const int64_t task = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp;
    if (task >= a.n_col * a.parts) return;
    const int64_t col = task / a.parts;
    const int part = (int)(task % a.parts);

    const WarpSmem<Real, FUSED> lay(L, M);
    unsigned char *base = smem_raw + (size_t)warp * lay.total;
    double *bacc_up = (double *)base;
    double *bacc_dn = bacc_up + L;
    double *nacc_up = bacc_dn + L;
    double *nacc_dn = nacc_up + L;
    long long *koff = (long long *)(base + lay.off_koff);
    Real *qg = (Real *)(base + lay.off_qg);
    Real *coef = (Real *)(base + lay.off_coef);
    Real *ts = (Real *)(base + lay.off_ts);

    const double Tg = a.T_g[col];
    const double albedo = a.albedo ? a.albedo[col] : a.albedo_scalar;
    const double solar = a.star_geom * (1.0 - albedo) / 4.0;  // (R*/d)^2 (1-albedo)/4, :640-642

    for (int i = lane; i < 4 * L; i += 32) bacc_up[i] = 0.0;
    const double *tau_col = nullptr;
    if (FUSED) {
        for (int idx = lane; idx < M * L; idx += 32) {
            int m = idx / L, k = idx - m * L;
            int t = a.tidx[((int64_t)m * a.n_col + col) * L + k];
            koff[idx] = (((long long)m * L + k) * a.nT_pad + t) * a.pitch;
            qg[idx] = (Real)(a.q[m * a.q_mol_stride + col * a.q_col_stride + k] / a.g);
        }
    } else {
        tau_col = a.tau + (int64_t)(a.tau_id ? a.tau_id[col] : 0) * L * a.N;
    }
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
                bacc_up[i] = 0.0;
                bacc_dn[i] = 0.0;
            }
        } else {
            ++slot;
            double *o = a.scratch + ((col * a.n_slots + slot) * 2) * (int64_t)L;
            for (int i = lane; i < L; i += 32) {
                o[i] = bacc_up[i];
                o[L + i] = bacc_dn[i];
                bacc_up[i] = 0.0;
                bacc_dn[i] = 0.0;
            }
        }
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
                    coef[k] = (Real)(g0[0] * bi0 + g0[nl] * bc + g0[2 * nl] * bi1);            // dn_last
                    coef[nl + k] = (Real)(g0[3 * nl] * bi0 + g0[4 * nl] * bc);                 // dn_gen
                    coef[2 * nl + k] = (Real)(g0[5 * nl] * bi0 + g0[6 * nl] * bc + g0[7 * nl] * bi1);  // up_first
                    coef[3 * nl + k] = (Real)(g0[8 * nl] * bc + g0[9 * nl] * bi1);             // up_gen
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
        const int ridx = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
        const bool rlane = (lane & 3) == 0;

        // ---- downward sweep, TOA -> surface, kU layers per block -------------------------------
        // raw[u][m]: table value (or tau) of interface k0+1+u, loaded one block ahead so the
        // L2/HBM latency hides behind the previous block's exp and recurrence work.
        Src raw[kU][MR];
        auto load_block = [&](int kfirst) {
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int kk = min(kfirst + u, L - 1);
                if (FUSED) {
#pragma unroll
                    for (int m = 0; m < MR; ++m) raw[u][m] = (Src)a.KL[koff[m * L + kk] + n];
                } else {
                    raw[u][0] = (Src)tau_col[(int64_t)kk * a.N + n];
                }
            }
        };
        Src Iprev;
        if (FUSED) {
            Real acc = Real(0);
#pragma unroll
            for (int m = 0; m < MR; ++m) acc += qg[m * L] * a.KL[koff[m * L] + n];
            Iprev = (Src)acc;
        } else {
            Iprev = (Src)tau_col[n];
        }
        load_block(1);
        Real E = Real(1), H = Real(0);
        for (int k0 = 0; k0 < nl; k0 += kU) {
            Src In[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                if (FUSED) {
                    const int kk = min(k0 + 1 + u, L - 1);
                    Real acc = Real(0);
#pragma unroll
                    for (int m = 0; m < MR; ++m) acc += qg[m * L + kk] * (Real)raw[u][m];
                    In[u] = (Src)acc;
                } else {
                    In[u] = raw[u][0];
                }
            }
            load_block(k0 + 1 + kU);
            Real t[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int k = k0 + u;
                const Src lo = u ? In[u - 1] : Iprev;
                Real dtau = FUSED ? (Real)((Real)a.hdP[min(k, nl - 1)] * (Real)(lo + In[u])) : (Real)(In[u] - lo);
                if (k >= nl) dtau = Real(0);
                t[u] = exp_neg(dtau);
                if (k < nl) ts[k * 32 + lane] = t[u];
            }
            Iprev = In[kU - 1];
            Real v[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int k = min(k0 + u, nl - 1);
                E *= t[u];
                Real val = bs * E;
                if (lw) {
                    const Real om = Real(1) - t[u];
                    const Real S = t[u] * H;
                    val += S + coef[k] * om;
                    H = S + coef[nl + k] * om;
                }
                v[u] = w * val;
            }
            const Real r = reduce8(v, lane);
            if (rlane && k0 + ridx < nl) bacc_dn[k0 + ridx + 1] += (double)r;
        }
        // ---- upward sweep, surface -> TOA ---------------------------------------------------------
        E = Real(1);
        Real S = Real(0);
        for (int k0 = nl - 1; k0 >= 0; k0 -= kU) {
            Real t[kU], v[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) t[u] = (k0 - u >= 0) ? ts[(k0 - u) * 32 + lane] : Real(1);
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int i = max(k0 - u, 0);
                E *= t[u];
                Real val = bg * E;
                if (lw) {
                    const Real om = Real(1) - t[u];
                    const Real y = t[u] * S;
                    val += y + coef[2 * nl + i] * om;
                    S = y + coef[3 * nl + i] * om;
                }
                v[u] = w * val;
            }
            const Real r = reduce8(v, lane);
            if (rlane && k0 - ridx >= 0) bacc_up[k0 - ridx] += (double)r;
        }
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
