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
// Mapping: one warp owns one (column, part) task; lanes hold kChains chains each
// (consecutive lanes = consecutive wavenumber bins, so table / tau rows are read coalesced).
// The downward sweep (TOA -> surface) computes t_k and parks it in shared memory
// ts[k][chain][lane]; the upward sweep (surface -> TOA) replays it, so each layer costs one
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

__device__ __forceinline__ double exp_neg(double x) { return exp(-x); }
__device__ __forceinline__ float exp_neg(float x) { return __expf(-x); }

template <typename Real, bool FUSED>
__global__ void __launch_bounds__(256, 1) flux_kernel(const FluxArgs<Real> a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int R = kChains;
    using Src = typename std::conditional<FUSED, Real, double>::type;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int L = a.L, nl = a.L - 1, M = a.M;
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
        // ---- this segment's chains -----------------------------------------------------------
        const int len = a.band_len[j];
        Real w[R];
        int64_t n[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            int ci = c0 + r * 32 + lane;
            bool act = ci < len;
            w[r] = act ? a.wchain[a.chain_off[j] + ci] : Real(0);
            n[r] = a.band_start[j] + (act ? ci : 0);
        }
        if (c0 == 0 && lane == 0) {   // boundary terms: Trans(i,i) = sum of the nu-weights
            bacc_up[L - 1] += (double)bg * a.band_sumw[j];
            bacc_dn[0] += (double)bs * a.band_sumw[j];
        }
        // ---- downward sweep, TOA -> surface ----------------------------------------------------
        Real E[R], H[R];
        Src Icur[R], Inext[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            E[r] = Real(1);
            H[r] = Real(0);
            if (FUSED) {
                Real acc = Real(0);
                for (int m = 0; m < M; ++m) acc += qg[m * L] * a.KL[koff[m * L] + n[r]];
                Icur[r] = (Src)acc;
                acc = Real(0);
                for (int m = 0; m < M; ++m) acc += qg[m * L + 1] * a.KL[koff[m * L + 1] + n[r]];
                Inext[r] = (Src)acc;
            } else {
                Icur[r] = (Src)tau_col[n[r]];
                Inext[r] = (Src)tau_col[a.N + n[r]];
            }
        }
        for (int k = 0; k < nl; ++k) {
            Src Iload[R];
            const int kk = (k + 2 < L) ? k + 2 : L - 1;   // prefetch one layer ahead
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (FUSED) {
                    Real acc = Real(0);
                    for (int m = 0; m < M; ++m) acc += qg[m * L + kk] * a.KL[koff[m * L + kk] + n[r]];
                    Iload[r] = (Src)acc;
                } else {
                    Iload[r] = (Src)tau_col[(int64_t)kk * a.N + n[r]];
                }
            }
            const Real dl = lw ? coef[k] : Real(0), dg = lw ? coef[nl + k] : Real(0);
            const Real hdp = (Real)a.hdP[k];
            Real v = Real(0);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                Real dtau = FUSED ? (Real)(hdp * (Real)(Icur[r] + Inext[r])) : (Real)(Inext[r] - Icur[r]);
                Real t = exp_neg(dtau);
                ts[(k * R + r) * 32 + lane] = t;
                E[r] *= t;
                Real val = bs * E[r];
                if (lw) {
                    Real om = Real(1) - t;
                    Real S = t * H[r];
                    val += S + dl * om;
                    H[r] = S + dg * om;
                }
                v += w[r] * val;
                Icur[r] = Inext[r];
                Inext[r] = Iload[r];
            }
            v = warp_sum(v);
            if (lane == 0) bacc_dn[k + 1] += (double)v;
        }
        // ---- upward sweep, surface -> TOA -------------------------------------------------------
        Real S[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            E[r] = Real(1);
            S[r] = Real(0);
        }
        for (int i = nl - 1; i >= 0; --i) {
            const Real uf = lw ? coef[2 * nl + i] : Real(0), ug = lw ? coef[3 * nl + i] : Real(0);
            Real v = Real(0);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                Real t = ts[(i * R + r) * 32 + lane];
                E[r] *= t;
                Real val = bg * E[r];
                if (lw) {
                    Real om = Real(1) - t;
                    Real y = t * S[r];
                    val += y + uf * om;
                    S[r] = y + ug * om;
                }
                v += w[r] * val;
            }
            v = warp_sum(v);
            if (lane == 0) bacc_up[i] += (double)v;
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
