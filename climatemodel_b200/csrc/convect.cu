// Dry convective adjustment, batched over columns on the device (convective_adjustment_single,
// convective_adjustment.py:7-135).  One warp per column; the data-dependent
// control flow (groups of unstable levels, grow-until-stable searches, accept / ignore) is warp
// uniform, the per-level arithmetic is spread over the lanes.
//
// Results are meant to be EQUAL to the NumPy reference, so the operations are rounded one by one
// in NumPy's order and its reductions are reproduced: np.trapz = pairwise sum (8 interleaved
// partial sums combined as a tree, then the tail) of d_i (y_{i+1} + y_i) / 2; np.median by rank.
// (p / p_ref)^alpha is taken from the caller (pw, evaluated once on the host), because pow is the
// one operation that may differ in the last place.
//
// Index r counts levels from the surface like the reference; the device arrays of this library
// start at the top of the atmosphere, so level r lives at k = nl - 1 - r.
#include "internal.cuh"

namespace cmb {

// NumPy's pairwise summation of a[0..n) for n <= 128 (numpy/_core/src/umath/loops_utils.h.src).
__device__ double np_sum_block(const double *a, int n, int lane) {
    double res = 0.0;
    if (n < 8) {
        if (lane == 0)
            for (int i = 0; i < n; ++i) res = __dadd_rn(res, a[i]);
        return __shfl_sync(0xffffffffu, res, 0);
    }
    const int body = n - n % 8;
    double r = 0.0;
    if (lane < 8) {
        r = a[lane];
        for (int i = 8; i < body; i += 8) r = __dadd_rn(r, a[i + lane]);
    }
    r = __dadd_rn(r, __shfl_down_sync(0xffffffffu, r, 1));      // (r0+r1) at lane 0, (r2+r3) at lane 2, ...
    r = __dadd_rn(r, __shfl_down_sync(0xffffffffu, r, 2));      // ((r0+r1)+(r2+r3)) at lane 0, ... at lane 4
    r = __dadd_rn(r, __shfl_down_sync(0xffffffffu, r, 4));
    if (lane == 0) {
        res = r;
        for (int i = body; i < n; ++i) res = __dadd_rn(res, a[i]);
    }
    return __shfl_sync(0xffffffffu, res, 0);
}
__device__ double np_sum(const double *a, int n, int lane) {
    if (n <= 128) return np_sum_block(a, n, lane);
    int n2 = n / 2;
    n2 -= n2 % 8;
    return __dadd_rn(np_sum(a, n2, lane), np_sum(a + n2, n - n2, lane));
}

// -np.trapz(y, p): terms d_i (y_{i+1} + y_i) / 2 written to `terms`, then summed like NumPy.
__device__ double enthalpy_of(const double *y, const double *p, double *terms, int nl, int lane) {
    for (int i = lane; i < nl - 1; i += 32)
        terms[i] = __ddiv_rn(__dmul_rn(__dsub_rn(p[i + 1], p[i]), __dadd_rn(y[i + 1], y[i])), 2.0);
    __syncwarp();
    const double s = np_sum(terms, nl - 1, lane);
    __syncwarp();
    return -s;
}

__global__ void convective_adjust_kernel(int nl, int64_t n_col, const double *__restrict__ p_ref,
                                         const double *__restrict__ pw_ref, double small, double conv_thresh,
                                         double *__restrict__ T_dev, const uint8_t *__restrict__ active,
                                         uint8_t *__restrict__ conv_level, int32_t *__restrict__ status,
                                         const uint8_t *__restrict__ frozen) {
    extern __shared__ double smem[];
    const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t col = (int64_t)blockIdx.x * warps + warp;
    const unsigned full = 0xffffffffu;
    double *p = smem, *pw = p + nl;                                          // shared by the CTA's warps
    double *T = pw + nl + (size_t)warp * 5 * nl, *th = T + nl, *lo = th + nl, *up = lo + nl, *terms = up + nl;
    uint8_t *flags = (uint8_t *)(smem + 2 * nl + (size_t)warps * 5 * nl) + (size_t)warp * 2 * nl;
    uint8_t *cur = flags, *ign = flags + nl;
    for (int r = threadIdx.x; r < nl; r += blockDim.x) {
        p[r] = p_ref[r];
        pw[r] = pw_ref[r];
    }
    __syncthreads();
    if (col >= n_col || (frozen && frozen[col])) return;
    double *Tcol = T_dev + col * nl;
    for (int r = lane; r < nl; r += 32) {
        T[r] = Tcol[nl - 1 - r];
        ign[r] = 0;
    }
    __syncwarp();

    // delta_T_thresh = median(T) / 4  (np.median: middle of the sorted values, mean of two when even)
    double m_lo = 0.0, m_hi = 0.0;
    {
        const int k_lo = (nl - 1) / 2, k_hi = nl / 2;
        for (int j = lane; j < nl; j += 32) {
            const double x = T[j];
            int rank = 0;
            for (int i = 0; i < nl; ++i) rank += (T[i] < x) || (T[i] == x && i < j);
            if (rank == k_lo) m_lo = x;
            if (rank == k_hi) m_hi = x;
        }
        m_lo = warp_sum(m_lo);
        m_hi = warp_sum(m_hi);
    }
    const double median = (nl & 1) ? m_lo : __ddiv_rn(__dadd_rn(__dadd_rn(0.0, m_lo), m_hi), 2.0);
    const double thresh = __ddiv_rn(median, 4.0);

    auto refresh_theta = [&]() {
        for (int r = lane; r < nl; r += 32) th[r] = __ddiv_rn(T[r], pw[r]);
        __syncwarp();
    };
    // unstable where potential temperature decreases with height; the last level copies its neighbour
    auto find_unstable = [&]() -> bool {
        bool any = false;
        for (int r0 = 0; r0 < nl; r0 += 32) {
            const int r = r0 + lane;
            bool u = false;
            if (r < nl) {
                const int q = min(r, nl - 2);
                u = __dsub_rn(th[q + 1], th[q]) < -small && !ign[r];
                cur[r] = u;
            }
            any |= __any_sync(full, u);
        }
        __syncwarp();
        return any;
    };
    refresh_theta();
    int st = 0;
    for (int sweep = 0; find_unstable(); ++sweep) {
        if (sweep > 8 * nl) {             // the reference would loop on; give up and say so
            st = 2;
            break;
        }
        const double enthalpy = enthalpy_of(T, p, terms, nl, lane);
        int a = 0;
        while (a < nl) {
            if (!cur[a]) {                                   // warp-uniform scan for the next group a..b
                ++a;
                continue;
            }
            int b = a;
            while (b + 1 < nl && cur[b + 1]) ++b;
            const int top = b + 1;
            if (top >= nl) {             // the reference indexes one level above the group: no such level
                for (int r = a + lane; r <= b; r += 32) ign[r] = 1;
                st = 1;
                __syncwarp();
                a = b + 1;
                continue;
            }
            // (a) pull the region down to the theta just above it, grown downwards until stable
            const double theta_low = th[top];
            int first = -1;
            for (int r = lane; r < top; r += 32)
                if (th[r] < theta_low) first = max(first, r);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) first = max(first, __shfl_xor_sync(full, first, o));
            first += 1;
            // (b) lift the region up to the theta at its base, grown upwards until stable
            const double theta_up = th[a];
            int last = nl;
            for (int r = a + 1 + lane; r < nl; r += 32)
                if (th[r] > theta_up) last = min(last, r);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) last = min(last, __shfl_xor_sync(full, last, o));
            if (last == nl) last = nl - 1;
            for (int r = lane; r < nl; r += 32) {
                lo[r] = __dmul_rn((r >= first && r <= top) ? theta_low : th[r], pw[r]);
                up[r] = __dmul_rn((r >= a && r <= last) ? theta_up : th[r], pw[r]);
            }
            __syncwarp();
            const double h_lo = enthalpy_of(lo, p, terms, nl, lane), h_up = enthalpy_of(up, p, terms, nl, lane);
            const double beta = __ddiv_rn(__dsub_rn(enthalpy, h_lo), __dsub_rn(h_up, h_lo));
            const double one_minus = __dsub_rn(1.0, beta);
            double worst = 0.0;
            bool nan_seen = false;
            for (int r = lane; r < nl; r += 32) {
                const double t_new = __dadd_rn(__dmul_rn(beta, up[r]), __dmul_rn(one_minus, lo[r]));
                lo[r] = t_new;
                const double d = fabs(__dsub_rn(t_new, T[r]));
                nan_seen |= isnan(d);
                worst = fmax(worst, d);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) worst = fmax(worst, __shfl_xor_sync(full, worst, o));
            nan_seen = __any_sync(full, nan_seen);
            __syncwarp();
            if (!nan_seen && worst < thresh) {
                for (int r = lane; r < nl; r += 32) T[r] = lo[r];
            } else {
                for (int r = a + lane; r <= b; r += 32) ign[r] = 1;
            }
            __syncwarp();
            refresh_theta();
            a = b + 1;
        }
    }
    for (int r = lane; r < nl; r += 32) {
        const int k = nl - 1 - r;
        const double before = Tcol[k];
        if (conv_level) conv_level[col * nl + r] = (!active || active[col * nl + r]) && fabs(__dsub_rn(T[r], before)) > conv_thresh;
        Tcol[k] = T[r];
    }
    if (status && lane == 0) status[col] = st;
}

// frozen[n_col] (device, may be NULL): columns to leave alone.
int launch_convective_adjust(int32_t n_levels, int64_t n_col, const double *p_centre, const double *p_power,
                             double conv_thresh, double *T, const uint8_t *active, uint8_t *convective_levels,
                             int32_t *status, const uint8_t *frozen, void *stream) {
    CMB_CHECK_ARG(p_centre && p_power && T, "null pointer");
    CMB_CHECK_ARG(n_levels >= 3 && n_levels <= 256, "between 3 and 256 levels supported, got %d", n_levels);
    CMB_CHECK_ARG(n_col >= 0, "negative column count");
    if (n_col == 0) return CMB200_OK;
    const int warps = 4;
    const size_t smem = (size_t)(2 + 5 * warps) * n_levels * sizeof(double) + (size_t)warps * 2 * n_levels;
    CMB_CUDA(cudaFuncSetAttribute(convective_adjust_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    convective_adjust_kernel<<<(unsigned)((n_col + warps - 1) / warps), warps * 32, smem, (cudaStream_t)stream>>>(
        n_levels, n_col, p_centre, p_power, 1e-10, conv_thresh, T, active, convective_levels, status, frozen);
    CMB_LAUNCHED();
    return CMB200_OK;
}

}  // namespace cmb

using namespace cmb;

extern "C" int cmb200_convective_adjust(int32_t n_levels, int64_t n_col, const double *p_centre,
                                        const double *p_power, double conv_thresh, double *T,
                                        const uint8_t *active, uint8_t *convective_levels, int32_t *status,
                                        void *stream) {
    return launch_convective_adjust(n_levels, n_col, p_centre, p_power, conv_thresh, T, active, convective_levels,
                                    status, nullptr, stream);
}
