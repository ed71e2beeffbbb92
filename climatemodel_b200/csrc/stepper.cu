// The adaptive time step that consumes the flux path's net flux, batched over columns on the
// device (Atmosphere.update_temp / update_time_step, base.py:166-246, without convective
// adjustment).  The reference keeps one column's state in the dict
// time_step_info (base.py:137-141); here it is a structure of arrays, one entry per column, and
// one warp advances one column:
//   tendency = g/c_p d(net)/dp per layer                                         (base.py:166-168)
//   updated layers = |net flux at the layer's first interface in the caller's level order| >
//   threshold, minus frozen layers                                               (base.py:171-172)
//   delta_net_flux = percentile of |net - previous net| (NumPy's linear method)  (base.py:173)
//   dt = DeltaT / |largest tendency among updated layers|, DeltaT adapted to oscillation /
//   stagnation of the layer that sets it, persistent offenders frozen            (base.py:197-246)
//   T += dt * tendency on the updated layers,  t += dt
// Levels are stored like everywhere in this library (index 0 = top of atmosphere).  The reference
// works on the arrays in the caller's order: RealGas keeps p_interface ascending (index 0 = top,
// real_gas.py:281), the grey model descending (index 0 = surface).  That order decides which
// interface gates layer i (net[i], the layer's first interface), the first-index tie-break of the
// arg-max and the numbering of MaxTendInd / RemoveInd; `top_first` selects it
// (cmb200_stepper_set_level_order; default 1 = RealGas).  The per-layer byte arrays (removed,
// active, conv_level) are always stored surface first, which is what the convective adjustment
// kernel reads.
#include "internal.cuh"

struct cmb200_stepper {
    cmb200_plan *plan = nullptr;
    int64_t n_col = 0;
    cmb::DevBuf<double> delta_T, max_tend, dt, prev_net;
    cmb::DevBuf<int32_t> max_tend_ind, n_same, n_same2, n_removed;
    cmb::DevBuf<uint8_t> removed;      // [n_col][L-1], reference level index
    cmb::DevBuf<uint8_t> active, conv_level;   // layers updated in the last step; layers the last convective adjustment moved
    cmb::DevBuf<int32_t> conv_status;
    cmb::DevBuf<double> p_centre, p_power;     // reference order (surface first): p and (p / p_ref)^alpha
    double temp_change = 1.0, delta_temp_change = 0.01;
    int top_first = 1;                         // reference level numbering: 1 = RealGas (index 0 = top), 0 = surface first
    const uint8_t *frozen = nullptr;           // device, [n_col]: columns the updates leave alone (cmb200_stepper_set_frozen)
};

namespace cmb {

struct StepperArgs {
    int L;
    int64_t n_col;
    const double *Pint;
    double g_over_cp, thresh, quantile, max_delta_T, delta_T_step;
    int changing_tau, convective, top_first;
    double conv_mult;
    const uint8_t *conv_level;
    uint8_t *active;
    const uint8_t *frozen;
    const double *net;
    double *T, *t, *dnf;
    double *delta_T, *max_tend, *dt, *prev_net;
    int32_t *max_tend_ind, *n_same, *n_same2, *n_removed;
    uint8_t *removed;
};

__device__ __forceinline__ double sign_of(double x) { return (x > 0.0) - (x < 0.0); }

__global__ void stepper_update_kernel(const StepperArgs a) {
    extern __shared__ double smem[];
    const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t col = (int64_t)blockIdx.x * warps + warp;
    if (col >= a.n_col || (a.frozen && a.frozen[col])) return;
    const int L = a.L, nl = L - 1;
    double *diff = smem + (size_t)warp * 2 * L, *tend = diff + L;       // |net - prev| ; tendency per layer
    const double *net = a.net + col * L;
    double *prev = a.prev_net + col * L;
    uint8_t *removed = a.removed + col * nl;
    const double t_now = a.t[col];
    const bool normal = t_now > 0.0 && !a.changing_tau;

    for (int i = lane; i < L; i += 32) diff[i] = fabs(net[i] - prev[i]);
    for (int k = lane; k < nl; k += 32)
        tend[k] = a.g_over_cp * (net[k] - net[k + 1]) / (a.Pint[k] - a.Pint[k + 1]);
    __syncwarp();

    // ---- delta_net_flux: NumPy percentile, method 'linear' (rank by counting, then _lerp) --------------
    double dnf = 1e6;
    if (normal) {
        const double virt = (L - 1) * a.quantile;
        const int lo = (int)floor(virt), hi = min(lo + 1, L - 1);
        const double gamma = virt - lo;
        double vlo = 0.0, vhi = 0.0;
        for (int j = lane; j < L; j += 32) {
            const double x = diff[j];
            int rank = 0;
            for (int i = 0; i < L; ++i) rank += (diff[i] < x) || (diff[i] == x && i < j);
            if (rank == lo) vlo = x;
            if (rank == hi) vhi = x;
        }
        vlo = warp_sum(vlo);            // exactly one lane holds each of them, the others add zeros
        vhi = warp_sum(vhi);
        const double d = vhi - vlo;
        dnf = gamma >= 0.5 ? __dsub_rn(vhi, __dmul_rn(d, 1.0 - gamma)) : __dadd_rn(vlo, __dmul_rn(d, gamma));   // no FMA: NumPy rounds twice
    }

    // ---- layers to update and the one with the largest |tendency| (first in the reference's order) ----
    const int tf = a.top_first;
    double best = -1.0;
    int best_r = -1, n_active = 0;
    for (int k = lane; k < nl; k += 32) {
        const int rs = nl - 1 - k;                                      // storage index of the byte arrays
        const int r = tf ? k : rs;                                      // reference level index
        const bool act = normal ? (fabs(net[tf ? k : k + 1]) > a.thresh && !removed[rs]) : true;
        if (act) {
            ++n_active;
            const double m = fabs(tend[k]);
            if (m > best || (m == best && r < best_r)) {
                best = m;
                best_r = r;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int orr = __shfl_xor_sync(0xffffffffu, best_r, o);
        n_active += __shfl_xor_sync(0xffffffffu, n_active, o);
        if (orr >= 0 && (best_r < 0 || ob > best || (ob == best && orr < best_r))) {
            best = ob;
            best_r = orr;
        }
    }

    // ---- update_time_step: one lane runs the scalar state machine, all lanes get dt --------------------
    double dt = a.dt[col];
    if (n_active > 0 && lane == 0) {
        double DeltaT = a.delta_T[col];
        int n1 = a.n_same[col], n2 = a.n_same2[col], n_rem = a.n_removed[col];
        const int best_k = tf ? best_r : nl - 1 - best_r, best_rs = nl - 1 - best_k;
        const double tnd = tend[best_k];
        const bool same_level = best_r == a.max_tend_ind[col];
        const bool same_sign = sign_of(tnd) == sign_of(a.max_tend[col]);
        if (same_level && !same_sign) {
            DeltaT = DeltaT - a.delta_T_step;
            n1 += 1;
            if (DeltaT < a.delta_T_step) DeltaT = a.delta_T_step;
        } else if (same_level && same_sign && DeltaT < a.max_delta_T) {
            DeltaT = DeltaT + a.delta_T_step;
            n2 += 1;
            if (DeltaT > a.max_delta_T) DeltaT = a.max_delta_T;
        } else {
            n1 = 0;
            n2 = 0;
        }
        if ((n1 > 1 && n2 > 10) || (n1 > 20 && n2 == 0) || (n_rem > 3 && n1 + n2 > 0)) {
            removed[best_rs] = 1;
            n_rem += 1;
            n1 = 0;
            n2 = 0;
        }
        dt = DeltaT / fabs(tnd);
        if (isinf(dt)) dt = 24.0 * 60.0 * 60.0;
        if (a.convective && a.conv_level[col * nl + best_rs]) dt = dt * a.conv_mult;    // base.py:181-184
        a.delta_T[col] = DeltaT;
        a.n_same[col] = n1;
        a.n_same2[col] = n2;
        a.n_removed[col] = n_rem;
        a.max_tend_ind[col] = best_r;
        a.max_tend[col] = tnd;
        a.dt[col] = dt;
    }
    dt = __shfl_sync(0xffffffffu, dt, 0);
    __syncwarp();

    // ---- T += dt * tendency on the layers chosen above (a layer frozen in this very step still moves) --
    double *T = a.T + col * nl;
    const int best_k_all = best_r < 0 ? -1 : (tf ? best_r : nl - 1 - best_r);
    for (int k = lane; k < nl; k += 32) {
        const int rs = nl - 1 - k;
        const bool was_removed_before = removed[rs] && k != best_k_all;
        const bool act = normal ? (fabs(net[tf ? k : k + 1]) > a.thresh && !was_removed_before) : true;
        if (act) T[k] = __dadd_rn(T[k], __dmul_rn(dt, tend[k]));
        if (a.active) a.active[col * nl + rs] = act;
    }
    for (int i = lane; i < L; i += 32) prev[i] = net[i];
    if (lane == 0) {
        a.t[col] = t_now + dt;
        if (a.dnf) a.dnf[col] = dnf;
    }
}

}  // namespace cmb

using namespace cmb;

extern "C" int cmb200_stepper_create(cmb200_stepper **stepper, cmb200_plan *plan, int64_t n_col,
                                     double temp_change, double delta_temp_change) {
    CMB_CHECK_ARG(stepper && plan, "null pointer");
    CMB_CHECK_ARG(n_col >= 0, "negative column count");
    CMB_CHECK_ARG(temp_change > 0 && delta_temp_change > 0, "temperature changes must be positive");
    CMB_CUDA(cudaSetDevice(plan->device));
    cmb200_stepper *s = new cmb200_stepper();
    s->plan = plan;
    s->n_col = n_col;
    s->temp_change = temp_change;
    s->delta_temp_change = delta_temp_change;
    *stepper = s;
    return cmb200_stepper_reset(s);
}

extern "C" int cmb200_stepper_reset(cmb200_stepper *s) {
    CMB_CHECK_ARG(s, "null pointer");
    CMB_CUDA(cudaSetDevice(s->plan->device));
    const size_t n = (size_t)s->n_col, L = (size_t)s->plan->L;
    if (n == 0) return CMB200_OK;
    std::vector<double> dT(n, s->temp_change), zero(n, 0.0);
    std::vector<int32_t> minus1(n, -1), izero(n, 0);
    CMB_CUDA(s->delta_T.upload(dT.data(), n));
    CMB_CUDA(s->max_tend.upload(zero.data(), n));
    CMB_CUDA(s->dt.upload(zero.data(), n));
    CMB_CUDA(s->max_tend_ind.upload(minus1.data(), n));
    CMB_CUDA(s->n_same.upload(izero.data(), n));
    CMB_CUDA(s->n_same2.upload(izero.data(), n));
    CMB_CUDA(s->n_removed.upload(izero.data(), n));
    if (s->prev_net.n < n * L) CMB_CUDA(s->prev_net.alloc(n * L));
    if (s->removed.n < n * (L - 1)) CMB_CUDA(s->removed.alloc(n * (L - 1)));
    if (s->active.n < n * (L - 1)) CMB_CUDA(s->active.alloc(n * (L - 1)));
    if (s->conv_level.n < n * (L - 1)) CMB_CUDA(s->conv_level.alloc(n * (L - 1)));
    if (s->conv_status.n < n) CMB_CUDA(s->conv_status.alloc(n));
    CMB_CUDA(cudaMemset(s->active.p, 0, n * (L - 1)));
    CMB_CUDA(cudaMemset(s->conv_level.p, 0, n * (L - 1)));
    CMB_CUDA(cudaMemset(s->conv_status.p, 0, n * sizeof(int32_t)));
    CMB_CUDA(cudaMemset(s->prev_net.p, 0, n * L * sizeof(double)));
    CMB_CUDA(cudaMemset(s->removed.p, 0, n * (L - 1)));
    CMB_CUDA(cudaDeviceSynchronize());
    return CMB200_OK;
}

extern "C" int cmb200_stepper_destroy(cmb200_stepper *s) {
    if (!s) return CMB200_OK;
    cudaSetDevice(s->plan->device);
    delete s;
    return CMB200_OK;
}

static int stepper_update(cmb200_stepper *s, const double *net, double *T, double *t, double *delta_net_flux,
                          double net_flux_thresh, double net_flux_percentile, int changing_tau, int convective,
                          double conv_mult, void *stream) {
    CMB_CHECK_ARG(s && net && T && t, "null pointer");
    CMB_CHECK_ARG(net_flux_percentile >= 0 && net_flux_percentile <= 100, "percentile must be within [0, 100]");
    if (s->n_col == 0) return CMB200_OK;
    CMB_CUDA(cudaSetDevice(s->plan->device));
    StepperArgs a;
    a.L = s->plan->L;
    a.n_col = s->n_col;
    a.Pint = s->plan->P.p;
    a.g_over_cp = s->plan->c.g / s->plan->c.c_p_dry;
    a.thresh = net_flux_thresh;
    a.quantile = net_flux_percentile / 100.0;
    a.max_delta_T = s->temp_change;
    a.delta_T_step = s->delta_temp_change;
    a.changing_tau = changing_tau;
    a.convective = convective;
    a.top_first = s->top_first;
    a.conv_mult = conv_mult;
    a.conv_level = s->conv_level.p;
    a.active = s->active.p;
    a.frozen = s->frozen;
    a.net = net;
    a.T = T;
    a.t = t;
    a.dnf = delta_net_flux;
    a.delta_T = s->delta_T.p;
    a.max_tend = s->max_tend.p;
    a.dt = s->dt.p;
    a.prev_net = s->prev_net.p;
    a.max_tend_ind = s->max_tend_ind.p;
    a.n_same = s->n_same.p;
    a.n_same2 = s->n_same2.p;
    a.n_removed = s->n_removed.p;
    a.removed = s->removed.p;
    const int warps = 4;
    const size_t smem = (size_t)warps * 2 * a.L * sizeof(double);
    stepper_update_kernel<<<(unsigned)((s->n_col + warps - 1) / warps), warps * 32, smem, (cudaStream_t)stream>>>(a);
    CMB_LAUNCHED();
    return CMB200_OK;
}

extern "C" int cmb200_stepper_update(cmb200_stepper *s, const double *net, double *T, double *t,
                                     double *delta_net_flux, double net_flux_thresh,
                                     double net_flux_percentile, int changing_tau, void *stream) {
    return stepper_update(s, net, T, t, delta_net_flux, net_flux_thresh, net_flux_percentile, changing_tau, 0, 1.0, stream);
}

extern "C" int cmb200_stepper_set_convection(cmb200_stepper *s, const double *p_centre, const double *p_power) {
    CMB_CHECK_ARG(s && p_centre && p_power, "null pointer");
    CMB_CUDA(cudaSetDevice(s->plan->device));
    const int nl = s->plan->L - 1;
    for (int r = 1; r < nl; ++r) CMB_CHECK_ARG(p_centre[r] < p_centre[r - 1], "p_centre must descend (surface first)");
    CMB_CUDA(s->p_centre.upload(p_centre, nl));
    CMB_CUDA(s->p_power.upload(p_power, nl));
    CMB_CUDA(cudaDeviceSynchronize());
    return CMB200_OK;
}

extern "C" int cmb200_stepper_update_convective(cmb200_stepper *s, const double *net, double *T, double *t,
                                                double *delta_net_flux, double net_flux_thresh,
                                                double net_flux_percentile, int changing_tau, double conv_thresh,
                                                double conv_t_multiplier, void *stream) {
    CMB_CHECK_ARG(s, "null pointer");
    if (s->n_col == 0) return CMB200_OK;
    if (!s->p_centre.p) {
        set_error("cmb200_stepper_set_convection has not been called");
        return CMB200_ERR_STATE;
    }
    int rc = stepper_update(s, net, T, t, delta_net_flux, net_flux_thresh, net_flux_percentile, changing_tau, 1,
                            conv_t_multiplier, stream);
    if (rc) return rc;
    return cmb::launch_convective_adjust(s->plan->L - 1, s->n_col, s->p_centre.p, s->p_power.p, conv_thresh, T,
                                         s->active.p, s->conv_level.p, s->conv_status.p, s->frozen, stream);
}

extern "C" int cmb200_stepper_set_level_order(cmb200_stepper *s, int top_first) {
    CMB_CHECK_ARG(s, "null pointer");
    s->top_first = top_first ? 1 : 0;
    return CMB200_OK;
}

extern "C" int cmb200_stepper_set_frozen(cmb200_stepper *s, const uint8_t *frozen) {
    CMB_CHECK_ARG(s, "null pointer");
    s->frozen = frozen;
    return CMB200_OK;
}

extern "C" int cmb200_stepper_state(const cmb200_stepper *s, double *delta_T, double *dt, int32_t *max_tend_ind,
                                    int32_t *n_removed) {
    CMB_CHECK_ARG(s, "null pointer");
    if (s->n_col == 0) return CMB200_OK;
    CMB_CUDA(cudaSetDevice(s->plan->device));
    CMB_CUDA(cudaDeviceSynchronize());
    const size_t n = (size_t)s->n_col;
    if (delta_T) CMB_CUDA(cudaMemcpy(delta_T, s->delta_T.p, n * sizeof(double), cudaMemcpyDeviceToHost));
    if (dt) CMB_CUDA(cudaMemcpy(dt, s->dt.p, n * sizeof(double), cudaMemcpyDeviceToHost));
    if (max_tend_ind) CMB_CUDA(cudaMemcpy(max_tend_ind, s->max_tend_ind.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost));
    if (n_removed) CMB_CUDA(cudaMemcpy(n_removed, s->n_removed.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost));
    return CMB200_OK;
}
