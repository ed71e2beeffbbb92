// Host drivers of the flux path: level-table construction, plan-grid optical depth, and the
// cmb200_flux / cmb200_flux_host entry points (RealGas.get_flux + net flux + heating rate).
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "flux_kernels.cuh"

namespace cmb {

static int sm_count(int device) {
    static int cached[64] = {0};
    if (device < 64 && cached[device]) return cached[device];
    int n = 148;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
    if (const char *e = getenv("CMB200_SM_LIMIT")) n = std::max(1, std::min(n, atoi(e)));     // measurement aid: a smaller persistent grid
    if (device < 64) cached[device] = n;
    return n;
}

#ifndef CMB_FLUX_SLICE_ONLY      // (launches a kernel of tables.cu: main translation unit only)
static int ensure_tidx(cmb200_plan *p, int64_t n_col, const double *T_tau, const double *T_g,
                       cudaStream_t s) {
    size_t need = (size_t)p->M * n_col * p->L;
    if (p->tidx.n < need) CMB_CUDA(p->tidx.alloc(need));
    for (int m = 0; m < p->M; ++m) {
        int64_t n = n_col * p->L;
        tidx_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(T_tau, T_g, n_col, p->L,
                                                               p->T_axes.p + (size_t)m * p->nT_pad,
                                                               p->nT_pad, p->tidx.p + (size_t)m * n);
        CMB_LAUNCHED();
    }
    return CMB200_OK;
}
#endif

// Relative cost of one pass of the flux kernel, used to balance the warps of a CTA: a long-wave
// pass carries the emission recurrences, and every (part, band) run ends with a replay-only pass.
static double cost_knob(const char *name, double dflt) {
    const char *e = getenv(name);
    return e && *e ? atof(e) : dflt;
}

// Work decomposition: W warps per CTA (from the shared-memory budget), C CTAs per column (more
// than one only when there are fewer columns than resident CTAs), parts = C * W contiguous runs
// of segments of equal estimated cost, and one partial-sum slot per (part, band) run.
// Pass costs in units of one short-wave fresh pass, fitted to per-warp cycle counts (CMB200_DEBUG_TIMING) of the
// kernel family: a long-wave fresh pass, and what a (part, band) run adds at its end (replay-only pass, write-out
// of the partial sums) for short- and long-wave bands.  Model 1 = direct short-wave passes with the tensor-memory
// lane sum (their short-wave pass is cheaper, so everything else weighs more) in FP64, model 2 = the same in FP32
// (a long-wave pass, still the round-1 form, weighs two short-wave ones), model 3 = FP32 with two segments per
// short-wave pass (the unit is then half a pair pass; part boundaries inside a short-wave band are moved so that a
// part's share of the band is a whole number of pairs), model 0 = the other kernels.
struct PassCosts {
    double lw, run_sw, run_lw;
};
static PassCosts pass_costs(int model) {
    PassCosts c = model == 1   ? PassCosts{1.7, 0.65, 0.7}
                  : model == 2 ? PassCosts{2.0, 0.65, 0.65}
                  : model == 3 ? PassCosts{1.7, 1.0, 1.8}
                               : PassCosts{1.5, 0.4, 0.6};
    c.lw = cost_knob("CMB200_COST_LW", c.lw);
    c.run_sw = cost_knob("CMB200_COST_RUN_SW", c.run_sw);
    c.run_lw = cost_knob("CMB200_COST_RUN_LW", c.run_lw);
    return c;
}

static int ensure_partition(cmb200_plan *p, int64_t n_col, int W, int resident_ctas, int cost_model) {
    Partition &pt = p->partition;
    // CTAs per column: with few columns a column is split over C CTAs so that the n_col * C tasks fill
    // the resident CTAs in whole waves -- pick the C with the least (waves / C), a little bias to small C.
    int C = 1;
    if (n_col < 4 * (int64_t)resident_ctas) {
        const int64_t most = std::max<int64_t>(1, p->n_seg / (2 * W));   // keep two segments per warp
        const int64_t top = std::min<int64_t>(most, std::max<int64_t>(4, 2 * ((resident_ctas + n_col - 1) / n_col)));
        double best = 0.0;
        for (int64_t c = 1; c <= top; ++c) {
            const int64_t waves = (n_col * c + resident_ctas - 1) / resident_ctas;
            const double t = (double)waves / c * (1.0 + 0.005 * c);
            if (c == 1 || t < best) {
                best = t;
                C = (int)c;
            }
        }
    }
    if (pt.n_col >= 0 && pt.warps == W && pt.ctas_per_col == C && pt.cost_model == cost_model) return CMB200_OK;   // same decomposition
    const int parts = C * W, n_seg = p->n_seg;
    const PassCosts pc = pass_costs(cost_model);
    std::vector<double> prefix(n_seg + 1, 0.0);
    int band_first = 0;
    for (int s = 0; s < n_seg; ++s) {
        const int j = p->hseg_band[s];
        const bool band_start = s == 0 || p->hseg_band[s - 1] != j;
        if (band_start) band_first = s;
        double c = p->hsw[j] ? 1.0 : pc.lw;
        if (cost_model == 3 && p->hsw[j]) c = ((s - band_first) & 1) ? 0.0 : 2.0;     // a pair pass starts at even offsets
        if (band_start) c += p->hsw[j] ? pc.run_sw : pc.run_lw;
        prefix[s + 1] = prefix[s] + c;
    }
    std::vector<int32_t> seg0(parts + 1), slot0(parts), bslot(p->B + 1, 0);
    for (int i = 0; i <= parts; ++i) {
        const double want = prefix[n_seg] * i / parts;
        int s = (int)(std::lower_bound(prefix.begin(), prefix.end(), want - 1e-9) - prefix.begin());
        if (s > 0 && want - prefix[s - 1] < prefix[s] - want) --s;       // nearer boundary
        seg0[i] = std::min(std::max(s, i ? seg0[i - 1] : 0), n_seg);
    }
    seg0[0] = 0;
    seg0[parts] = n_seg;
    if (cost_model == 3)
        for (int i = 1; i < parts; ++i) {
            const int b = seg0[i];
            if (b <= seg0[i - 1] || b >= n_seg) continue;
            const int j = p->hseg_band[b];
            if (!p->hsw[j] || p->hseg_band[b - 1] != j) continue;        // not inside a short-wave band
            int first = b;                                              // start of the previous part's share of band j
            while (first > seg0[i - 1] && p->hseg_band[first - 1] == j) --first;
            if ((b - first) & 1) seg0[i] = std::min(b + 1, seg0[i + 1]);
        }
    std::vector<int32_t> slot_band;
    for (int i = 0; i < parts; ++i) {
        slot0[i] = (int32_t)slot_band.size();
        int last = -1;
        for (int s = seg0[i]; s < seg0[i + 1]; ++s)
            if (p->hseg_band[s] != last) {
                last = p->hseg_band[s];
                slot_band.push_back(last);
            }
    }
    for (int j = 0; j <= p->B; ++j)
        bslot[j] = (int32_t)(std::lower_bound(slot_band.begin(), slot_band.end(), j) - slot_band.begin());
    pt.warps = W;
    pt.cost_model = cost_model;
    pt.ctas_per_col = C;
    pt.parts = parts;
    pt.n_slots = (int)slot_band.size();
    CMB_CUDA(pt.part_seg0.upload(seg0.data(), seg0.size()));
    CMB_CUDA(pt.part_slot0.upload(slot0.data(), slot0.size()));
    CMB_CUDA(pt.band_slot0.upload(bslot.data(), bslot.size()));
    std::vector<double> sdelta(slot_band.size() + 1, 0.0);
    for (size_t q = 0; q < slot_band.size(); ++q) sdelta[q] = p->hdelta[slot_band[q]];
    CMB_CUDA(pt.slot_delta.upload(sdelta.data(), sdelta.size()));
    CMB_CUDA(cudaDeviceSynchronize());
    pt.n_col = n_col;
    return CMB200_OK;
}

// Warps per CTA of kernel variant V for L interfaces: bounded by shared memory (t-store, accumulators,
// coefficient rows) and, for the tensor-memory t-store, by the 512 columns of a lane quarter shared
// by the warps w, w + 4, ...  0 = the variant cannot run this L.
template <typename Real, int MM, int V>
static int variant_warps(int L, int max_smem, int *tmem_cpw, int *tmem_cols) {
    const int Lp = padded_layers(L), wpr = variant_tstore_words<Real>(V);
    *tmem_cpw = *tmem_cols = 0;
    int most = 8;
    if (V & kVarTmem) {
        const int per_quarter = std::min(variant_threads<Real>(V) / 128, 512 / (Lp * wpr + ((V & kVarRedTmem) ? 2 * kU : 0)));
        most = 4 * per_quarter;
        const int cpw = Lp * wpr;
        int cols = 32;
        while (cols < per_quarter * cpw) cols *= 2;
        if ((V & kVarRedTmem) && cols < 512 && cols < per_quarter * (cpw + 2 * kU)) cols *= 2;   // exchange columns
        *tmem_cpw = cpw;
        *tmem_cols = cols;
    }
    const CtaSmem<Real, MM, V> lay(L);
    int W = (size_t)max_smem > lay.shared_total
                ? (int)std::min<size_t>(most, ((size_t)max_smem - lay.shared_total) / lay.warp_total)
                : 0;
    if (V & kVarTmem) W = W / 4 * 4;                     // whole lane quarters
    return W;
}

// Kernel variants (flux_kernels.cuh).  Every (precision, molecule count) has the plain kernel (0) and
// the default one; further variants are experiments, compiled for FP64 with 0 or 2 molecules only
// when the build names them (CMB200_EXPERIMENTS="17,21" python -m climatemodel_b200.csrc.build) and
// selected at run time with CMB200_FLUX_VARIANT (tools/ab_flux_variants.py).
// Default variant per arithmetic type and optical-depth source.  Resident tau: the round-1 kernel (kReferenceVariant:
// t-store in tensor memory, staged coefficient rows; bit-identical to the plain kernel 0), in FP64 with the shared-memory
// exp table.  Table-fused path: the direct short-wave passes with the tensor-memory lane sums on top of that, in FP32
// carrying two segments per pass (kVarPair).
constexpr int kReferenceVariant = kVarTmem | kVarCoefSmem;
template <typename Real, int MM>
constexpr int default_variant() {
    constexpr int direct = kVarSwDirect | kVarRedTmem | kVarSmallSt;      // table-fused path: direct short-wave passes
    if (!std::is_same<Real, double>::value) return MM > 0 ? (kReferenceVariant | direct | kVarPair) : kReferenceVariant;
    return MM > 0 ? (kReferenceVariant | kVarExpLds | direct) : (kReferenceVariant | kVarExpLds);
}
#ifndef CMB_FLUX_EXPERIMENTS
#define CMB_FLUX_EXPERIMENTS(X)
#endif
static int flux_variant() {          // -1: the default of the (type, source) at hand
    const char *e = getenv("CMB200_FLUX_VARIANT");
    return e && *e ? atoi(e) : -1;
}

// The level table regathered into segment space (internal.cuh) for the fused path; CMB200_SEGMENT_TABLE=0
// keeps the kernel on the bin-space table (A/B), and so does a failed allocation.
bool ensure_segment_table_f64(cmb200_plan *p, cudaStream_t s);      // defined in the main translation unit (they launch a kernel of tables.cu)
bool ensure_segment_table_f32(cmb200_plan *p, cudaStream_t s);
#ifndef CMB_FLUX_SLICE_ONLY
template <typename Real>
static bool ensure_segment_table(cmb200_plan *p, cudaStream_t s) {
    const char *e = getenv("CMB200_SEGMENT_TABLE");
    if ((e && *e == '0') || p->n_seg == 0) return false;
    if (p->ks_valid) return true;
    // one segment of zero padding per row: a pair pass (kVarPair) reads the segment after its last one, with weight 0
    const int64_t spitch = ((int64_t)p->n_seg + 1) * 32, rows = (int64_t)p->M * p->L * p->nT_pad;
    const Real *KL;
    Real *KS;
    if constexpr (std::is_same<Real, double>::value) {
        if (p->KS64.alloc((size_t)rows * spitch) != cudaSuccess) return cudaGetLastError(), false;
        KL = p->KL64.p, KS = p->KS64.p;
        if (cudaMemsetAsync(KS, 0, (size_t)rows * spitch * sizeof(Real), s) != cudaSuccess) return false;
    } else {
        if (p->KS32.alloc((size_t)rows * spitch) != cudaSuccess) return cudaGetLastError(), false;
        KL = p->KL32.p, KS = p->KS32.p;
        if (cudaMemsetAsync(KS, 0, (size_t)rows * spitch * sizeof(Real), s) != cudaSuccess) return false;
    }
    dim3 grid((unsigned)((spitch + 255) / 256), (unsigned)std::min<int64_t>(rows, 65535));
    segment_table_kernel<Real><<<grid, 256, 0, s>>>(KL, rows, p->pitch, p->seg_bin.p, (int64_t)p->n_seg * 32, spitch, KS);
    if (cudaGetLastError() != cudaSuccess) return false;
    g_launches.fetch_add(1);
    p->spitch = spitch;
    p->ks_valid = true;
    return true;
}
bool ensure_segment_table_f64(cmb200_plan *p, cudaStream_t s) { return ensure_segment_table<double>(p, s); }
bool ensure_segment_table_f32(cmb200_plan *p, cudaStream_t s) { return ensure_segment_table<float>(p, s); }
#endif

template <typename Real, int MM, int V>
static int launch_flux(cmb200_plan *p, const cmb200_flux_io *io, const double *Tint, cudaStream_t s) {
    const int L = p->L, Lp = padded_layers(L);
    const bool seg_space = MM > 0 && (std::is_same<Real, double>::value ? ensure_segment_table_f64(p, s) : ensure_segment_table_f32(p, s));
    const int64_t pitch = seg_space ? p->spitch : p->pitch;
    const double span = MM > 0 ? (double)MM * L * p->nT_pad * (double)pitch * sizeof(Real) / 16.0 : (double)L * (double)p->N;
    if (span >= 4294967296.0) {
        set_error("table / tau rows exceed the 32-bit row offsets of the flux kernel (%.3g units)", span);
        return CMB200_ERR_LIMIT;
    }
    int max_smem = 0;
    CMB_CUDA(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, p->device));
    int tmem_cpw = 0, tmem_cols = 0;
    const int W = variant_warps<Real, MM, V>(L, max_smem, &tmem_cpw, &tmem_cols);
    if constexpr ((V & kVarPair) != 0) {      // pair passes: segment-space table and two warps per lane quarter, else one segment per pass
        if (!seg_space || W < 8 || (W / 4) * (tmem_cpw + 2 * kU) > tmem_cols) return launch_flux<Real, MM, V & ~kVarPair>(p, io, Tint, s);
    }
    if constexpr (V != 0) {      // the variant has to offer at least the warps of the plain kernel, else that one runs
        int c0, c1;
        if (W < variant_warps<Real, MM, 0>(L, max_smem, &c0, &c1)) return launch_flux<Real, MM, 0>(p, io, Tint, s);
    }
    const CtaSmem<Real, MM, V> lay(L, -1, p->B);
    if (W < 1) {
        set_error("%d levels need %zu bytes of shared memory per CTA; the device offers %d", L, lay.total(1), max_smem);
        return CMB200_ERR_LIMIT;
    }
    // A CTA that allocates tensor memory must be alone on its SM (a second one would wait for the
    // columns): ask for more than half of the shared memory.
    size_t smem = (V & kVarTmem) ? std::max(lay.total(W), (size_t)max_smem / 2 + 1024) : lay.total(W);
    if constexpr ((V & kVarRedTmem) != 0 && !std::is_same<Real, double>::value) {
        // FP32: no exp table, everything contiguous; the exchange columns follow the t-stores in tensor memory
        if (lay.total(W) + 1024 > (size_t)max_smem || (W / 4) * (tmem_cpw + 2 * kU) > tmem_cols)
            return launch_flux<Real, MM, V & ~(kVarSwDirect | kVarRedTmem | kVarSmallSt | kVarPair)>(p, io, Tint, s);
    } else if constexpr ((V & kVarSwDirect) != 0 && std::is_same<Real, double>::value) {
        // the exp table occupies shared-window addresses [xl_window, xl_window + 64 kB): the level tables and the
        // warps' regions have to end below it (the window starts at a small driver-reserved offset)
        constexpr unsigned win = xl_window(V);
        if (lay.total(W) + 2048 > win || (size_t)max_smem < (size_t)win + kXlBytes)
            return launch_flux<Real, MM, V & ~(kVarSwDirect | kVarRedTmem | kVarSmallSt | kVarPair)>(p, io, Tint, s);
        smem = (size_t)win + kXlBytes;
        if constexpr ((V & kVarRedTmem) != 0) {
            // the warps' coefficient rows sit above the table; the exchange columns follow the t-stores in tensor memory
            smem += (size_t)W * lay.coef_bytes;
            if (smem + 1024 > (size_t)max_smem || (W / 4) * (tmem_cpw + 2 * kU) > tmem_cols)
                return launch_flux<Real, MM, V & ~(kVarRedTmem | kVarSmallSt)>(p, io, Tint, s);
        }
    }
    CMB_CUDA(cudaFuncSetAttribute(flux_kernel<Real, MM, V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int ctas_per_sm = 1;
    CMB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, flux_kernel<Real, MM, V>, W * 32, smem));
    const int resident = sm_count(p->device) * std::max(1, ctas_per_sm);
    int rc = ensure_partition(p, io->n_col, W, resident,
                              (V & kVarPair) ? 3 : (V & kVarRedTmem) ? (std::is_same<Real, double>::value ? 1 : 2) : 0);
    if (rc) return rc;
    const Partition &pt = p->partition;
    if (pt.n_slots == 0) {     // no segment in this plan's band / segment range: every sum is empty
        const size_t n = (size_t)io->n_col;
        if (io->up) CMB_CUDA(cudaMemsetAsync(io->up, 0, n * L * p->B * sizeof(double), s));
        if (io->down) CMB_CUDA(cudaMemsetAsync(io->down, 0, n * L * p->B * sizeof(double), s));
        if (io->net) CMB_CUDA(cudaMemsetAsync(io->net, 0, n * L * sizeof(double), s));
        if (io->heating) CMB_CUDA(cudaMemsetAsync(io->heating, 0, n * (L - 1) * sizeof(double), s));
        return CMB200_OK;
    }
    const int C = pt.ctas_per_col;
    const int64_t n_tasks = io->n_col * C;
    const int64_t grid = std::min<int64_t>(n_tasks, resident);
    {
        const size_t regions = C == 1 ? (size_t)grid : (size_t)io->n_col;
        const size_t need = regions * pt.n_slots * 2 * L;
        if (p->scratch.n < need) CMB_CUDA(p->scratch.alloc(need));
        const size_t cneed = ((size_t)grid * p->B * 4 * Lp * sizeof(Real) + 7) / 8;
        if (p->coefbuf.n < cneed) {
            CMB_CUDA(p->coefbuf.alloc(cneed));
            CMB_CUDA(cudaMemsetAsync(p->coefbuf.p, 0, cneed * sizeof(double), s));   // the padding stays zero
        }
    }
    FluxArgs<Real> a;
    memset(&a, 0, sizeof(a));
    a.L = L;
    a.B = p->B;
    a.M = p->M;
    a.nT_pad = p->nT_pad;
    a.n_slots = pt.n_slots;
    a.band_lo = p->band_lo;
    a.band_hi = p->band_hi;
    a.ctas_per_col = C;
    a.N = p->N;
    a.pitch = pitch;
    a.seg_space = seg_space;
    a.dbg_row0 = getenv("CMB200_DEBUG_ROW0") != nullptr;
    a.n_col = io->n_col;
    a.n_tasks = n_tasks;
    a.geo = p->geo.p;
    a.Pint = p->P.p;
    a.hdP = p->hdP.p;
    a.band_start = p->band_start.p;
    a.band_len = p->band_len.p;
    a.band_centre = p->band_centre.p;
    a.band_delta = p->band_delta.p;
    a.band_sumw = p->band_sumw.p;
    a.band_sw = p->band_sw.p;
    a.wchain = std::is_same<Real, double>::value ? (const Real *)p->wchain64.p : (const Real *)p->wchain32.p;
    a.chain_off = p->chain_off.p;
    a.seg_w = std::is_same<Real, double>::value ? (const Real *)p->seg_w64.p : (const Real *)p->seg_w32.p;
    a.seg_bin = p->seg_bin.p;
    a.seg_band = p->seg_band.p;
    a.seg_c0 = p->seg_c0.p;
    a.part_seg0 = pt.part_seg0.p;
    a.part_slot0 = pt.part_slot0.p;
    a.band_slot0 = pt.band_slot0.p;
    a.slot_delta = pt.slot_delta.p;
    a.pc = PlanckConst{p->c.h_planck, p->c.speed_of_light, p->c.k_boltzmann};
    a.T_star = p->T_star;
    a.star_geom = p->R_star * p->R_star / (p->dist * p->dist);
    a.g = p->c.g;
    a.g_over_cp = p->c.g / p->c.c_p_dry;
    a.T = io->T;
    a.Tint = Tint;
    a.T_g = io->T_g;
    a.albedo = io->albedo;
    a.albedo_scalar = io->albedo_scalar;
    a.tau = io->tau;
    a.tau_id = io->tau_id;
    if (seg_space)
        a.KL = std::is_same<Real, double>::value ? (const Real *)p->KS64.p : (const Real *)p->KS32.p;
    else
        a.KL = std::is_same<Real, double>::value ? (const Real *)p->KL64.p : (const Real *)p->KL32.p;
    a.tidx = p->tidx.p;
    a.q = io->q;
    a.q_col_stride = io->q_per_column ? L : 0;
    a.q_mol_stride = io->q_per_column ? io->n_col * L : L;
    a.up = io->up;
    a.down = io->down;
    a.net = io->net;
    a.heating = io->heating;
    a.scratch = p->scratch.p;
    a.coefbuf = (Real *)p->coefbuf.p;
    if (!p->task_counter.p) CMB_CUDA(p->task_counter.alloc(1));
    CMB_CUDA(cudaMemsetAsync(p->task_counter.p, 0, sizeof(unsigned long long), s));
    a.task_counter = p->task_counter.p;

    const bool debug_timing = getenv("CMB200_DEBUG_TIMING") != nullptr;
    DevBuf<long long> dbg;
    if (debug_timing) {
        CMB_CUDA(dbg.alloc(64));
        a.dbg = dbg.p;
    }
    if constexpr ((V & (kVarExpLds | kVarSwDirect)) != 0) {
        if (!p->exp_tab.p) {      // 2^(j/512): exp2l in long double (64-bit significand), rounded once to double
            std::vector<double> tab(kXlN);
            for (int j = 0; j < kXlN; ++j) tab[j] = xl_table_entry((double)exp2l((long double)j / kXlN), j);
            CMB_CUDA(p->exp_tab.upload(tab.data(), tab.size()));
        }
        a.exp_tab = p->exp_tab.p;
    }
    a.tmem_cpw = tmem_cpw;
    a.tmem_cols = tmem_cols;
    flux_kernel<Real, MM, V><<<(unsigned)grid, W * 32, smem, s>>>(a);
    CMB_LAUNCHED();
    if (debug_timing) {   // per-warp sweep cycles of CTA 0's first column next to what the partition gave each warp
        long long h[64];
        CMB_CUDA(cudaStreamSynchronize(s));
        CMB_CUDA(cudaMemcpy(h, dbg.p, sizeof(h), cudaMemcpyDeviceToHost));
        std::vector<int32_t> seg0(pt.parts + 1);
        CMB_CUDA(cudaMemcpy(seg0.data(), pt.part_seg0.p, seg0.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
        for (int w = 0; w < W; ++w) {
            int n_sw = 0, n_lw = 0, r_sw = 0, r_lw = 0, last = -1;
            for (int q = seg0[w]; q < seg0[w + 1]; ++q) {
                const int j = p->hseg_band[q];
                (p->hsw[j] ? n_sw : n_lw)++;
                if (j != last) (p->hsw[j] ? r_sw : r_lw)++;
                last = j;
            }
            if (w == 0) fprintf(stderr, "cmb200 timing: prologue split: level tables (warp 0) %lld, coefficients (warp 1) %lld, both (warp 2) %lld, level tables (warp 3) %lld\n", h[60], h[61], h[62], h[63]);
            fprintf(stderr, "cmb200 timing: warp %d sweep %lld (prologue %lld, whole task %lld)  fresh sw %d lw %d  runs sw %d lw %d\n", w, h[w], h[32 + w], h[48 + w], n_sw, n_lw, r_sw, r_lw);
        }
    }
    if (C > 1) {
        if (p->band_hi > p->band_lo) {
            flux_band_sums_kernel<<<dim3((unsigned)io->n_col, (unsigned)(p->band_hi - p->band_lo)), std::min(512, (2 * L + 31) / 32 * 32), 0, s>>>(
                L, p->B, p->band_lo, pt.n_slots, pt.band_slot0.p, p->scratch.p, io->up, io->down);
            CMB_LAUNCHED();
        }
        flux_net_kernel<<<(unsigned)io->n_col, 128, L * sizeof(double), s>>>(
            L, p->band_lo, p->band_hi, pt.n_slots, pt.band_slot0.p, p->band_delta.p, p->scratch.p, p->P.p,
            a.g_over_cp, io->net, io->heating);
        CMB_LAUNCHED();
    }
    return CMB200_OK;
}

#ifndef CMB_FLUX_SLICE_ONLY
__global__ void exp_neg_test_kernel(const double *__restrict__ x, int64_t n, double *__restrict__ out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) out[i] = exp_neg(x[i]);
}
// The table routine of the direct short-wave passes (direct_exp), eight values per thread, its table built in the
// same place of the shared window as the flux kernel builds it.
__global__ void exp_neg_table_test_kernel(const double *__restrict__ x, int64_t n, double *__restrict__ out,
                                          const double *__restrict__ exp_tab) {
    constexpr int V = kVarExpLds | kVarSwDirect;
    const unsigned window0 = (unsigned)__cvta_generic_to_shared(smem_raw);
    double *xtab = (double *)(smem_raw + (xl_window(V) - window0));
    for (int i = threadIdx.x; i < kXlN * kXlCopies; i += blockDim.x) xtab[i] = exp_tab[i / kXlCopies];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const ExpAux xaux{xtab + (lane & (kXlCopies - 1)), xl_window(V) | ((unsigned)(lane & (kXlCopies - 1)) << 3)};
    for (int64_t base = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * kU; base < n; base += (int64_t)gridDim.x * blockDim.x * kU) {
        double xv[kU], t[kU], v[kU], c1 = 0.0;
#pragma unroll
        for (int u = 0; u < kU; ++u) xv[u] = x[base + u < n ? base + u : n - 1], v[u] = 0.0;
        direct_exp<V, false>(xv, t, xaux, lane, v, c1);
#pragma unroll
        for (int u = 0; u < kU; ++u)
            if (base + u < n) out[base + u] = t[u];
    }
}

#endif

template <typename Real, int MM>
static int launch_flux_variant(int variant, cmb200_plan *plan, const cmb200_flux_io *io, const double *Tint, cudaStream_t s) {
    if constexpr (std::is_same<Real, double>::value && (MM == 0 || MM == 2)) {
#define CMB_FLUX_X(V_) \
    case V_: return launch_flux<Real, MM, V_>(plan, io, Tint, s);
        switch (variant) {
            CMB_FLUX_EXPERIMENTS(CMB_FLUX_X)
            default: break;
        }
#undef CMB_FLUX_X
    }
    if (variant == 0) return launch_flux<Real, MM, 0>(plan, io, Tint, s);
    if (variant == kReferenceVariant) return launch_flux<Real, MM, kReferenceVariant>(plan, io, Tint, s);
    // the default's own fall-back (shuffle reduction instead of the tensor-memory transpose) is compiled anyway: A/B
    if constexpr ((default_variant<Real, MM>() & kVarRedTmem) != 0 && std::is_same<Real, double>::value)
        if (variant == (default_variant<Real, MM>() & ~(kVarRedTmem | kVarSmallSt)))
            return launch_flux<Real, MM, default_variant<Real, MM>() & ~(kVarRedTmem | kVarSmallSt)>(plan, io, Tint, s);
    if constexpr ((default_variant<Real, MM>() & kVarPair) != 0)         // FP32: one segment per pass on 12 warps (A/B)
        if (variant == (default_variant<Real, MM>() & ~kVarPair))
            return launch_flux<Real, MM, default_variant<Real, MM>() & ~kVarPair>(plan, io, Tint, s);
    return launch_flux<Real, MM, default_variant<Real, MM>()>(plan, io, Tint, s);
}

// The flux kernels of one (precision, molecule count) as a translation unit of their own (flux_slice.cu), so that
// the build compiles them in parallel; CMB_FLUX_SLICES (set by build.py) makes this file call them instead of
// instantiating the kernels here.
#define CMB_FLUX_SLICE_DECL(NAME) int NAME(int variant, cmb200_plan *plan, const cmb200_flux_io *io, const double *Tint, cudaStream_t s);
CMB_FLUX_SLICE_DECL(launch_flux_slice_d0) CMB_FLUX_SLICE_DECL(launch_flux_slice_d1) CMB_FLUX_SLICE_DECL(launch_flux_slice_d2)
CMB_FLUX_SLICE_DECL(launch_flux_slice_d3) CMB_FLUX_SLICE_DECL(launch_flux_slice_d4)
CMB_FLUX_SLICE_DECL(launch_flux_slice_f0) CMB_FLUX_SLICE_DECL(launch_flux_slice_f1) CMB_FLUX_SLICE_DECL(launch_flux_slice_f2)
CMB_FLUX_SLICE_DECL(launch_flux_slice_f3) CMB_FLUX_SLICE_DECL(launch_flux_slice_f4)
#undef CMB_FLUX_SLICE_DECL

#ifndef CMB_FLUX_SLICE_ONLY
static int validate_flux_io(const cmb200_plan *p, const cmb200_flux_io *io) {
    CMB_CHECK_ARG(p && io, "null pointer");
    CMB_CHECK_ARG(io->n_col >= 0, "negative column count");
    if (io->n_col == 0) return CMB200_OK;
    CMB_CHECK_ARG(io->T && io->T_g, "T and T_g are required");
    if (io->tau) {
        CMB_CHECK_ARG(io->n_tau >= 1, "n_tau must be >= 1 when tau is given");
        CMB_CHECK_ARG(io->n_tau == 1 || io->tau_id, "tau_id is required when n_tau > 1");
    } else {
        if (p->M == 0) {
            set_error("no tau given and the plan has no absorption tables (cmb200_plan_set_tables)");
            return CMB200_ERR_STATE;
        }
        CMB_CHECK_ARG(io->q, "q is required for the table-fused path");
    }
    return CMB200_OK;
}

}  // namespace cmb

using namespace cmb;

extern "C" int cmb200_plan_set_tables(cmb200_plan *plan, int32_t M, const cmb200_table *tables) {
    CMB_CHECK_ARG(plan && tables, "null pointer");
    CMB_CHECK_ARG(M >= 1 && M <= kMaxMolecules, "between 1 and %d molecules supported, got %d", kMaxMolecules, M);
    CMB_CUDA(cudaSetDevice(plan->device));
    int nT_pad = 0;
    for (int m = 0; m < M; ++m) {
        int rc = check_table(&tables[m]);
        if (rc) return rc;
        nT_pad = std::max(nT_pad, (int)tables[m].n_T);
    }
    const int L = plan->L;
    const int64_t N = plan->N, pitch = (N + 15) & ~(int64_t)15;
    const size_t total = (size_t)M * L * nT_pad * pitch;
    plan->KL64.release();
    plan->KL32.release();
    if (plan->precision == CMB200_F64)
        CMB_CUDA(plan->KL64.alloc(total));
    else
        CMB_CUDA(plan->KL32.alloc(total));
    std::vector<double> axes((size_t)M * nT_pad, 1e300);  // padding rows are never nearest
    plan->nT.assign(M, 0);
    DevBuf<double> dnu;
    CMB_CUDA(dnu.upload(plan->hnu.data(), N));
    for (int m = 0; m < M; ++m) {
        const cmb200_table *t = &tables[m];
        plan->nT[m] = t->n_T;
        for (int i = 0; i < t->n_T; ++i) axes[(size_t)m * nT_pad + i] = t->T[i];
        TableOnDevice dev;
        int rc = dev.upload(t, 0);
        if (rc) return rc;
        DevBuf<int32_t> ip, inu;
        CMB_CUDA(ip.alloc(L));
        CMB_CUDA(inu.alloc(N));
        if ((rc = launch_nearest(plan->P.p, L, dev.p.p, t->n_p, false, false, 0, 0, ip.p, 0))) return rc;
        if ((rc = launch_nearest(dnu.p, N, dev.nu.p, t->n_nu, dev.nu_ascending, true, dev.nu_lo, dev.nu_hi, inu.p, 0)))
            return rc;
        dim3 grid((unsigned)((pitch + 255) / 256), (unsigned)L, (unsigned)nT_pad);
        const size_t off = (size_t)m * L * nT_pad * pitch;
        if (plan->precision == CMB200_F64)
            level_table_kernel<double><<<grid, 256>>>(dev.k.p, t->n_T, t->n_nu, ip.p, inu.p, L, nT_pad, N, pitch, plan->KL64.p + off);
        else
            level_table_kernel<float><<<grid, 256>>>(dev.k.p, t->n_T, t->n_nu, ip.p, inu.p, L, nT_pad, N, pitch, plan->KL32.p + off);
        CMB_LAUNCHED();
        CMB_CUDA(cudaDeviceSynchronize());
    }
    CMB_CUDA(plan->T_axes.upload(axes.data(), axes.size()));
    CMB_CUDA(cudaDeviceSynchronize());
    plan->M = M;
    plan->nT_pad = nT_pad;
    plan->pitch = pitch;
    plan->ks_valid = false;
    plan->partition.n_col = -1;
    return CMB200_OK;
}

extern "C" int cmb200_plan_optical_depth(cmb200_plan *plan, int64_t n_col, const double *T_tau,
                                         const double *T_g, const double *q, int q_per_column,
                                         double *tau, void *stream) {
    CMB_CHECK_ARG(plan && q && tau, "null pointer");
    CMB_CHECK_ARG(T_tau || T_g, "either T_tau or T_g is required");
    CMB_CHECK_ARG(n_col >= 0, "negative column count");
    if (plan->M == 0) {
        set_error("the plan has no absorption tables (cmb200_plan_set_tables)");
        return CMB200_ERR_STATE;
    }
    if (n_col == 0) return CMB200_OK;
    CMB_CUDA(cudaSetDevice(plan->device));
    cudaStream_t s = (cudaStream_t)stream;
    int rc = ensure_tidx(plan, n_col, T_tau, T_g, s);
    if (rc) return rc;
    dim3 grid((unsigned)((plan->N + 127) / 128), (unsigned)n_col);
    CMB_CHECK_ARG(n_col <= 65535, "at most 65535 columns per optical-depth call");
    if (plan->precision == CMB200_F64)
        plan_optical_depth_kernel<double><<<grid, 128, 0, s>>>(plan->KL64.p, plan->M, plan->L, plan->nT_pad, plan->N,
                                                              plan->pitch, plan->tidx.p, q, q_per_column ? plan->L : 0,
                                                              n_col, plan->dx.p, plan->c.g, tau);
    else
        plan_optical_depth_kernel<float><<<grid, 128, 0, s>>>(plan->KL32.p, plan->M, plan->L, plan->nT_pad, plan->N,
                                                             plan->pitch, plan->tidx.p, q, q_per_column ? plan->L : 0,
                                                             n_col, plan->dx.p, plan->c.g, tau);
    CMB_LAUNCHED();
    return CMB200_OK;
}

extern "C" int cmb200_flux_workspace_bytes(cmb200_plan *plan, int64_t n_col, int64_t *bytes) {
    CMB_CHECK_ARG(plan && bytes && n_col >= 0, "bad argument");
    *bytes = (int64_t)n_col * plan->L * 8 + (int64_t)plan->M * n_col * plan->L;
    return CMB200_OK;
}

extern "C" int cmb200_flux(cmb200_plan *plan, const cmb200_flux_io *io, void *stream) {
    int rc = validate_flux_io(plan, io);
    if (rc) return rc;
    if (io->n_col == 0) return CMB200_OK;
    CMB_CUDA(cudaSetDevice(plan->device));
    cudaStream_t s = (cudaStream_t)stream;
    const int L = plan->L;
    double *Tint = io->T_interface;
    if (!Tint) {
        size_t need = (size_t)io->n_col * L;
        if (plan->Tint.n < need) CMB_CUDA(plan->Tint.alloc(need));
        Tint = plan->Tint.p;
    }
    {
        const int warps = 4;
        const int64_t items = io->n_col * ((L + 31) / 32);
        interface_temperature_kernel<<<(unsigned)((items + warps - 1) / warps), warps * 32, 0, s>>>(
            L, io->n_col, plan->splineT.p, io->T, io->T_g, Tint);
        CMB_LAUNCHED();
    }
    const bool fused = io->tau == nullptr;
    if (fused) {
        rc = ensure_tidx(plan, io->n_col, io->T_tau, io->T_g, s);
        if (rc) return rc;
    }
    const int mm = fused ? plan->M : 0;
    const int variant = flux_variant();
#ifdef CMB_FLUX_SLICES
#define CMB_FLUX_SLICE_double(M_) launch_flux_slice_d##M_
#define CMB_FLUX_SLICE_float(M_) launch_flux_slice_f##M_
#define CMB_FLUX_CASE(REAL, M_) \
    case M_: return CMB_FLUX_SLICE_##REAL(M_)(variant, plan, io, Tint, s);
#else
#define CMB_FLUX_CASE(REAL, M_) \
    case M_: return launch_flux_variant<REAL, M_>(variant, plan, io, Tint, s);
#endif
#ifdef CMB_DEV_BUILD      /* kernel development: only the bench shapes, so that a rebuild takes seconds */
    if (plan->precision == CMB200_F64) {
        switch (mm) {
            CMB_FLUX_CASE(double, 0)
            CMB_FLUX_CASE(double, 2)
        }
    } else {
        switch (mm) {
            CMB_FLUX_CASE(float, 2)
        }
    }
#else
    if (plan->precision == CMB200_F64) {
        switch (mm) {
            CMB_FLUX_CASE(double, 0)
            CMB_FLUX_CASE(double, 1)
            CMB_FLUX_CASE(double, 2)
            CMB_FLUX_CASE(double, 3)
            CMB_FLUX_CASE(double, 4)
        }
    } else {
        switch (mm) {
            CMB_FLUX_CASE(float, 0)
            CMB_FLUX_CASE(float, 1)
            CMB_FLUX_CASE(float, 2)
            CMB_FLUX_CASE(float, 3)
            CMB_FLUX_CASE(float, 4)
        }
    }
#endif
#undef CMB_FLUX_CASE
    set_error("unsupported molecule count %d", mm);
    return CMB200_ERR_LIMIT;
}

// Host-pointer entry point: columns are streamed through the device in chunks, with the copy of
// the next chunk's inputs and of the previous chunk's outputs overlapping the kernels of the
// current one (three streams, double-buffered staging kept on the plan between calls).  The
// kernels of successive chunks run in order on one stream because they share the plan's scratch.
struct cmb200_host_pipe {
    cudaStream_t h2d = nullptr, comp = nullptr, d2h = nullptr;
    cudaEvent_t in_ready[2] = {nullptr, nullptr}, computed[2] = {nullptr, nullptr}, out_done[2] = {nullptr, nullptr};
    cmb::DevBuf<double> T[2], Tg[2], alb[2], Ttau[2], q[2], up[2], down[2], net[2], heat[2], Tint[2], tau;
    cmb::DevBuf<int32_t> tid[2];
    ~cmb200_host_pipe() {
        for (int b = 0; b < 2; ++b) {
            if (in_ready[b]) cudaEventDestroy(in_ready[b]);
            if (computed[b]) cudaEventDestroy(computed[b]);
            if (out_done[b]) cudaEventDestroy(out_done[b]);
        }
        if (h2d) cudaStreamDestroy(h2d);
        if (comp) cudaStreamDestroy(comp);
        if (d2h) cudaStreamDestroy(d2h);
    }
};
void cmb::destroy_host_pipe(cmb200_host_pipe *p) { delete p; }

template <typename T>
static cudaError_t grow(cmb::DevBuf<T> &b, size_t n) {
    return b.n >= n ? cudaSuccess : b.alloc(n);
}

extern "C" int cmb200_flux_host(cmb200_plan *plan, const cmb200_flux_io *io) {
    int rc = validate_flux_io(plan, io);
    if (rc) return rc;
    if (io->n_col == 0) return CMB200_OK;
    CMB_CUDA(cudaSetDevice(plan->device));
    const int64_t n = io->n_col;
    const int L = plan->L, B = plan->B, M = plan->M;
    if (io->tau && io->tau_id)
        for (int64_t i = 0; i < n; ++i)
            CMB_CHECK_ARG(io->tau_id[i] >= 0 && io->tau_id[i] < io->n_tau, "tau_id[%lld] out of range", (long long)i);
    if (!plan->host_pipe) {
        cmb200_host_pipe *hp = new cmb200_host_pipe();
        plan->host_pipe = hp;
        CMB_CUDA(cudaStreamCreateWithFlags(&hp->h2d, cudaStreamNonBlocking));
        CMB_CUDA(cudaStreamCreateWithFlags(&hp->comp, cudaStreamNonBlocking));
        CMB_CUDA(cudaStreamCreateWithFlags(&hp->d2h, cudaStreamNonBlocking));
        for (int b = 0; b < 2; ++b) {
            CMB_CUDA(cudaEventCreateWithFlags(&hp->in_ready[b], cudaEventDisableTiming));
            CMB_CUDA(cudaEventCreateWithFlags(&hp->computed[b], cudaEventDisableTiming));
            CMB_CUDA(cudaEventCreateWithFlags(&hp->out_done[b], cudaEventDisableTiming));
        }
    }
    cmb200_host_pipe &hp = *plan->host_pipe;
    // chunk: a multiple of the resident CTA count, ~16 chunks for large batches, one for small ones
    const int64_t quantum = (int64_t)sm_count(plan->device) * 8;
    int64_t chunk = n;
    if (n > 4 * quantum) chunk = std::min<int64_t>(std::max<int64_t>(((n / 16 + quantum - 1) / quantum) * quantum, 4 * quantum), 64 * quantum);
    const bool fused = io->tau == nullptr, q_cols = fused && io->q_per_column;
    const int n_buf = chunk < n ? 2 : 1;
    for (int b = 0; b < n_buf; ++b) {
        CMB_CUDA(grow(hp.T[b], (size_t)chunk * (L - 1)));
        CMB_CUDA(grow(hp.Tg[b], (size_t)chunk));
        if (io->albedo) CMB_CUDA(grow(hp.alb[b], (size_t)chunk));
        if (io->tau && io->tau_id) CMB_CUDA(grow(hp.tid[b], (size_t)chunk));
        if (fused && io->T_tau) CMB_CUDA(grow(hp.Ttau[b], (size_t)chunk * L));
        if (q_cols) CMB_CUDA(grow(hp.q[b], (size_t)M * chunk * L));
        if (io->up) CMB_CUDA(grow(hp.up[b], (size_t)chunk * L * B));
        if (io->down) CMB_CUDA(grow(hp.down[b], (size_t)chunk * L * B));
        if (io->net) CMB_CUDA(grow(hp.net[b], (size_t)chunk * L));
        if (io->heating) CMB_CUDA(grow(hp.heat[b], (size_t)chunk * (L - 1)));
        if (io->T_interface) CMB_CUDA(grow(hp.Tint[b], (size_t)chunk * L));
    }
    // column-independent inputs: once, ahead of the first chunk
    if (io->tau) {
        CMB_CUDA(grow(hp.tau, (size_t)io->n_tau * L * plan->N));
        CMB_CUDA(cudaMemcpyAsync(hp.tau.p, io->tau, (size_t)io->n_tau * L * plan->N * sizeof(double), cudaMemcpyHostToDevice, hp.h2d));
    } else if (!q_cols) {
        CMB_CUDA(grow(hp.q[0], (size_t)M * L));
        CMB_CUDA(cudaMemcpyAsync(hp.q[0].p, io->q, (size_t)M * L * sizeof(double), cudaMemcpyHostToDevice, hp.h2d));
    }
    auto h2d = [&](double *dst, const double *src, size_t count) {
        return cudaMemcpyAsync(dst, src, count * sizeof(double), cudaMemcpyHostToDevice, hp.h2d);
    };
    auto d2h = [&](double *dst, const double *src, size_t count) {
        return cudaMemcpyAsync(dst, src, count * sizeof(double), cudaMemcpyDeviceToHost, hp.d2h);
    };
    // Host order is software-pipelined too: chunk k's inputs and kernels are queued before chunk
    // k-1's outputs are fetched, so a copy that blocks the host (pageable memory) does not starve
    // the device.
    const int64_t n_chunks = (n + chunk - 1) / chunk;
    auto fetch_outputs = [&](int64_t k) -> int {
        const int b = (int)(k % n_buf);
        const int64_t c0 = k * chunk, m = std::min(chunk, n - c0);
        // ---- outputs ------------------------------------------------------------------------------------
        CMB_CUDA(cudaStreamWaitEvent(hp.d2h, hp.computed[b], 0));
        if (io->up) CMB_CUDA(d2h(io->up + c0 * L * B, hp.up[b].p, (size_t)m * L * B));
        if (io->down) CMB_CUDA(d2h(io->down + c0 * L * B, hp.down[b].p, (size_t)m * L * B));
        if (io->net) CMB_CUDA(d2h(io->net + c0 * L, hp.net[b].p, (size_t)m * L));
        if (io->heating) CMB_CUDA(d2h(io->heating + c0 * (L - 1), hp.heat[b].p, (size_t)m * (L - 1)));
        if (io->T_interface) CMB_CUDA(d2h(io->T_interface + c0 * L, hp.Tint[b].p, (size_t)m * L));
        CMB_CUDA(cudaEventRecord(hp.out_done[b], hp.d2h));
        return CMB200_OK;
    };
    for (int64_t k = 0; k < n_chunks; ++k) {
        const int64_t c0 = k * chunk;
        const int b = (int)(k % n_buf);
        const int64_t m = std::min(chunk, n - c0);
        // ---- inputs of this chunk (the buffer's previous kernels must be done with them) ----------
        if (k >= n_buf) CMB_CUDA(cudaStreamWaitEvent(hp.h2d, hp.computed[b], 0));
        CMB_CUDA(h2d(hp.T[b].p, io->T + c0 * (L - 1), (size_t)m * (L - 1)));
        CMB_CUDA(h2d(hp.Tg[b].p, io->T_g + c0, (size_t)m));
        if (io->albedo) CMB_CUDA(h2d(hp.alb[b].p, io->albedo + c0, (size_t)m));
        if (io->tau && io->tau_id)
            CMB_CUDA(cudaMemcpyAsync(hp.tid[b].p, io->tau_id + c0, (size_t)m * sizeof(int32_t), cudaMemcpyHostToDevice, hp.h2d));
        if (fused && io->T_tau) CMB_CUDA(h2d(hp.Ttau[b].p, io->T_tau + c0 * L, (size_t)m * L));
        if (q_cols)
            for (int mol = 0; mol < M; ++mol)
                CMB_CUDA(h2d(hp.q[b].p + (size_t)mol * m * L, io->q + ((size_t)mol * n + c0) * L, (size_t)m * L));
        CMB_CUDA(cudaEventRecord(hp.in_ready[b], hp.h2d));
        // ---- kernels ------------------------------------------------------------------------------------
        cmb200_flux_io d = *io;
        d.n_col = m;
        d.T = hp.T[b].p;
        d.T_g = hp.Tg[b].p;
        d.albedo = io->albedo ? hp.alb[b].p : nullptr;
        d.tau = io->tau ? hp.tau.p : nullptr;
        d.tau_id = io->tau && io->tau_id ? hp.tid[b].p : nullptr;
        d.T_tau = fused && io->T_tau ? hp.Ttau[b].p : nullptr;
        d.q = fused ? (q_cols ? hp.q[b].p : hp.q[0].p) : nullptr;
        d.up = io->up ? hp.up[b].p : nullptr;
        d.down = io->down ? hp.down[b].p : nullptr;
        d.net = io->net ? hp.net[b].p : nullptr;
        d.heating = io->heating ? hp.heat[b].p : nullptr;
        d.T_interface = io->T_interface ? hp.Tint[b].p : nullptr;
        CMB_CUDA(cudaStreamWaitEvent(hp.comp, hp.in_ready[b], 0));
        if (k >= n_buf) CMB_CUDA(cudaStreamWaitEvent(hp.comp, hp.out_done[b], 0));   // outputs of chunk k-2 copied out
        // bands outside the plan's band / segment range are not written by the kernels: they read as zero
        if (d.up) CMB_CUDA(cudaMemsetAsync(d.up, 0, (size_t)m * L * B * sizeof(double), hp.comp));
        if (d.down) CMB_CUDA(cudaMemsetAsync(d.down, 0, (size_t)m * L * B * sizeof(double), hp.comp));
        rc = cmb200_flux(plan, &d, hp.comp);
        if (rc) {
            cudaDeviceSynchronize();
            return rc;
        }
        CMB_CUDA(cudaEventRecord(hp.computed[b], hp.comp));
        if (k >= 1 && (rc = fetch_outputs(k - 1))) return rc;
    }
    if ((rc = fetch_outputs(n_chunks - 1))) return rc;
    CMB_CUDA(cudaStreamSynchronize(hp.h2d));
    CMB_CUDA(cudaStreamSynchronize(hp.comp));
    CMB_CUDA(cudaStreamSynchronize(hp.d2h));
    return CMB200_OK;
}

extern "C" int cmb200_test_exp_neg_host(const double *x, int64_t n, double *out) {
    CMB_CHECK_ARG(x && out && n >= 0, "bad argument");
    if (n == 0) return CMB200_OK;
    DevBuf<double> dx, dout;
    CMB_CUDA(dx.upload(x, n));
    CMB_CUDA(dout.alloc(n));
    exp_neg_test_kernel<<<(unsigned)((n + 255) / 256), 256>>>(dx.p, n, dout.p);
    CMB_LAUNCHED();
    CMB_CUDA(cudaMemcpy(out, dout.p, n * sizeof(double), cudaMemcpyDeviceToHost));
    return CMB200_OK;
}

extern "C" int cmb200_test_exp_neg_table_host(const double *x, int64_t n, double *out) {
    CMB_CHECK_ARG(x && out && n >= 0, "bad argument");
    if (n == 0) return CMB200_OK;
    DevBuf<double> dx, dout, dtab;
    CMB_CUDA(dx.upload(x, n));
    CMB_CUDA(dout.alloc(n));
    std::vector<double> tab(kXlN);
    for (int j = 0; j < kXlN; ++j) tab[j] = xl_table_entry((double)exp2l((long double)j / kXlN), j);
    CMB_CUDA(dtab.upload(tab.data(), tab.size()));
    constexpr int V = kVarExpLds | kVarSwDirect;
    const int smem = (int)(xl_window(V) + kXlBytes);
    CMB_CUDA(cudaFuncSetAttribute(exp_neg_table_test_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    exp_neg_table_test_kernel<<<(unsigned)std::min<int64_t>(1024, (n + 256 * kU - 1) / (256 * kU)), 256, smem>>>(dx.p, n, dout.p, dtab.p);
    CMB_LAUNCHED();
    CMB_CUDA(cudaMemcpy(out, dout.p, n * sizeof(double), cudaMemcpyDeviceToHost));
    return CMB200_OK;
}
#endif  // CMB_FLUX_SLICE_ONLY
#ifdef CMB_FLUX_SLICE_ONLY
}  // namespace cmb
#endif
