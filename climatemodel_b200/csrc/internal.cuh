// Library-internal declarations: the plan object and the launchers shared between .cu files.
#pragma once
#include <vector>

#include "common.cuh"

namespace cmb {

constexpr int kMaxMolecules = 4;
constexpr int kChains = 1;  // wavenumber bins per lane and pass of the flux kernel (R)

int check_table(const cmb200_table *t);
int launch_nearest(const double *d_values, int64_t n, const double *d_axis, int n_axis,
                   bool axis_ascending, bool crop, double lo, double hi, int32_t *d_out,
                   cudaStream_t s);

// Work decomposition of one flux evaluation, cached per (n_col) on the plan.
struct Partition {
    int64_t n_col = -1;
    int warps = 0;          // W: warps per CTA
    int ctas_per_col = 1;   // C: CTAs sharing a column
    int parts = 0;          // C * W contiguous, cost-balanced runs of segments per column
    int n_slots = 0;        // partial-sum slots per column: one per (part, band) run
    int cost_model = -1;    // which kernel family the segment costs were fitted to (flux.cu:pass_costs)
    DevBuf<int32_t> part_seg0, part_slot0, band_slot0;
    DevBuf<double> slot_delta;
};

}  // namespace cmb

struct cmb200_host_pipe;
namespace cmb {
void destroy_host_pipe(cmb200_host_pipe *p);
}

struct cmb200_plan {
    int device = 0;
    int precision = CMB200_F64;
    cmb200_constants c{};
    int L = 0, B = 0;
    int64_t N = 0;
    // host copies
    std::vector<double> hP, hnu, hcentre, hdelta;
    std::vector<int32_t> hstart, hlen;
    std::vector<uint8_t> hsw;
    int band_lo = 0, band_hi = 0;
    int64_t seg_lo = 0, seg_hi = (int64_t)1 << 62, n_seg_total = 0;
    // star (real_gas.py:246)
    double T_star = 5778.0, R_star = 6.96340e8, dist = 1.495978707e11;
    // device: grid
    cmb::DevBuf<double> P, dx, hdP, geo, splineT;
    // device: bands and chains
    cmb::DevBuf<int32_t> band_start, band_len;
    cmb::DevBuf<double> band_centre, band_delta, band_sumw, wchain64;
    cmb::DevBuf<float> wchain32;
    cmb::DevBuf<uint8_t> band_sw;
    cmb::DevBuf<int64_t> chain_off;
    int64_t n_chains = 0;
    // device: segments (runs of 32*kChains bins inside one band) for bands [band_lo, band_hi)
    cmb::DevBuf<int32_t> seg_band, seg_c0, seg_bin;   // seg_bin[n_seg][32]: wavenumber bin of each lane's chain
    cmb::DevBuf<double> seg_w64;                      // seg_w[n_seg][32]: its nu-weight (0 for idle lanes)
    cmb::DevBuf<float> seg_w32;
    std::vector<int32_t> hseg_band, hseg_c0;
    std::vector<double> hwchain;      // nu-weights of all chains, band after band
    std::vector<int64_t> hchain_off;
    int n_seg = 0;
    // device: level table KL[M][L][nT_pad][pitch]
    int M = 0, nT_pad = 0;
    int64_t pitch = 0;
    std::vector<int> nT;
    cmb::DevBuf<double> KL64, T_axes;  // T_axes[M][nT_pad]
    cmb::DevBuf<float> KL32;
    // the same table in segment space for the flux kernel: KS[M][L][nT_pad][n_seg * 32], column
    // s * 32 + lane = bin of lane `lane` of segment s, so that every row a warp loads is one aligned
    // 256-byte (FP32: 128-byte) run; built on first use, dropped when segments or tables change
    cmb::DevBuf<double> KS64;
    cmb::DevBuf<float> KS32;
    int64_t spitch = 0;
    bool ks_valid = false;
    cmb::Partition partition;
    // per-call scratch kept between calls (grown on demand)
    cmb::DevBuf<double> Tint, scratch, coefbuf;
    cmb::DevBuf<uint8_t> tidx;
    cmb::DevBuf<unsigned long long> task_counter;
    cmb::DevBuf<double> exp_tab;        // 2^(j/512), the flux kernel's shared-memory exp table (kVarExpLds)
    cmb200_host_pipe *host_pipe = nullptr;   // streams and staging buffers of cmb200_flux_host
    ~cmb200_plan() {
        if (host_pipe) cmb::destroy_host_pipe(host_pipe);
    }
};

namespace cmb {
int build_segments(cmb200_plan *p);
int launch_convective_adjust(int32_t n_levels, int64_t n_col, const double *p_centre, const double *p_power,
                             double conv_thresh, double *T, const uint8_t *active, uint8_t *convective_levels,
                             int32_t *status, const uint8_t *frozen, void *stream);
int spline_operator_host(const double *P, int L, std::vector<double> &At);
}  // namespace cmb
