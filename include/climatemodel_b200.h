/*
 * climatemodel_b200 -- C ABI of the B200-native real-gas radiative-transfer hot path.
 *
 * This is the drop-in boundary.  The reference (jduffield65/ClimateModel)
 * is pure Python, so the "FFI" a maintainer binds is ctypes; INTEGRATION.md shows the stub.
 * Every entry point names the reference function it replaces (file:line under the reference
 * tree).  All signatures are plain C: pointers, sizes and POD structs -- no torch / C++ types.
 *
 * Conventions
 *   - interface index 0 = top of atmosphere, L-1 = surface (real_gas.py:281); pressures ascending
 *   - cell centres p_k = (P_k + P_{k+1})/2 carry T_k, k = 0..L-2         (real_gas.py:282-284)
 *   - units as Model/constants.py: Pa, K, cm^-1, m^2/kg, kg/kg, W m^-2 / cm^-1, W m^-2, K/s
 *   - all floating-point data crossing this boundary is FP64 (the reference's dtype); a plan
 *     created with CMB200_F32 computes the wavenumber-resolved stage in FP32 internally
 *   - functions ending in _host take HOST pointers and do their own host<->device copies;
 *     the others take DEVICE pointers (on the plan's device) and a cudaStream_t passed as void*
 *   - every function returns CMB200_OK (0) or a negative error code; cmb200_last_error() gives
 *     the message of the calling thread's last failure.  There is no CPU fallback: without a
 *     CUDA device every compute entry point fails with CMB200_ERR_CUDA.
 */
#ifndef CLIMATEMODEL_B200_H
#define CLIMATEMODEL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CMB200_OK            0
#define CMB200_ERR_ARG      -1   /* bad argument (ValueError on the Python side)            */
#define CMB200_ERR_CUDA     -2   /* CUDA runtime error / no device                          */
#define CMB200_ERR_STATE    -3   /* plan not ready for this call (e.g. tables not set)      */
#define CMB200_ERR_LIMIT    -4   /* problem does not fit (shared memory / index range)      */

#define CMB200_F64 0
#define CMB200_F32 1

typedef struct cmb200_plan cmb200_plan;     /* opaque: grids, bands, tables resident in HBM */

/* Physical constants the path uses -- same names/units as Model/constants.py:3-16. */
typedef struct {
    double g;                /* gravitational_acceleration           m s^-2      */
    double c_p_dry;          /* heat capacity of dry air             J kg^-1 K^-1 */
    double h_planck;         /* J s                                               */
    double speed_of_light;   /* m s^-1                                            */
    double k_boltzmann;      /* J K^-1                                            */
} cmb200_constants;

/* One absorption-coefficient lookup table, the dict saved by hitran.make_table
 * (hitran.py:315-356; loaded at real_gas.py:77): axes p[n_p], T[n_T], nu[n_nu] and
 * k[n_p][n_T][n_nu] in C order (nu fastest).  HOST pointers. */
typedef struct {
    int32_t n_p, n_T, n_nu;
    const double *p, *T, *nu, *k;
} cmb200_table;

/* ---- library ---------------------------------------------------------------------------- */
const char *cmb200_last_error(void);
int cmb200_version(void);
int cmb200_device_count(int *count);

/* ---- A1  Planck function ---------------------------------------------------------------- */
/* B_wavenumber(nu, T) (real_gas.py:30-42, via B_freq :14-27): out[i_T][i_nu], broadcast of
 * nu[n_nu] against T[n_T].  No overflow guard, like the reference. */
int cmb200_planck_wavenumber_host(const cmb200_constants *c, const double *nu, int64_t n_nu,
                                  const double *T, int64_t n_T, double *out);

/* ---- A2  nearest-neighbour table gather --------------------------------------------------- */
/* get_absorption_coef / load_absorption_coef (real_gas.py:45-83): k_out[n_lev][n_nu];
 * argmin |x - axis| with first-index tie-break on p, T and nu; when zero_outside != 0 bins
 * outside [table nu.min(), nu.max()] give 0 (load_absorption_coef), else they clamp to the
 * nearest table bin (get_absorption_coef). */
int cmb200_absorption_coef_host(const cmb200_table *table, const double *p, const double *T,
                                int64_t n_lev, const double *nu, int64_t n_nu, int zero_outside,
                                double *k_out);

/* ---- A3  optical depth --------------------------------------------------------------------- */
/* optical_depth (real_gas.py:86-127) with the q_funcs already evaluated by the caller:
 * p_sorted[n_lev] ascending, T[n_lev], q[M][n_lev]; tau_out[n_lev][n_nu].  Bit-exact with the
 * reference's trapezoid (same operation order, no FMA contraction). */
int cmb200_optical_depth_host(const cmb200_constants *c, int32_t M, const cmb200_table *tables,
                              const double *p_sorted, const double *T, int64_t n_lev,
                              const double *nu, int64_t n_nu, const double *q, double *tau_out);

/* ---- A4 / A5  band-averaged transmission --------------------------------------------------- */
/* transmission (real_gas.py:130-154).  The exact-equality lookups of p1, p2, nu_band in p_all,
 * nu_all (:143-147) are done by the caller, who passes row indices rows1[n1], rows2[n2] into
 * tau[n_lev][n_nu] and the bin run [bin0, bin0+n_bins) with its wavenumbers nu_band[n_bins].
 * out[n1][n2] = trapz_nu exp(tau[rows1] - tau[rows2]) / delta. */
int cmb200_transmission_host(const double *tau, int64_t n_lev, int64_t n_nu,
                             const int32_t *rows1, int32_t n1, const int32_t *rows2, int32_t n2,
                             const int32_t *bins, const double *nu_band, int32_t n_bins,
                             double delta, double *out);
/* dtransmission_dq (real_gas.py:157-184): as above times (p1 - p2) k_nu / g inside the
 * integral; scale[n1][n2] holds the caller's (p1 - p2) factor per output entry. */
int cmb200_dtransmission_dq_host(const double *tau, int64_t n_lev, int64_t n_nu,
                                 const int32_t *rows1, int32_t n1, const int32_t *rows2, int32_t n2,
                                 const int32_t *bins, const double *nu_band, int32_t n_bins,
                                 double delta, const double *k_band, const double *scale,
                                 double g, double *out);

/* ---- f3  line-by-line table builder --------------------------------------------------------- */
/* get_absorption_coefficient (real_gas_data/hitran.py:223-247): Lorentz lines summed onto the
 * grid nu[n_nu] (evenly spaced) for the levels (p[i], T[i]), i < n_p.  Line l has centre
 * line_nu (cm^-1), strength line_sw at T_reference (already in m^2/kg units, s_conversion :122),
 * air-broadened half-width line_gamma_air at (p_reference, T_reference) and temperature exponent
 * line_n_air; each line is kept within n_line_widths of its widest half-width
 * (wavenumbers_near_line :182-199).  k[n_p][n_nu] in m^2/kg. */
int cmb200_line_absorption_host(const cmb200_constants *c, double p_reference, double T_reference,
                                int64_t n_p, const double *p, const double *T, int64_t n_nu,
                                const double *nu, int64_t n_lines, const double *line_nu,
                                const double *line_sw, const double *line_gamma_air,
                                const double *line_n_air, double n_line_widths, double *k);

/* ---- plan: grids, bands and tables resident on one device ---------------------------------- */
typedef struct cmb200_stepper cmb200_stepper;

/* Replaces the state RealGas.__init__ builds (real_gas.py:275-292): interface pressures
 * P[L] ascending, wavenumber grid nu[N] (get_wavenumber_array :300), bands
 * (get_wavenumber_bands :336-388) as bin runs band_start/len[B] with centre, delta, sw flag. */
int cmb200_plan_create(cmb200_plan **plan, int device, int precision, const cmb200_constants *c,
                       int32_t L, const double *P, int64_t N, const double *nu, int32_t B,
                       const int32_t *band_start, const int32_t *band_len,
                       const double *band_centre, const double *band_delta,
                       const uint8_t *band_sw);
int cmb200_plan_destroy(cmb200_plan *plan);
/* star as self.star (real_gas.py:246): incoming short-wave uses B(nu, T_star) (R_star/dist)^2 */
int cmb200_plan_set_star(cmb200_plan *plan, double T_star, double R_star, double dist);
/* Gathers each table onto the plan's (P, nu) grid -- the integer nearest-neighbour work of
 * real_gas.py:57-62, :76-82 done once -- into the level table KL[M][L][nT][N] kept in HBM. */
int cmb200_plan_set_tables(cmb200_plan *plan, int32_t M, const cmb200_table *tables);
/* Restrict the plan to bands [band_lo, band_hi): wavenumber-band sharding across GPUs
 * (SURVEY.md 8e).  Outputs net/heating then hold this shard's partial band sums. */
int cmb200_plan_set_band_range(cmb200_plan *plan, int32_t band_lo, int32_t band_hi);
/* Finer shard unit: a segment is a run of bins_per_segment consecutive bins inside one band
 * (bands are enumerated in order, each split into ceil(len/bins_per_segment) segments).
 * Restricting a plan to segments [seg_lo, seg_hi) lets ranks balance work when one band holds
 * a large share of the bins; up/down/net are then partial sums to be added across ranks. */
int cmb200_plan_segments(const cmb200_plan *plan, int32_t *bins_per_segment, int64_t *n_segments);
int cmb200_plan_set_segment_range(cmb200_plan *plan, int64_t seg_lo, int64_t seg_hi);
int cmb200_plan_info(const cmb200_plan *plan, int32_t *L, int64_t *N, int32_t *B, int32_t *M,
                     int32_t *nT, int32_t *precision, int64_t *n_chains);

/* ---- A3 on the plan grid (device) ---------------------------------------------------------- */
/* tau[n_col][L][N] from the level table: T_tau[n_col][L] (NULL: isothermal at T_g[col]),
 * q[M][n_col|1][L].  Same arithmetic as cmb200_optical_depth_host. */
int cmb200_plan_optical_depth(cmb200_plan *plan, int64_t n_col, const double *T_tau,
                              const double *T_g, const double *q, int q_per_column,
                              double *tau, void *stream);

/* cmb200_plan_optical_depth without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, same arithmetic, same results as
 * cmb200_plan_optical_depth; the one scratch buffer of that call (the table-temperature row index per
 * molecule, column and level) lives in a file-scope device array of 1 MiB that every GPU loading the
 * library holds until exit, so plan M x n_col x L must be <= 1,048,576 (CMB200_ERR_ARG with a message
 * otherwise; n_col <= 65535 as for the regular call).  Calls no cudaMalloc / cudaFree / allocating
 * library routine; the first call on a device looks the array's address up (cudaGetSymbolAddress).
 * All calls on a device share the array, which keeps the last call's contents (every entry a call reads
 * is rewritten by that call first, so nothing needs clearing): calls whose kernels overlap on the GPU
 * race -- keep them on one stream or order them with events. */
int cmb200_plan_optical_depth_noalloc(cmb200_plan *plan, int64_t n_col, const double *T_tau,
                                      const double *T_g, const double *q, int q_per_column,
                                      double *tau, void *stream);

/* ---- A6-A9  fluxes, net flux, heating rate -------------------------------------------------- */
typedef struct {
    int64_t n_col;
    const double *T;        /* [n_col][L-1]  cell-centre temperatures (self.T)                 */
    const double *T_g;      /* [n_col]       surface temperature (self.T_g)                    */
    const double *albedo;   /* [n_col] or NULL (then albedo_scalar)                             */
    double albedo_scalar;
    /* optical-depth source A: resident tau (self.tau_interface, real_gas.py:291)              */
    const double *tau;      /* [n_tau][L][N] or NULL                                            */
    const int32_t *tau_id;  /* [n_col] index into tau, or NULL (all columns use tau[0])         */
    int32_t n_tau;
    /* optical-depth source B: fused from the plan's level table (stage 1 inside the kernel)    */
    const double *T_tau;    /* [n_col][L] lookup temperatures, or NULL (isothermal at T_g)      */
    const double *q;        /* [M][n_col|1][L] specific humidity on the interfaces              */
    int32_t q_per_column;
    /* outputs -- any may be NULL                                                                */
    double *up;             /* [n_col][L][B]   up_flux   (real_gas.py:638-663)                  */
    double *down;           /* [n_col][L][B]   down_flux                                        */
    double *net;            /* [n_col][L]      net flux  (real_gas.py:712-713)                  */
    double *heating;        /* [n_col][L-1]    g/c_p dnet/dp (base.py:166-168)                  */
    double *T_interface;    /* [n_col][L]      spline temperatures (real_gas.py:634-636)        */
} cmb200_flux_io;

/* RealGas.get_flux + net flux + heating rate for n_col independent columns; DEVICE pointers. */
int cmb200_flux(cmb200_plan *plan, const cmb200_flux_io *io, void *stream);
/* Same with HOST pointers: copies inputs H2D, runs, copies the requested outputs D2H. */
int cmb200_flux_host(cmb200_plan *plan, const cmb200_flux_io *io);
/* Bytes of device scratch cmb200_flux would allocate for n_col columns (0 if none). */
int cmb200_flux_workspace_bytes(cmb200_plan *plan, int64_t n_col, int64_t *bytes);
/* Number of kernel launches issued by this library since load (bench.py's gpu_launches). */
int64_t cmb200_launch_count(void);

/* ---- f1  adaptive time step, batched over columns on the device ----------------------------- */
/* Per-column state of Atmosphere.time_step_info (base.py:137-141: DeltaT, MaxTend, MaxTendInd,
 * nSameMaxInd, nSameMaxInd2, RemoveInd, dt) and the previous net flux, for n_col columns on the
 * plan's device.  temp_change / delta_temp_change as in Atmosphere.__init__ (base.py:113). */
int cmb200_stepper_create(cmb200_stepper **stepper, cmb200_plan *plan, int64_t n_col,
                          double temp_change, double delta_temp_change);
int cmb200_stepper_reset(cmb200_stepper *stepper);
int cmb200_stepper_destroy(cmb200_stepper *stepper);
/* Level numbering of the caller the stepper stands in for.  The reference works on its arrays in
 * the order the model stores them: RealGas ascending in pressure (index 0 = top of atmosphere,
 * real_gas.py:281), the grey model surface first (grey.py:207-249).  update_temp gates layer i on
 * |net_flux[i]| (base.py:171), np.argmax breaks ties by first index (base.py:209) and MaxTendInd /
 * RemoveInd hold indices in that order.  top_first = 1 (default): RealGas numbering;
 * top_first = 0: surface-first numbering.  The device arrays stay top first either way. */
int cmb200_stepper_set_level_order(cmb200_stepper *stepper, int top_first);
/* Atmosphere.update_temp with convective_adjust=False (base.py:166-196) and update_time_step
 * (base.py:197-246) for every column at once.  DEVICE pointers: net[n_col][L] from cmb200_flux,
 * T[n_col][L-1] updated in place, t[n_col] (seconds) advanced by each column's own dt,
 * delta_net_flux[n_col] (may be NULL) = percentile of |net - previous net|, 1e6 on a column's
 * first step or when changing_tau is set.  Runs on `stream`, no synchronisation. */
int cmb200_stepper_update(cmb200_stepper *stepper, const double *net, double *T, double *t,
                          double *delta_net_flux, double net_flux_thresh,
                          double net_flux_percentile, int changing_tau, void *stream);
/* ---- f2  dry convective adjustment, batched over columns on the device ----------------------- */
/* convective_adjustment_single (convective_adjustment.py:7-135) for every column.  DEVICE
 * pointers: p_centre[n_levels] cell-centre pressures, SURFACE FIRST (descending, the reference's
 * order), p_power[n_levels] = (p_centre / p_reference)^(R lapse / g) evaluated by the caller;
 * T[n_col][n_levels] in this library's order (top of atmosphere first), adjusted in place.
 * Optional: active[n_col][n_levels] (levels to report on, surface first; NULL = all),
 * convective_levels[n_col][n_levels] out = active levels moved by more than conv_thresh
 * (base.py:189-192), status[n_col] out (0 ok, 1 an unstable run reached the top level -- where
 * the reference raises IndexError -- and was left alone, 2 iteration cap). */
int cmb200_convective_adjust(int32_t n_levels, int64_t n_col, const double *p_centre,
                             const double *p_power, double conv_thresh, double *T,
                             const uint8_t *active, uint8_t *convective_levels, int32_t *status,
                             void *stream);
/* Host arrays p_centre / p_power as above, copied to the stepper; then
 * cmb200_stepper_update_convective = update_temp with convective_adjust=True (base.py:166-196):
 * the step, times conv_t_multiplier when the level that sets it was convecting, then the
 * adjustment. */
int cmb200_stepper_set_convection(cmb200_stepper *stepper, const double *p_centre, const double *p_power);
int cmb200_stepper_update_convective(cmb200_stepper *stepper, const double *net, double *T, double *t,
                                     double *delta_net_flux, double net_flux_thresh,
                                     double net_flux_percentile, int changing_tau, double conv_thresh,
                                     double conv_t_multiplier, void *stream);
/* frozen[n_col] (DEVICE, read at every update; NULL clears): columns with a non-zero entry are
 * skipped by cmb200_stepper_update / _update_convective -- T, t and the step state stay as they
 * are.  This is how a batch follows Atmosphere.evolve_to_equilibrium (base.py:266-335), where a
 * column stops stepping once it has reached equilibrium. */
int cmb200_stepper_set_frozen(cmb200_stepper *stepper, const uint8_t *frozen);
/* Copies of the per-column state to HOST arrays of n_col entries (any may be NULL);
 * max_tend_ind is numbered in the order chosen by cmb200_stepper_set_level_order. */
int cmb200_stepper_state(const cmb200_stepper *stepper, double *delta_T, double *dt,
                         int32_t *max_tend_ind, int32_t *n_removed);

/* ---- grey two-stream fluxes (the reference's other radiation scheme) ------------------------- */
/* GreyGas.get_lw_flux (grey.py:251-275) and get_sw_flux (:277-294) for ny columns at once.
 * HOST arrays in the reference's layout [level][column], level 0 at the surface:
 * T, dtau [nz-1][ny]; toa_up_lw[ny] = (1 - albedo_mod) solar_latitude_factor F / 4 (:265);
 * up_sw0 / down_sw0 [ny] = the short-wave streams without atmosphere (:287-290, may be NULL);
 * tau_sw [nz][ny] short-wave optical depth on the interfaces (NULL: transparent or
 * isothermal=True).  Outputs [nz][ny]: up_lw, down_lw, and optionally up_sw, down_sw,
 * net = up_lw - down_lw + up_sw - down_sw (:341). */
int cmb200_grey_flux_host(int32_t nz, int64_t ny, double sigma, const double *T, const double *dtau,
                          const double *toa_up_lw, const double *up_sw0, const double *down_sw0,
                          const double *tau_sw, double *up_lw, double *down_lw, double *up_sw,
                          double *down_sw, double *net);

/* ---- measurement helpers -------------------------------------------------------------------- */
/* Issue-rate microbenchmarks on the plan-independent current device: dependent-chain-free
 * DFMA and FFMA loops (TFLOP/s, 2 flop per FMA) and a device copy (GB/s, read+write). */
int cmb200_measure_peaks(int device, double *fp64_tflops, double *fp32_tflops, double *copy_gbs);
/* Denominators measured the way the flux kernel runs (one CTA of 8 warps per SM): DFMA / FFMA loops
 * whose multiplier and addend sit in registers (TFLOP/s), and the L2 -> SM read bandwidth -- a 64 MB
 * L2-resident buffer read repeatedly with coalesced 16-byte loads at full occupancy (GB/s), the ceiling of
 * the flux kernel's table stream.  Any pointer may be NULL. */
int cmb200_measure_peaks2(int device, double *fp64_reg_tflops, double *fp32_reg_tflops, double *l2_stream_gbs);

/* Test hooks: the flux kernel's own exp(-x) routines evaluated on n values (HOST pointers) -- the degree-11
 * polynomial of the plain and long-wave passes (<= 2 ulp), and the shared-memory table routine of the FP64 default's
 * short-wave passes (2^(j/512) table + cubic, error <= 1.1e-14 for x <= 40, <= 1e-13 up to the clamp at 708).  Both stand in for np.exp in real_gas.py:159-160. */
int cmb200_test_exp_neg_host(const double *x, int64_t n, double *out);
int cmb200_test_exp_neg_table_host(const double *x, int64_t n, double *out);

#ifdef __cplusplus
}
#endif
#endif
