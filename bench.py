"""Benchmark of the real-gas flux hot path (BASELINE.json metric: column.level.wavenumber-bin / s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
                    [--columns C] [--precision f64|f32] [--dnu 5] [--nz 100] [--bands 40]

Workload (BASELINE.json configs[2]): batched real-gas fluxes over a synthetic lat-lon grid,
1,000,000 columns x 100 levels x ~10.4k wavenumber bins (d_nu = 5 cm^-1), 40 bands, two synthetic
molecules; each column has its own temperature profile, table-temperature indices and water-vapour
profile, so optical depth (stage 1) is rebuilt per column inside the fused kernel.  One "step" =
one flux evaluation of all columns: interface temperatures -> optical depth -> transmissivity +
Planck -> up/down sweeps -> band sums -> net flux and heating rate.

`value`  : inputs resident in HBM, CUDA events on the launch stream.
`e2e`    : the same evaluation through the host-buffer C-ABI call (cmb200_flux_host): pinned host
           inputs copied H2D and net flux + heating rate copied D2H inside the timed region.
N > 1    : N = G column groups x S wavenumber shards.  With this workload's column count the columns
           alone are split (S = 1: no data-path collective); --wavenumber-shards S makes
           the S ranks of a group evaluate the same columns over cost-balanced shares of the wavenumber
           segments with one NCCL all-reduce of net flux / heating rate inside the group (the regime of
           configs[3], tools/bench_spectral_sweep.py).  Total work is fixed as N grows (strong scaling).
parity   : at every N, columns of every group are compared with the CPU checker after the timed loop.
cpu_baseline / --impl reference: the UNMODIFIED reference staged under oracle/_ref (oracle/make_ref.py),
           optical_depth + RealGas.get_flux per column; the oracle port only if the staging is missing.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'column-level-wavenumber-bins/sec (real-gas flux evaluation)'
UNIT = 'column*level*bin/s'

# ---- algorithmic work per unit (one column * level * bin) of the shipped kernels ----------------------
# Counted from the kernel source (csrc/flux_kernels.cuh), FMA = 2 flop, every other FP instruction 1:
#   stage 1 (fused)  I = sum_m q_m/g k_m (1 mul + (M-1) fma), x = hdP (I_lo + I_hi) (1 add + 1 mul)
#   exp  FP64        table routine of sw_fresh_pass / sweep_pass (kVarExpLds): 5 DFMA + 1 DADD + 1 DMUL
#        FP32        ex2.approx + 1 mul
#   short-wave       E_fresh *= t, E_replay *= t, their sum; lane sum of a block of 8 units: FP64 8 adds through the
#                    tensor-memory transpose (6 of four lanes + 2 into the class accumulator), FP32 9 adds of the shuffle tree
#   long-wave        additionally, per sweep, w(1-t) (fma), t H (mul), E + S + c w(1-t) (fma + add), H (fma)
# Every FP64 instruction -- DFMA, DMUL or DADD -- holds the FP64 pipe for the same two cycles per warp, so the pipe
# view counts INSTRUCTIONS (views.<pipe>.issue_frac = what ncu calls pipe utilisation, restricted to the algorithm's
# own instructions); the flop view (FMA = 2) is the conventional roofline and the line's roofline.frac.
# The table stream is the compulsory memory traffic of a unit: one table value per molecule crosses L2 -> SM
# per (column, level, bin); table rows are shared by nobody (every column picks its own temperature row).
def work_model(precision, M):
    exp_instr, exp_flop = (7, 12) if precision == 'f64' else (2, 2)
    stage1_instr, stage1_flop = M + 2, 2 * M + 1
    sw_instr = sw_flop = 3 + (8.0 if precision == 'f64' else 9.0) / 8.0
    lw_instr, lw_flop = 2 * 5, 2 * 8
    word = 8 if precision == 'f64' else 4
    return {'stage1': {'instr': stage1_instr, 'flop': stage1_flop}, 'exp': {'instr': exp_instr, 'flop': exp_flop},
            'short_wave': {'instr': sw_instr, 'flop': sw_flop}, 'long_wave_extra': {'instr': lw_instr, 'flop': lw_flop},
            'table_bytes': word * M}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--columns', type=int, default=1_000_000)
    ap.add_argument('--precision', default='f64', choices=['f64', 'f32'])
    ap.add_argument('--dnu', type=float, default=5.0)
    ap.add_argument('--nz', type=int, default=100)
    ap.add_argument('--bands', type=int, default=40)
    ap.add_argument('--cpu-sample', type=int, default=12, help='columns timed for cpu_baseline (one process)')
    ap.add_argument('--ref-columns-per-core', type=int, default=3, help='--impl reference: columns per process and step')
    ap.add_argument('--no-f32', action='store_true', help='skip the FP32-mode sub-record')
    ap.add_argument('--no-rce', action='store_true', help='skip the RCE time-stepping sub-record')
    ap.add_argument('--f32-columns', type=int, default=118400)
    ap.add_argument('--rce-columns', type=int, default=100000)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--wavenumber-shards', type=int, default=0,
                    help='N>1: ranks sharing one set of columns (default: plan.wavenumber_shards rule); '
                         'the remaining factor of N splits the columns')
    return ap.parse_args()


def config_dict(a, w, extra=None):
    cfg = {'workload': 'batched real-gas fluxes, synthetic lat-lon grid (BASELINE configs[2])',
           'columns': a.columns, 'levels': a.nz, 'bins': int(w['nu'].size), 'bands': a.bands,
           'molecules': len(w['tables']), 'd_nu_cm-1': a.dnu,
           'optical_depth': 'rebuilt per column inside the kernel (per-column T indices and q)',
           'l2': 'per-step inputs and outputs (%.1f GB) exceed the 126 MB L2' % (a.columns * a.nz * 8 * 5 / 1e9)}
    if extra:
        cfg.update(extra)
    return cfg


# ---- CPU reference arm ------------------------------------------------------------------------
# The unmodified reference (oracle/_ref, staged by oracle/make_ref.py; /root/reference in the build
# container) when it is there -- kind "reference": its own optical_depth + RealGas.get_flux per column,
# oracle/ref_runner.py -- else the oracle's restatement of the same algorithm (kind "port").
_W = {}


def _cpu_init(w):
    _W.update(w)


def _cpu_column(args):
    import warnings
    warnings.filterwarnings('ignore')
    import oracle.realgas_oracle as orc
    T, T_g, T_tau, q, albedo = args
    w = _W
    tau = orc.optical_depth(w['P'], T_tau, w['nu'], w['tables'], list(q))
    up, dn = orc.get_flux(w['P'], T, T_g, tau, w['nu'], w['bands'], albedo=albedo)
    net = orc.net_flux(up, dn, w['bands']['delta'])
    return net, orc.heating_rate(net, w['P'])


def cpu_workload(a):
    import oracle.realgas_oracle as orc
    from climatemodel_b200 import synthetic
    return synthetic.workload(a.dnu, a.nz, a.bands, planck=orc.planck_wavenumber)


def reference_kind():
    try:
        from oracle import ref_runner
        return 'reference' if ref_runner.available() else 'port'
    except Exception:
        return 'port'


def cpu_time_columns(w, cols, pool=None, kind='port'):
    """Seconds for the given columns and their (net, heating) results, in this process or over `pool`."""
    if kind == 'reference':
        from oracle import ref_runner
        return ref_runner.time_columns(w, cols, pool)
    T, T_g, T_tau, q, albedo = cols
    jobs = [(T[i], T_g[i], T_tau[i], q[:, i], albedo[i]) for i in range(T.shape[0])]
    t0 = time.perf_counter()
    if pool is None:
        _cpu_init(w)
        res = [_cpu_column(j) for j in jobs]
    else:
        res = pool.map(_cpu_column, jobs)
    return time.perf_counter() - t0, res


def cpu_pool(w, cores, kind):
    import multiprocessing as mp
    if kind == 'reference':
        from oracle import ref_runner
        return mp.Pool(cores, initializer=ref_runner._init, initargs=(w,))
    return mp.Pool(cores, initializer=_cpu_init, initargs=(w,))


def baseline_note(kind):
    if kind == 'reference':
        from oracle import ref_runner
        return ('unmodified reference: optical_depth (real_gas.py:86-127) + RealGas.get_flux (:629-669) per column, '
                'from ' + ref_runner.which())
    return 'oracle restatement of the reference algorithm (real_gas.py:564-669); the reference tree is not staged'


def run_reference(a):
    """--impl reference: the reference's own CPU implementation of the path on every host core, one column
    per task, a bounded sample of the workload's columns per step; the rate is what the line reports,
    ms_per_step is the linear extrapolation to all columns and is labelled so."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    from climatemodel_b200 import synthetic
    kind = reference_kind()
    w = cpu_workload(a)
    cores = os.cpu_count() or 1
    n = cores * max(1, a.ref_columns_per_core)
    cols = synthetic.columns(n, w['P'], seed=0)
    first = [c[:, :cores] if c.ndim == 3 else c[:cores] for c in cols]
    with cpu_pool(w, cores, kind) as pool:
        for _ in range(max(a.warmup, 1)):
            cpu_time_columns(w, first, pool, kind)
        t = 0.0
        for _ in range(a.steps):
            dt, _ = cpu_time_columns(w, cols, pool, kind)
            t += dt
    units = n * a.nz * w['nu'].size
    value = units * a.steps / t
    sample = '%d of %d columns per step on %d processes; %s' % (n, a.columns, cores, baseline_note(kind))
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': a.gpus, 'steps': a.steps,
            'warmup': a.warmup, 'ms_per_step': 1e3 * t / a.steps * a.columns / n, 'higher_is_better': True,
            'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'extrapolated': True, 'sample_columns_per_step': n, 'measured_s_per_step': t / a.steps,
            'ms_per_step_note': 'measured on %d columns, scaled linearly to %d' % (n, a.columns),
            'config': config_dict(a, w), 'gpu_launches': 0,
            'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': kind, 'sample': sample,
                             'extrapolated': True},
            'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    print(json.dumps(line))


# ---- clocks ---------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    REASONS = {0x4: 'sw_power_cap', 0x8: 'hw_slowdown', 0x20: 'sw_thermal_slowdown', 0x40: 'hw_thermal_slowdown',
               0x80: 'hw_power_brake_slowdown'}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        while not self.stop_flag and self.nv is not None:
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                bits = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for b, name in self.REASONS.items():
                    if bits & b:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.1)

    def result(self):
        self.stop_flag = True
        return {'sm_mhz': float(np.median(self.samples)) if self.samples else None, 'sm_max_mhz': self.max_mhz,
                'reasons': sorted(self.reasons), 'samples': len(self.samples)}


# ---- GPU arm ------------------------------------------------------------------------------------------
def pick_shards(a, n_segments, world):
    """N ranks = G column groups x S wavenumber shards.  With plenty of columns per GPU (configs[2]: more
    than four per warp of the persistent grid) the columns alone are split -- no exchange, no duplicated
    per-column prologue; few columns x many bins (configs[3], tools/bench_spectral_sweep.py)
    shard the spectrum instead.  --wavenumber-shards overrides."""
    from climatemodel_b200.plan import wavenumber_shards
    if a.wavenumber_shards:
        return a.wavenumber_shards
    if a.columns // world >= 4 * 148 * 8:
        return 1
    return wavenumber_shards(n_segments, world)


def time_flux(torch, plan, dcols, n, reps, out=None):
    """Device-resident flux evaluation of the first n columns: mean ms per launch (CUDA events) and outputs."""
    dT, dTg, dTt, dq, dal = dcols
    args = (dT[:n], dTg[:n])
    kw = dict(T_tau=dTt[:n], q=dq[:, :n].contiguous(), albedo=dal[:n], outputs=('net', 'heating'))
    res = plan.flux_device(*args, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        res = plan.flux_device(*args, **kw)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, res


def heating_within(got, ref, net_scale, P, tol, eps):
    """Norm-wise tolerance or the rounding floor of a difference of net fluxes (tests/test_gpu_parity.py)."""
    from climatemodel_b200 import constants as k
    err = np.abs(got - ref)
    floor = 16 * eps * net_scale * k.g / k.c_p_dry / np.diff(P)
    return bool(np.all((err <= tol * np.abs(ref).max()) | (err <= floor)))


def rce_record(torch, plan, w, dev, n_col, n_steps=24):
    """BASELINE's second metric: RCE time-steps per second -- flux kernel + adaptive step + dry convective
    adjustment on the device (ColumnStepper), every column on its own clock; and the configs[4] ensemble
    (64 latitudes x {1x, 2x CO2}) stepped to steady state."""
    from climatemodel_b200 import synthetic, _lib
    from climatemodel_b200.radiation.columns import ColumnStepper
    P, L = w['P'], w['P'].size
    T, T_g, T_tau, q, albedo = synthetic.columns(n_col, P, seed=5)
    t = lambda x: torch.tensor(np.ascontiguousarray(x), device=dev)
    cols = ColumnStepper(plan, t(T), t(T_g), q=t(q), T_tau=t(T_tau), albedo=t(albedo))
    for _ in range(2):
        cols.step(convective_adjust=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_steps):
        cols.step(convective_adjust=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n_steps
    flux_ms, _ = time_flux(torch, plan, [t(x) for x in (T, T_g, T_tau, q, albedo)], n_col, 3)
    rec = {'metric': 'RCE column time-steps/sec (flux + adaptive step + convective adjustment on the device)',
           'columns': n_col, 'value': n_col / (ms * 1e-3), 'unit': 'column*step/s', 'ms_per_step': ms,
           'bare_flux_ms': flux_ms, 'step_overhead_over_flux': ms / flux_ms - 1.0}
    # configs[4]: the ensemble to steady state
    lat = np.degrees(np.arcsin((np.arange(64) + 0.5) / 64 * 2 - 1))
    lat_all, ppmv = np.repeat(lat, 2), np.tile([370.0, 740.0], 64)
    Tg = 300.0 - 50.0 * np.sin(np.radians(lat_all)) ** 2
    n = Tg.size
    qe = np.empty((2, n, L))
    qe[0] = synthetic.co2_mass_ratio(ppmv)[:, None]
    qe[1] = 0.005 * (P[None, :] / P[-1]) ** 3
    for trial in range(2):                              # the first run pays the one-time set-up of this column count
        ens = ColumnStepper(plan, t(np.repeat(Tg[:, None], L - 1, axis=1)), t(Tg), q=t(qe),
                            T_tau=t(np.repeat(Tg[:, None], L, axis=1)), albedo=0.3)
        torch.cuda.synchronize()
        l0, t0 = _lib.launch_count(), time.perf_counter()
        steps = ens.evolve_to_equilibrium(flux_thresh=1e-3, convective_adjust=True, t_end=4.0, max_steps=200000)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    surf = ens.T[:, -1].cpu().numpy()
    flux128, _ = time_flux(torch, plan, [t(x) for x in (np.repeat(Tg[:, None], L - 1, axis=1), Tg, np.repeat(Tg[:, None], L, axis=1),
                                                       qe, np.full(n, 0.3))], n, 5)
    rec['ensemble'] = {'workload': 'RCE CO2-doubling ensemble over latitude (BASELINE configs[4])', 'columns': int(n),
                       'wall_s': wall, 'column_steps': int(steps.sum()), 'max_steps_of_a_column': int(steps.max()),
                       'column_steps_per_s': float(steps.sum()) / wall, 'ms_per_ensemble_step': 1e3 * wall / int(steps.max()),
                       'bare_flux_ms_128_columns': flux128,
                       'bound': 'the flux evaluation of one column per SM (128 columns on 148 SMs): a step cannot be shorter than '
                                'one column on one SM; launches are not the limit (a CUDA-graph replay of the whole step measures the same)',
                       'gpu_launches': int(_lib.launch_count() - l0),
                       'dT_lowest_layer_2xCO2_K_mean': float((surf[1::2] - surf[0::2]).mean())}
    # the same ensemble with the flux kernel in FP32 mode (stepper, convection and equilibrium test stay FP64)
    from climatemodel_b200.plan import FluxPlan
    plan32 = FluxPlan(w['P'], w['nu'], w['bands'], tables=w['tables'], precision='f32', device=dev.index or 0)
    for trial in range(2):
        ens = ColumnStepper(plan32, t(np.repeat(Tg[:, None], L - 1, axis=1)), t(Tg), q=t(qe),
                            T_tau=t(np.repeat(Tg[:, None], L, axis=1)), albedo=0.3)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        steps32 = ens.evolve_to_equilibrium(flux_thresh=1e-3, convective_adjust=True, t_end=4.0, max_steps=200000)
        torch.cuda.synchronize()
        wall32 = time.perf_counter() - t0
    surf32 = ens.T[:, -1].cpu().numpy()
    plan32.close()
    rec['ensemble']['f32_flux'] = {'wall_s': wall32, 'column_steps': int(steps32.sum()), 'max_steps_of_a_column': int(steps32.max()),
                                   'dT_lowest_layer_2xCO2_K_mean': float((surf32[1::2] - surf32[0::2]).mean()),
                                   'lowest_layer_T_max_abs_diff_vs_f64_K': float(np.abs(surf32 - surf).max())}
    return rec


def run_b200(a):
    import torch
    import torch.distributed as dist
    from climatemodel_b200 import _lib, synthetic
    from climatemodel_b200.plan import FluxPlan

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device (there is no CPU fallback)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    w = synthetic.workload(a.dnu, a.nz, a.bands)
    plan = FluxPlan(w['P'], w['nu'], w['bands'], tables=w['tables'], precision=a.precision, device=local)
    L, N, M = plan.L, plan.N, plan.M
    S, G, group = 1, 1, None
    if world > 1:
        S = pick_shards(a, plan.segments()[1].size, world)
        if world % S:
            raise SystemExit('--wavenumber-shards must divide the number of ranks')
        G = world // S
        if S > 1:
            for g in range(G):                                  # every rank creates every group
                grp = dist.new_group(list(range(g * S, (g + 1) * S)))
                if g == rank // S:
                    group = grp
            plan.shard_by_wavenumber(rank % S, S, **({'lw_cost': float(os.environ['CMB200_SHARD_LW_COST'])}
                                                     if 'CMB200_SHARD_LW_COST' in os.environ else {}))
    n_total = a.columns
    n_col = n_total * (rank // S + 1) // G - n_total * (rank // S) // G
    T, T_g, T_tau, q, albedo = synthetic.columns(n_col, w['P'], seed=rank // S)

    def pinned(x):
        t = torch.empty(x.shape, dtype=torch.float64, pin_memory=True)
        t.numpy()[...] = x
        return t
    hT, hTg, hTt, hq, hal = (pinned(x) for x in (T, T_g, T_tau, q, albedo))
    dT, dTg, dTt, dq, dal = (t.to(dev) for t in (hT, hTg, hTt, hq, hal))
    out_buf = torch.zeros(n_col * (2 * L - 1), dtype=torch.float64, device=dev)
    out = {'net': out_buf[:n_col * L].view(n_col, L), 'heating': out_buf[n_col * L:].view(n_col, L - 1)}

    def step():
        plan.flux_device(dT, dTg, T_tau=dTt, q=dq, albedo=dal, outputs=('net', 'heating'), out=out)
        if S > 1:
            dist.all_reduce(out_buf, group=group)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 0)):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    n0 = _lib.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)]
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    ev[0].record()
    for i in range(a.steps):
        kev[i][0].record()
        plan.flux_device(dT, dTg, T_tau=dTt, q=dq, albedo=dal, outputs=('net', 'heating'), out=out)
        kev[i][1].record()
        if S > 1:
            dist.all_reduce(out_buf, group=group)
        ev[i + 1].record()
    barrier()
    launches = _lib.launch_count() - n0
    clocks = sampler.result()
    total_ms = ev[0].elapsed_time(ev[-1])
    kernel_ms = float(np.mean([s.elapsed_time(e) for s, e in kev]))
    per_rank_kernel_ms = None
    if world > 1:
        t = torch.tensor([total_ms, kernel_ms], dtype=torch.float64, device=dev)
        per_rank = [torch.zeros(2, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(per_rank, t.clone())
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        per_rank_kernel_ms = [round(float(x[1]), 3) for x in per_rank]
        total_ms, kernel_ms = t.tolist()
    ms_per_step = total_ms / a.steps
    units = n_total * L * N
    value = units / (ms_per_step * 1e-3)

    # ---- parity of what the timed loop produced, at every N: a few (reduced) columns of every column group
    # against the CPU checker (the unmodified reference when it is staged, else the oracle) ---------------
    import oracle.realgas_oracle  # noqa: F401  (checker only)
    kind = reference_kind()
    ns = a.cpu_sample if world == 1 else 2
    err_net = err_heat = 0.0
    heat_ok = True
    cpu_dt = None
    if not a.no_cpu_baseline and rank % S == 0 and a.precision == 'f64':
        wc = cpu_workload(a) if kind == 'port' else w
        cols_s = [x[:, :ns] if x.ndim == 3 else x[:ns] for x in (T, T_g, T_tau, q, albedo)]
        cpu_dt, res = cpu_time_columns(wc, cols_s, None, kind)
        net_ref = np.stack([r[0] for r in res])
        heat_ref = np.stack([r[1] for r in res])
        got_n, got_h = out['net'][:ns].cpu().numpy(), out['heating'][:ns].cpu().numpy()
        err_net = float(np.abs(got_n - net_ref).max() / np.abs(net_ref).max())
        err_heat = float(np.abs(got_h - heat_ref).max() / np.abs(heat_ref).max())
        heat_ok = all(heating_within(got_h[i], heat_ref[i], 1.5e3, w['P'], 1e-10, np.finfo(float).eps) for i in range(ns))
    if world > 1:
        e = torch.tensor([err_net, err_heat, 0.0 if heat_ok else 1.0], dtype=torch.float64, device=dev)
        dist.all_reduce(e, op=dist.ReduceOp.MAX)
        err_net, err_heat, heat_ok = float(e[0]), float(e[1]), float(e[2]) == 0.0

    # ---- end to end through the host-buffer C ABI -----------------------------------------------------
    e2e = None
    if not a.no_e2e:
        h_out = {'net': torch.empty((n_col, L), dtype=torch.float64, pin_memory=True).numpy(),
                 'heating': torch.empty((n_col, L - 1), dtype=torch.float64, pin_memory=True).numpy()}
        args = dict(T_tau=hTt.numpy(), q=hq.numpy(), albedo=hal.numpy(), outputs=('net', 'heating'), out=h_out)

        def e2e_step():
            if S > 1:                     # copies in, kernel, all-reduce on the device, copies out -- per chunk
                plan.flux_sharded(hT.numpy(), hTg.numpy(), hTt.numpy(), hq.numpy(), hal.numpy(), h_out, group=group)
            else:
                plan.flux(hT.numpy(), hTg.numpy(), **args)
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        k_e2e = min(a.steps, 3)
        for _ in range(k_e2e):
            e2e_step()
        barrier()
        dt = (time.perf_counter() - t0) / k_e2e
        if world > 1:
            tt = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = tt.item()
        h2d = sum(x.numel() * 8 for x in (hT, hTg, hTt, hq, hal)) * world          # every rank copies its columns in
        d2h = n_col * (2 * L - 1) * 8 * world                                       # ... and the result out
        e2e = {'value': units / dt, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
               'ms_per_step': dt * 1e3,
               'api': 'FluxPlan.flux_sharded: pinned host buffers, cmb200_flux + NCCL all-reduce per chunk' if S > 1
               else 'cmb200_flux_host via FluxPlan.flux (pinned host buffers)'}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (flux_kernel), every input of the arithmetic in the line ---------
    peaks = _lib.measure_peaks(local)
    wm = work_model(a.precision, M)
    sw = w['bands']['sw'].astype(bool)
    chains_lw, chains_sw = int(w['bands']['len'][~sw].sum()), int(w['bands']['len'][sw].sum())
    launch_units = n_col * (L - 1) * (chains_sw + chains_lw) / S                    # (chain, layer) units of this rank's launch
    lw_share = chains_lw / (chains_sw + chains_lw)
    flop_unit = wm['stage1']['flop'] + wm['exp']['flop'] + wm['short_wave']['flop'] + lw_share * wm['long_wave_extra']['flop']
    instr_unit = wm['stage1']['instr'] + wm['exp']['instr'] + wm['short_wave']['instr'] + lw_share * wm['long_wave_extra']['instr']
    sec = kernel_ms * 1e-3
    pipe = 'fp64' if a.precision == 'f64' else 'fp32'
    pipe_peak = peaks[pipe + '_reg_tflops']
    pipe_ach = flop_unit * launch_units / sec / 1e12
    issue_peak = pipe_peak * 1e12 / 2.0                                              # FMA-pipe instructions (lanes) per second
    l2_ach = wm['table_bytes'] * launch_units / sec / 1e9
    views = {
        pipe: {'achieved': pipe_ach, 'peak': pipe_peak, 'unit': 'TFLOP/s', 'frac': pipe_ach / pipe_peak,
               'flop_per_unit': flop_unit, 'instr_per_unit': instr_unit,
               'issue_frac': instr_unit * launch_units / sec / issue_peak,
               'peak_source': 'cmb200_measure_peaks2: DFMA/FFMA loop, register operands, 8 warps per SM, on this GPU '
                              '(constant-operand loop of round 1: %.2f TFLOP/s)' % peaks[pipe + '_tflops']},
        'l2': {'achieved': l2_ach, 'peak': peaks['l2_stream_gbs'], 'unit': 'GB/s', 'frac': l2_ach / peaks['l2_stream_gbs'],
               'bytes_per_unit': wm['table_bytes'],
               'peak_source': 'cmb200_measure_peaks2: 64 MB L2-resident buffer read repeatedly, coalesced 16-byte loads, '
                              'full occupancy, on this GPU'},
        'hbm': {'achieved': n_col * ((L - 1) + 1 + L + M * L + 1 + 2 * L - 1) * 8 / sec / 1e9, 'unit': 'GB/s',
                'note': 'compulsory HBM traffic is per column (inputs and outputs), not per bin: not the bound'}}
    try:
        mp = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        views['hbm']['peak'] = mp.get('hbm_gbs')
        views['hbm']['frac'] = views['hbm']['achieved'] / mp['hbm_gbs']
    except Exception:
        pass
    # The arithmetic pipe is the roofline the north_star names for this regime (the table rows never leave L2 and
    # L2 -> SM bandwidth is far from its ceiling); what ncu shows closest to its limit is the SM's load/store data
    # path (shared-memory exp table, level data, shuffles, table rows: profiles/r02_flux_kernel_ncu_full.txt).
    bound = pipe
    traffic, ncu_rec = None, None
    try:        # one ncu capture of this kernel per round, committed under profiles/ (tools/ncu_summary.py); never measured here
        ncu_rec = json.load(open(os.path.join(ROOT, 'profiles', 'r02_flux_kernel_ncu.json')))
        dr = ncu_rec.get('dram_1M_columns')
        if dr and a.precision == 'f64' and dr.get('columns') == n_col and S == 1:
            traffic = {'bytes_per_launch': dr['dram_bytes_read'] + dr['dram_bytes_write'], 'from': dr['from']}
    except Exception:
        pass
    roofline = {'bound': bound, 'achieved': views[bound]['achieved'], 'peak': views[bound]['peak'],
                'unit': views[bound]['unit'], 'frac': views[bound]['frac'], 'traffic': traffic,
                'pipe_issue_frac': views[bound]['issue_frac'],
                'kernel': 'flux_kernel<%s, fused>' % ('double' if a.precision == 'f64' else 'float'),
                'kernel_ms': kernel_ms, 'launch_units': launch_units, 'long_wave_share_of_units': lw_share,
                'work_model': wm, 'views': views,
                'how': 'achieved = per-unit figure of work_model x launch_units / kernel_ms; frac = achieved / peak of the view named '
                       'by bound; views.l2 = the table stream against the measured L2 -> SM read bandwidth; DESIGN.md section 5',
                'kernel_variant': 'FP64: shared-memory exp table, direct short-wave pass, t-store and lane sums in tensor memory '
                                  '(LDTM / STTM, transposing LDTM.16x256b); CMB200_FLUX_VARIANT=5123 shuffle lane sums, =3 round-1 kernel, '
                                  '=0 plain kernel',
                **({'per_rank_kernel_ms': per_rank_kernel_ms,
                    'per_rank_kernel_ms_spread': (max(per_rank_kernel_ms) - min(per_rank_kernel_ms)) / max(per_rank_kernel_ms)}
                   if world > 1 else {}),
                'ncu': (ncu_rec or {}).get('f64' if a.precision == 'f64' else 'f32')}

    cpu = None
    if cpu_dt is not None:
        cpu = {'value': ns * L * N / cpu_dt, 'unit': UNIT, 'cores': 1, 'kind': kind,
               'sample': '%d of %d columns, single process; %s' % (ns, n_col, baseline_note(kind)),
               'extrapolated': False}
    parity = {'checker': kind, 'columns_checked_per_group': ns, 'groups': G,
              'net_flux_rel_err_vs_oracle': err_net, 'heating_rel_err_vs_oracle': err_heat,
              'heating_within_1e-10_or_rounding_floor': heat_ok, 'tolerance': 1e-10}
    if cpu is not None:
        cpu['net_flux_rel_err_vs_gpu'] = err_net

    # ---- sub-records (N = 1): FP32 mode and the RCE time-stepper -------------------------------------------
    f32 = rce = None
    if world == 1 and a.precision == 'f64' and not a.no_f32:
        n32 = min(a.f32_columns, n_col)
        plan32 = FluxPlan(w['P'], w['nu'], w['bands'], tables=w['tables'], precision='f32', device=local)
        ms32, o32 = time_flux(torch, plan32, (dT, dTg, dTt, dq, dal), n32, 3)
        ms64, o64 = time_flux(torch, plan, (dT, dTg, dTt, dq, dal), n32, 3)
        wm32 = work_model('f32', M)
        f32_flop = wm32['stage1']['flop'] + wm32['exp']['flop'] + wm32['short_wave']['flop'] + lw_share * wm32['long_wave_extra']['flop']
        u32 = n32 * (L - 1) * (chains_sw + chains_lw)
        n64, h64 = o64['net'].cpu().numpy(), o64['heating'].cpu().numpy()
        n32a, h32a = o32['net'].cpu().numpy(), o32['heating'].cpu().numpy()
        f32 = {'columns': n32, 'value': n32 * L * N / (ms32 * 1e-3), 'unit': UNIT, 'ms': ms32,
               'speedup_over_f64_same_columns': ms64 / ms32,
               'net_flux_rel_err_vs_f64': float(np.abs(n32a - n64).max() / np.abs(n64).max()),
               'heating_rel_err_vs_f64': float(np.abs(h32a - h64).max() / np.abs(h64).max()),
               'heating_within_1e-5_or_fp32_rounding_floor': all(
                   heating_within(h32a[i], h64[i], 1.5e3, w['P'], 1e-5, np.finfo(np.float32).eps) for i in range(min(n32, 256))),
               'heating_note': 'a heating rate is g/c_p times the DIFFERENCE of net fluxes of adjacent interfaces over dp: FP32 chains '
                               'leave ~1e-7 of the net flux (1.5e3 W m^-2 scale) in it, which the norm-wise 1e-5 cannot see through; the '
                               'test bound is tolerance OR 16 eps32 x flux scale x g/c_p/dp per layer (tests/test_gpu_parity.py:heating_ok)',
               'tolerance': 1e-5,
               'kernel_variant': 'FP32: direct short-wave passes carrying two segments per lane in packed FP32 arithmetic (FFMA2 / FMUL2 / '
                                 'FADD2), t-store and lane sums in tensor memory, 8 warps; CMB200_FLUX_VARIANT=28675 one segment per pass on 12 warps',
               'bound': 'issue slots 46 % used, neither FP32 pipe (fma 21 %, xu 25 %) nor the table stream (all table loads served from L1: '
                        '0.7 % faster): fixed-latency waits -- the XU serialises a block\'s sixteen MUFU.EX2 -- with two warps per scheduler to hide '
                        'them; DESIGN.md section 5',
               'roofline': {'fp32': {'achieved': f32_flop * u32 / (ms32 * 1e-3) / 1e12, 'peak': peaks['fp32_reg_tflops'],
                                     'unit': 'TFLOP/s', 'frac': f32_flop * u32 / (ms32 * 1e-3) / 1e12 / peaks['fp32_reg_tflops'],
                                     'flop_per_unit': f32_flop},
                            'l2': {'achieved': wm32['table_bytes'] * u32 / (ms32 * 1e-3) / 1e9, 'peak': peaks['l2_stream_gbs'],
                                   'unit': 'GB/s', 'frac': wm32['table_bytes'] * u32 / (ms32 * 1e-3) / 1e9 / peaks['l2_stream_gbs']}}}
        if cpu_dt is not None:
            k = min(ns, n32)
            f32['net_flux_rel_err_vs_oracle'] = float(np.abs(n32a[:k] - net_ref[:k]).max() / np.abs(net_ref[:k]).max())
            f32['heating_rel_err_vs_oracle'] = float(np.abs(h32a[:k] - heat_ref[:k]).max() / np.abs(heat_ref[:k]).max())
        plan32.close()
    if world == 1 and a.precision == 'f64' and not a.no_rce:
        rce = rce_record(torch, plan, w, dev, min(a.rce_columns, n_col))

    line = {'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': a.steps, 'warmup': a.warmup,
            'ms_per_step': ms_per_step, 'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None,
            'dtype': a.precision, 'data': 'synthetic',
            'config': config_dict(a, w, {'parallelism': ('%d column group(s) x %d wavenumber shard(s)%s' % (
                G, S, ', NCCL all-reduce of net/heating inside a group' if S > 1 else ': columns split over the ranks, no data-path collective'))
                if world > 1 else 'single GPU'}),
            'clocks': clocks, 'e2e': e2e, 'gpu_launches': launches, 'roofline': roofline, 'cpu_baseline': cpu,
            'parity': parity, 'net_flux_rel_err_vs_oracle': err_net, 'f32': f32, 'rce': rce,
            'peaks_measured_here': peaks}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    args = parse()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)
