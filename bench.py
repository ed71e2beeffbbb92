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
N > 1    : N = G column groups x S wavenumber shards.  The S ranks of a group evaluate the same
           columns, each over its cost-balanced share of the wavenumber segments, and one NCCL
           all-reduce inside the group adds the partial net flux / heating rate; S is the largest
           divisor of N that keeps the kernel's per-warp work split balanced
           (plan.wavenumber_shards; --wavenumber-shards overrides), the G groups split the columns.
           Total work is fixed as N grows (strong scaling); e2e goes through FluxPlan.flux_sharded.
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

# Algorithmic FP64 (or FP32) operations per (chain, layer) unit, FMA = 2:
#   stage 1 fused : M multiply-adds of k*q, the layer mean and dP scaling          2M + 2
#   exp           : range reduction + degree-11 polynomial + scaling (CUDA exp)     EXP_FLOPS
#   short-wave    : two running products, two source terms, two weighted sums       8
#   long-wave     : additionally both emission recurrences                          + 14
NCU_DRAM_BYTES_PER_COLUMN = (68.149760e6 + 39.287040e6) / 8880    # profiles/r01_flux_kernel_v7_ncu_full.txt
EXP_FLOPS = {'f64': 36, 'f32': 4}


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
    ap.add_argument('--cpu-sample', type=int, default=16, help='columns timed for cpu_baseline')
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


def cpu_time_columns(w, cols, pool=None):
    T, T_g, T_tau, q, albedo = cols
    jobs = [(T[i], T_g[i], T_tau[i], q[:, i], albedo[i]) for i in range(T.shape[0])]
    t0 = time.perf_counter()
    if pool is None:
        _cpu_init(w)
        res = [_cpu_column(j) for j in jobs]
    else:
        res = pool.map(_cpu_column, jobs)
    return time.perf_counter() - t0, res


def run_reference(a):
    """--impl reference: the oracle's restatement of the reference algorithm (band transmission
    matrices, real_gas.py:564-669) on all host cores, one column per task; a bounded sample of the
    workload's columns per step."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    import multiprocessing as mp
    from climatemodel_b200 import synthetic
    w = cpu_workload(a)
    cores = os.cpu_count() or 1
    n = max(cores, min(a.cpu_sample * cores // 4, 4 * cores))
    cols = synthetic.columns(n, w['P'], seed=0)
    with mp.Pool(cores, initializer=_cpu_init, initargs=(w,)) as pool:
        for _ in range(a.warmup):
            cpu_time_columns(w, [c[..., :cores, :] if c.ndim == 3 else c[:cores] for c in cols], pool)
        t = 0.0
        for _ in range(a.steps):
            dt, _ = cpu_time_columns(w, cols, pool)
            t += dt
    units = n * a.nz * w['nu'].size
    value = units * a.steps / t
    sample = '%d of %d columns per step on %d processes (oracle restatement of the reference algorithm)' % (n, a.columns, cores)
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': a.gpus, 'steps': a.steps,
            'warmup': a.warmup, 'ms_per_step': 1e3 * t / a.steps * a.columns / n, 'higher_is_better': True,
            'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': config_dict(a, w), 'gpu_launches': 0,
            'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
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
    # N ranks = G column groups x S wavenumber shards.  The S ranks of a group evaluate
    # the same columns, each over its share of the spectrum, and one all-reduce inside the group adds the
    # partial net flux / heating rate; groups are independent.  S is as large as the spectrum allows
    # (plan.wavenumber_shards: every warp keeps a run of segments long enough to balance).
    S, G, group = 1, 1, None
    if world > 1:
        from climatemodel_b200.plan import wavenumber_shards
        S = a.wavenumber_shards or wavenumber_shards(plan.segments()[1].size, world)
        if world % S:
            raise SystemExit('--wavenumber-shards must divide the number of ranks')
        G = world // S
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
    if world > 1:
        t = torch.tensor([total_ms, kernel_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        per_rank = [torch.zeros(2, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(per_rank, torch.tensor([total_ms, kernel_ms], dtype=torch.float64, device=dev))
        per_rank_kernel_ms = [round(float(x[1]), 3) for x in per_rank]
        total_ms, kernel_ms = t.tolist()
    ms_per_step = total_ms / a.steps
    units = n_total * L * N
    value = units / (ms_per_step * 1e-3)

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
        d2h = n_col * (2 * L - 1) * 8 * world                                       # ... and the reduced result out
        e2e = {'value': units / dt, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
               'ms_per_step': dt * 1e3,
               'api': 'FluxPlan.flux_sharded: pinned host buffers, cmb200_flux + NCCL all-reduce per chunk' if S > 1
               else 'cmb200_flux_host via FluxPlan.flux (pinned host buffers)'}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (flux_kernel) -----------------------------------------------------
    peaks = _lib.measure_peaks(local)
    sw = w['bands']['sw'].astype(bool)
    chains_lw = int(w['bands']['len'][~sw].sum())
    chains_sw = int(w['bands']['len'][sw].sum())
    per_sw = 2 * M + 2 + EXP_FLOPS[a.precision] + 8
    flops = n_col * (L - 1) * (chains_sw * per_sw + chains_lw * (per_sw + 14)) / S      # this rank's launch
    key = 'fp64_tflops' if a.precision == 'f64' else 'fp32_tflops'
    achieved = flops / (kernel_ms * 1e-3) / 1e12
    roofline = {'bound': 'fp64' if a.precision == 'f64' else 'fp32', 'achieved': achieved, 'peak': peaks[key],
                'unit': 'TFLOP/s', 'frac': achieved / peaks[key], 'traffic': None,
                'kernel': 'flux_kernel<%s, fused>' % ('double' if a.precision == 'f64' else 'float'),
                'kernel_ms': kernel_ms,
                'kernel_variant': 't-store in tensor memory (LDTM/STTM), coefficient rows staged in shared memory, '
                                  'segment-space table rows (csrc/flux_kernels.cuh; CMB200_FLUX_VARIANT=0 = plain kernel)',
                'fp64_pipe_util_ncu': 0.532 if a.precision == 'f64' else None,   # profiles/r01_flux_kernel_v7_ncu_full.txt
                **({'per_rank_kernel_ms': per_rank_kernel_ms} if world > 1 else {}),
                'peak_source': 'cmb200_measure_peaks: independent-FMA issue-rate microbenchmark on this GPU '
                               '(MEASURED_PEAKS.json has no FP64/FP32 vector peak)',
                'hbm': {'algorithmic_bytes_per_step': int(n_col * ((L - 1) + 1 + L + M * L + 1 + 2 * L - 1) * 8),
                        'note': 'compulsory HBM traffic is per column, not per bin: the kernel is pipe-bound'}}
    try:
        mp = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        roofline['hbm']['measured_copy_gbs'] = mp.get('hbm_gbs')
    except Exception:
        pass
    roofline['hbm']['achieved_gbs'] = roofline['hbm']['algorithmic_bytes_per_step'] / (kernel_ms * 1e-3) / 1e9
    # DRAM traffic of the flux kernel from the committed `ncu --set full` capture (dram__bytes_read.sum +
    # dram__bytes_write.sum = 107.4 MB for 8880 columns of this workload, profiles/r01_flux_kernel_v7_ncu_full.txt),
    # scaled per column to this launch; only meaningful for the default FP64 workload shape.
    if a.precision == 'f64' and (L, M, w['nu'].size) == (100, 2, 10400):
        roofline['traffic'] = NCU_DRAM_BYTES_PER_COLUMN * n_col
        roofline['traffic_source'] = 'ncu dram bytes per column (8880-column capture) x columns of this launch'

    cpu = None
    if world == 1 and not a.no_cpu_baseline:
        import oracle.realgas_oracle  # noqa: F401  (checker / baseline only)
        ns = a.cpu_sample
        cols = [x[:, :ns] if x.ndim == 3 else x[:ns] for x in (T, T_g, T_tau, q, albedo)]
        dt, res = cpu_time_columns(w, cols)
        net_ref = np.stack([r[0] for r in res])
        got = out['net'][:ns].cpu().numpy()
        err = float(np.abs(got - net_ref).max() / np.abs(net_ref).max())
        cpu = {'value': ns * L * N / dt, 'unit': UNIT, 'cores': 1, 'kind': 'port',
               'sample': '%d of %d columns, oracle restatement of the reference algorithm (real_gas.py:564-669), '
                         'single process' % (ns, n_col),
               'net_flux_rel_err_vs_gpu': err}

    line = {'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': a.steps, 'warmup': a.warmup,
            'ms_per_step': ms_per_step, 'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None,
            'dtype': a.precision, 'data': 'synthetic',
            'config': config_dict(a, w, {'parallelism': '%d column group(s) x %d wavenumber shards, NCCL all-reduce of net/heating inside a group'
                                         % (G, S) if world > 1 else 'single GPU'}),
            'clocks': clocks, 'e2e': e2e, 'gpu_launches': launches, 'roofline': roofline, 'cpu_baseline': cpu,
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
