"""BASELINE configs[3]: high-resolution spectral sweep -- few columns x 100 levels x ~1M wavenumber
bins (d_nu = 0.05 cm^-1), the table-streaming regime: every column reads its own
M x L table rows of N bins once, so the compulsory HBM traffic is sizeof(real) * M bytes per
column.level.bin.  Prints one JSON line with the achieved rate against both rooflines.

    python tools/bench_spectral_sweep.py [--columns 4] [--dnu 0.05] [--steps 10]
    python -m torch.distributed.run --nproc-per-node N ... tools/bench_spectral_sweep.py   (wavenumber shards)
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--columns', type=int, default=4)
    ap.add_argument('--dnu', type=float, default=0.05)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--precision', default='f64')
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from climatemodel_b200 import _lib, synthetic
    from climatemodel_b200.plan import FluxPlan
    world, rank = int(os.environ.get('WORLD_SIZE', '1')), int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    t0 = time.perf_counter()
    w = synthetic.workload(a.dnu, 100, 40)
    t_tables = time.perf_counter() - t0
    plan = FluxPlan(w['P'], w['nu'], w['bands'], tables=w['tables'], precision=a.precision, device=local)
    L, N, M = plan.L, plan.N, plan.M
    n_seg = plan.segments()[1].size
    if world > 1:
        plan.shard_by_wavenumber(rank, world)
    T, T_g, T_tau, q, albedo = (torch.from_numpy(np.ascontiguousarray(x)).to(dev)
                                for x in synthetic.columns(a.columns, w['P'], seed=0))
    out_buf = torch.zeros(a.columns * (2 * L - 1), dtype=torch.float64, device=dev)
    out = {'net': out_buf[:a.columns * L].view(a.columns, L), 'heating': out_buf[a.columns * L:].view(a.columns, L - 1)}

    def step():
        plan.flux_device(T, T_g, T_tau=T_tau, q=q, albedo=albedo, outputs=('net', 'heating'), out=out)
        if world > 1:
            dist.all_reduce(out_buf)
    for _ in range(a.warmup):
        step()
    # d_nu = 0.05: the table rows of one evaluation (1.7 GB per column) exceed the 126 MB L2
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = ms.item()
    if rank == 0:
        size = 8 if a.precision == 'f64' else 4
        units = a.columns * L * N
        table_bytes = units * M * size                       # every (column, level, molecule) row of N bins once
        peaks = _lib.measure_peaks(local)
        hbm_peak = 6467.0
        try:
            hbm_peak = float(json.load(open(os.path.join(os.path.dirname(__file__), '..', 'MEASURED_PEAKS.json')))['hbm_gbs'])
        except Exception:
            pass
        print(json.dumps({
            'metric': 'column-level-wavenumber-bins/sec (real-gas flux evaluation)', 'value': units / (ms * 1e-3),
            'unit': 'column*level*bin/s', 'n_gpus': world, 'ms_per_step': ms, 'steps': a.steps, 'warmup': a.warmup,
            'dtype': a.precision, 'scaling': 'strong',
            'config': {'workload': 'high-resolution spectral sweep (BASELINE configs[3])', 'columns': a.columns,
                       'levels': L, 'bins': N, 'segments': int(n_seg), 'molecules': M, 'd_nu_cm-1': a.dnu,
                       'parallelism': 'wavenumber shards x%d + all-reduce' % world if world > 1 else 'single GPU',
                       'table_build_s (host NumPy)': round(t_tables, 1)},
            'roofline_hbm': {'algorithmic_bytes': table_bytes, 'achieved_gbs': table_bytes / (ms * 1e-3) / 1e9 / world,
                             'peak_gbs': hbm_peak, 'frac': table_bytes / (ms * 1e-3) / 1e9 / world / hbm_peak,
                             'note': 'per GPU; sizeof(real) * M bytes per unit'},
            'fp64_peak_tflops': peaks.get('fp64_tflops') if isinstance(peaks, dict) else None}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
