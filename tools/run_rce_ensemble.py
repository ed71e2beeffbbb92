"""BASELINE configs[4]: radiative-convective equilibrium with convective adjustment, a CO2-doubling
ensemble over a latitude grid, time-stepped to steady state on the device.

    python tools/run_rce_ensemble.py [--latitudes 64] [--flux-thresh 1e-3] [--t-end 4.0]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tools/run_rce_ensemble.py ...

Columns = latitudes x {1x, 2x CO2} (370 / 740 ppmv), each the batched equivalent of one
``RealGas(ny=1)`` run of real_gas_script.py: surface temperature T_g(latitude) fixed, isothermal
start at T_g, optical depth from the isothermal profile (real_gas.py:285-292), then
``evolve_to_equilibrium(convective_adjust=True)`` (base.py:266-335) -- flux kernel, adaptive step,
dry convective adjustment and the per-column equilibrium test all on the GPU
(``ColumnStepper.evolve_to_equilibrium``).  Columns are independent, so N ranks split them with no
data-path collective; rank 0 gathers the final profiles for the report.  Prints one JSON line.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--latitudes', type=int, default=64)
    ap.add_argument('--flux-thresh', type=float, default=1e-3)
    ap.add_argument('--t-end', type=float, default=4.0)
    ap.add_argument('--max-steps', type=int, default=200000)
    ap.add_argument('--dnu', type=float, default=5.0)
    ap.add_argument('--nz', type=int, default=100)
    ap.add_argument('--no-convection', action='store_true')
    ap.add_argument('--precision', default='f64', choices=['f64', 'f32'], help='arithmetic of the flux kernel (the stepper stays FP64)')
    ap.add_argument('--repeat', type=int, default=2, help='runs of the whole ensemble; the last one is reported (the first pays one-time set-up)')
    a = ap.parse_args()

    import torch
    import torch.distributed as dist
    from climatemodel_b200 import _lib, synthetic
    from climatemodel_b200.plan import FluxPlan
    from climatemodel_b200.radiation.columns import ColumnStepper, column_share, gather_columns

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    w = synthetic.workload(a.dnu, a.nz, 40)
    P = w['P']
    L = P.size
    # the ensemble: Gaussian-like latitudes, two CO2 amounts per latitude
    lat = np.degrees(np.arcsin((np.arange(a.latitudes) + 0.5) / a.latitudes * 2 - 1))
    lat_all = np.repeat(lat, 2)
    ppmv_all = np.tile([370.0, 740.0], a.latitudes)
    n_all = lat_all.size
    mine = column_share(n_all, rank, world)               # round-robin: neighbours in latitude cost the same
    T_g = 300.0 - 50.0 * np.sin(np.radians(lat_all[mine])) ** 2
    n_col = mine.size
    T0 = np.repeat(T_g[:, None], L - 1, axis=1)            # isothermal start (real_gas.py:285-289)
    T_tau = np.repeat(T_g[:, None], L, axis=1)
    q = np.empty((2, n_col, L))
    q[0] = synthetic.co2_mass_ratio(ppmv_all[mine])[:, None]
    q[1] = 0.005 * (P[None, :] / P[-1]) ** 3              # H2O-like, decaying with height
    t = lambda x: torch.tensor(np.ascontiguousarray(x), device=dev)
    plan = FluxPlan(P, w['nu'], w['bands'], tables=w['tables'], precision=a.precision, device=local)
    for _ in range(max(1, a.repeat)):
        cols = ColumnStepper(plan, t(T0), t(T_g), q=t(q), T_tau=t(T_tau), albedo=0.3)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        l0 = _lib.launch_count()
        t0 = time.perf_counter()
        n_steps = cols.evolve_to_equilibrium(flux_thresh=a.flux_thresh, convective_adjust=not a.no_convection,
                                             t_end=a.t_end, max_steps=a.max_steps)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        launches = _lib.launch_count() - l0

    out = torch.cat((cols.T, cols.t[:, None], n_steps.to(torch.float64)[:, None],
                     cols.net_flux.abs().amax(dim=1, keepdim=True)), dim=1)          # [n_col, L-1 + 3]
    stats = torch.tensor([wall, float(n_steps.sum()), float(n_steps.max()), float(launches)], dtype=torch.float64, device=dev)
    full = gather_columns(out, n_all, rank, world).cpu().numpy()
    if world > 1:
        all_stats = [torch.zeros_like(stats) for _ in range(world)]
        dist.all_gather(all_stats, stats)
        st = torch.stack(all_stats).cpu().numpy()
        wall, col_steps, max_steps, launches = st[:, 0].max(), st[:, 1].sum(), st[:, 2].max(), st[:, 3].sum()
    else:
        wall, col_steps, max_steps, launches = stats.cpu().numpy()
    if rank == 0:
        T_fin = full[:, :L - 1]
        surf_air = T_fin[:, -1]                            # lowest layer (index 0 = top of atmosphere)
        d_co2 = surf_air[1::2] - surf_air[0::2]
        strat = T_fin[:, L // 8]
        line = {'metric': 'RCE column time-steps/sec (flux + adaptive step + convective adjustment + equilibrium test on the device)',
                'value': col_steps / wall, 'unit': 'column*step/s', 'n_gpus': world, 'scaling': 'strong',
                'config': {'workload': 'RCE CO2-doubling ensemble over latitude (BASELINE configs[4])',
                           'latitudes': a.latitudes, 'members': 2, 'columns': int(n_all), 'levels': int(L),
                           'bins': int(w['nu'].size), 'flux_thresh': a.flux_thresh, 't_end_years': a.t_end,
                           'convective_adjust': not a.no_convection, 'flux_precision': a.precision, 'runs': a.repeat},
                'wall_s': wall, 'column_steps': int(col_steps), 'max_steps_of_a_column': int(max_steps),
                'gpu_launches': int(launches),
                'model_years_mean': float(full[:, L - 1].mean() / (365 * 86400)),
                'max_abs_net_flux_at_end_W_m2': float(full[:, L + 1].max()),
                'lowest_layer_T_K': {'min': float(surf_air.min()), 'max': float(surf_air.max())},
                'dT_lowest_layer_2xCO2_K': {'mean': float(d_co2.mean()), 'min': float(d_co2.min()), 'max': float(d_co2.max())},
                'dT_upper_level_2xCO2_K_mean': float((strat[1::2] - strat[0::2]).mean())}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
