"""A few steps of the configs[4] ensemble (128 columns: flux + adaptive step + convective adjustment), for an
ncu launch list or a quick timing.   python tools/profile_rce_step.py [--columns 128] [--steps 20]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--columns', type=int, default=128)
    ap.add_argument('--steps', type=int, default=20)
    a = ap.parse_args()
    import torch
    from climatemodel_b200 import synthetic
    from climatemodel_b200.plan import FluxPlan
    from climatemodel_b200.radiation.columns import ColumnStepper
    dev = torch.device('cuda', 0)
    w = synthetic.workload(5.0, 100, 40)
    P, L = w['P'], w['P'].size
    n = a.columns
    Tg = np.linspace(250.0, 300.0, n)
    q = np.empty((2, n, L))
    q[0] = synthetic.co2_mass_ratio(np.full(n, 370.0))[:, None]
    q[1] = 0.005 * (P[None, :] / P[-1]) ** 3
    t = lambda x: torch.tensor(np.ascontiguousarray(x), device=dev)
    plan = FluxPlan(P, w['nu'], w['bands'], tables=w['tables'], device=0)
    cols = ColumnStepper(plan, t(np.repeat(Tg[:, None], L - 1, axis=1)), t(Tg), q=t(q),
                         T_tau=t(np.repeat(Tg[:, None], L, axis=1)), albedo=0.3)
    for _ in range(3):
        cols.step(convective_adjust=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        cols.step(convective_adjust=True)
    e1.record()
    torch.cuda.synchronize()
    print(json.dumps({'columns': n, 'steps': a.steps, 'ms_per_step': e0.elapsed_time(e1) / a.steps}))


if __name__ == '__main__':
    main()
