"""Column time-steps per second of the device-resident radiative time-stepping loop (flux kernel +
stepper kernel back to back, no host round trip) on the bench workload.
Prints one JSON line.

    python tools/bench_rce_steps.py [--columns 100000] [--steps 3]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from climatemodel_b200 import synthetic, _lib  # noqa: E402
from climatemodel_b200.plan import FluxPlan  # noqa: E402
from climatemodel_b200.radiation.columns import ColumnStepper  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--columns', type=int, default=100000)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=2)
    a = ap.parse_args()
    w = synthetic.workload()
    T, T_g, T_tau, q, albedo = synthetic.columns(a.columns, w['P'], seed=1)
    dev = torch.device('cuda', 0)
    plan = FluxPlan(w['P'], w['nu'], w['bands'], tables=w['tables'])
    t = lambda x: torch.tensor(np.ascontiguousarray(x), device=dev)
    cols = ColumnStepper(plan, t(T), t(T_g), q=t(q), T_tau=t(T_tau), albedo=t(albedo))
    for _ in range(a.warmup):
        cols.step()
    torch.cuda.synchronize()
    l0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        cols.step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    print(json.dumps({'metric': 'column time-steps/sec (flux + adaptive step on the device)', 'columns': a.columns,
                      'levels': int(w['P'].size), 'bins': int(w['nu'].size), 'steps': a.steps, 'ms_per_step': ms,
                      'value': a.columns / (ms * 1e-3), 'gpu_launches_per_step': (_lib.launch_count() - l0) / a.steps,
                      'mean_model_time_s': float(cols.t.mean()), 'T_mean_K': float(cols.T.mean())}))


if __name__ == '__main__':
    main()
