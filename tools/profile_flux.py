"""One short, fixed invocation of the flux kernel for ncu captures and quick timings.

    python tools/profile_flux.py [--columns 1480] [--reps 3] [--precision f64] [--variant V] [--check]

bench.py's workload shape (100 levels x 10.4k bins, 2 molecules, per-column T / q), a column count
that is a multiple of the SM count.  Prints one JSON line: ms per evaluation (CUDA events),
column-level-bins per second and, with --check, the error of net flux / heating rate against the
oracle on three columns.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--columns', type=int, default=1480)
    ap.add_argument('--reps', type=int, default=3)
    ap.add_argument('--precision', default='f64')
    ap.add_argument('--variant', default=None)
    ap.add_argument('--check', action='store_true')
    ap.add_argument('--levels', type=int, default=100)
    a = ap.parse_args()
    if a.variant is not None:
        os.environ['CMB200_FLUX_VARIANT'] = a.variant

    import torch
    from climatemodel_b200 import synthetic
    from climatemodel_b200.plan import FluxPlan

    dev = torch.device('cuda', 0)
    w = synthetic.workload(5.0, a.levels, 40)
    cols = synthetic.columns(a.columns, w['P'], seed=0)
    T, T_g, T_tau, q, albedo = cols
    dT, dTg, dTt, dq, dal = [torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in cols]
    plan = FluxPlan(w['P'], w['nu'], w['bands'], tables=w['tables'], precision=a.precision, device=0)

    def step():
        return plan.flux_device(dT, dTg, T_tau=dTt, q=dq, albedo=dal, outputs=('net', 'heating'))
    for _ in range(2):
        out = step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.reps
    rec = {'columns': a.columns, 'levels': a.levels, 'bins': int(w['nu'].size), 'precision': a.precision,
           'variant': a.variant, 'ms': ms, 'units_per_s': a.columns * a.levels * w['nu'].size / (ms * 1e-3)}
    if a.check:
        import oracle.realgas_oracle as orc
        net, heat = out['net'].cpu().numpy(), out['heating'].cpu().numpy()
        worst_n = worst_h = 0.0
        for i in range(3):
            tau = orc.optical_depth(w['P'], T_tau[i], w['nu'], w['tables'], list(q[:, i]))
            up, dn = orc.get_flux(w['P'], T[i], T_g[i], tau, w['nu'], w['bands'], albedo=albedo[i])
            ref = orc.net_flux(up, dn, w['bands']['delta'])
            worst_n = max(worst_n, float(np.abs(net[i] - ref).max() / np.abs(ref).max()))
            href = orc.heating_rate(ref, w['P'])
            worst_h = max(worst_h, float(np.abs(heat[i] - href).max() / np.abs(href).max()))
        rec['net_vs_oracle'] = worst_n
        rec['heating_vs_oracle'] = worst_h
    print(json.dumps(rec), flush=True)


if __name__ == '__main__':
    main()
