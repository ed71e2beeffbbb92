"""A/B of the flux kernel variants (CMB200_FLUX_VARIANT, csrc/flux_kernels.cuh) on one GPU.

    python tools/ab_flux_variants.py [--columns 17760] [--variants 0,3,5123,28675,default] [--reps 5] [--out file.json]

Same workload shape as bench.py (100 levels x 10.4k bins, 2 molecules, per-column T / q), a
column count that is a multiple of the SM count; every variant is timed with CUDA events and its
net flux / heating rate compared with variant 0's (same inputs) and, on a few columns, with the
oracle.  Both precisions, the resident-tau path and the split (few columns) mode are covered.
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
    ap.add_argument('--columns', type=int, default=17760)
    ap.add_argument('--variants', default='0,3,5123,28675,default',
                    help="CMB200_FLUX_VARIANT values; 'default' = unset (a value a precision does not know runs its default)")
    ap.add_argument('--reps', type=int, default=5)
    ap.add_argument('--out', default=None)
    a = ap.parse_args()
    variants = [v if v == 'default' else int(v) for v in a.variants.split(',')]

    import torch
    from climatemodel_b200 import synthetic
    from climatemodel_b200.plan import FluxPlan
    import oracle.realgas_oracle as orc

    dev = torch.device('cuda', 0)
    w = synthetic.workload(5.0, 100, 40)
    cols = synthetic.columns(a.columns, w['P'], seed=0)
    T, T_g, T_tau, q, albedo = cols
    dcols = [torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in cols]
    results = {'columns': a.columns, 'levels': 100, 'bins': int(w['nu'].size), 'runs': []}

    # oracle on a few columns
    ns = 3
    tau0 = [orc.optical_depth(w['P'], T_tau[i], w['nu'], w['tables'], list(q[:, i])) for i in range(ns)]
    ref_net = []
    for i in range(ns):
        up, dn = orc.get_flux(w['P'], T[i], T_g[i], tau0[i], w['nu'], w['bands'], albedo=albedo[i])
        ref_net.append(orc.net_flux(up, dn, w['bands']['delta']))
    ref_net = np.stack(ref_net)

    def rel(x, y):
        return float(np.abs(x - y).max() / np.abs(y).max())

    def run(plan, label, n, reps, **kw):
        base = None
        for v in variants:
            if v == 'default':
                os.environ.pop('CMB200_FLUX_VARIANT', None)
            else:
                os.environ['CMB200_FLUX_VARIANT'] = str(v)
            args = [t[:n] if t.dim() < 3 else t[:, :n].contiguous() for t in dcols]
            dT, dTg, dTt, dq, dal = args

            def step():
                if 'tau' in kw:
                    return plan.flux_device(dT, dTg, tau=kw['tau'], albedo=dal, outputs=('net', 'heating'))
                return plan.flux_device(dT, dTg, T_tau=dTt, q=dq, albedo=dal, outputs=('net', 'heating'))
            for _ in range(2):
                out = step()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                out = step()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            net, heat = out['net'].cpu().numpy(), out['heating'].cpu().numpy()
            rec = {'case': label, 'variant': v, 'columns': n, 'ms': ms, 'finite': bool(np.isfinite(net).all() and np.isfinite(heat).all())}
            if base is None:
                base = (net, heat)
            rec['net_vs_first'] = rel(net, base[0])
            rec['heating_vs_first'] = rel(heat, base[1])
            if 'tau' not in kw and n >= ns:
                rec['net_vs_oracle'] = rel(net[:ns], ref_net)
            results['runs'].append(rec)
            print(json.dumps(rec), flush=True)

    for prec in ('f64', 'f32'):
        plan = FluxPlan(w['P'], w['nu'], w['bands'], tables=w['tables'], precision=prec, device=0)
        run(plan, 'fused_' + prec, a.columns, a.reps)
        run(plan, 'fused_split_' + prec, 7, 3)          # fewer columns than SMs: several CTAs per column
        if prec == 'f64':
            tau = torch.from_numpy(np.ascontiguousarray(tau0[0])).to(dev)
            run(plan, 'resident_tau_' + prec, 1480, 3, tau=tau)
    os.environ.pop('CMB200_FLUX_VARIANT', None)
    if a.out:
        with open(a.out, 'w') as f:
            json.dump(results, f, indent=1)


if __name__ == '__main__':
    main()
