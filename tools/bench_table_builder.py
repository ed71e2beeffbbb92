"""Throughput of the line-by-line table builder on a HITRAN-sized synthetic
line list, next to a NumPy restatement of the reference loop (hitran.py:223-247) on a sample of
the lines.  Prints one JSON line.

    python tools/bench_table_builder.py [--lines 100000]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from climatemodel_b200.radiation.real_gas_data import hitran  # noqa: E402


def numpy_reference(p, T, nu, lines, n_line_widths):
    out = np.zeros((p.size, nu.size))
    for i in range(lines['nu'].size):
        g = hitran.gamma_extrapolate(p, T, lines['gamma_air'][i], lines['n_air'][i])
        s = hitran.s_extrapolate(T, lines['sw'][i], lines['nu'][i], lines['n_air'][i])
        i1, i2 = hitran.wavenumbers_near_line(nu, nu[1] - nu[0], lines['nu'][i], g.max(), n_line_widths)
        out[:, i1:i2] += s.reshape(-1, 1) * hitran.lorentzian_profile(nu[i1:i2].reshape(1, -1), lines['nu'][i],
                                                                      g.reshape(-1, 1))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--lines', type=int, default=100000)
    ap.add_argument('--cpu-lines', type=int, default=2000)
    a = ap.parse_args()
    rng = np.random.default_rng(0)
    lines = {'nu': np.sort(rng.uniform(0, 4500, a.lines)), 'sw': 10 ** rng.uniform(-6, 3, a.lines),
             'gamma_air': rng.uniform(0.04, 0.11, a.lines), 'n_air': rng.uniform(0.4, 0.8, a.lines)}
    p = hitran.table_p_values.copy()
    nu = np.arange(0, 4500 + 5, hitran.table_dnu, dtype=float)
    T = np.full(p.size, 270.0)
    hitran.get_absorption_coefficient(p[:2], T[:2], nu, 'custom', {k: v[:10] for k, v in lines.items()})   # warm-up
    t0 = time.perf_counter()
    k = hitran.get_absorption_coefficient(p, T, nu, 'custom', lines)
    dt = time.perf_counter() - t0
    sub = {key: v[:: a.lines // a.cpu_lines][: a.cpu_lines] for key, v in lines.items()}
    t0 = time.perf_counter()
    ref = numpy_reference(p, T, nu, sub, 1000)
    dt_cpu = time.perf_counter() - t0
    got = hitran.get_absorption_coefficient(p, T, nu, 'custom', sub)
    nz = ref != 0
    print(json.dumps({'metric': 'line-level evaluations/sec (one temperature slice of a lookup table)',
                      'lines': a.lines, 'levels': int(p.size), 'bins': int(nu.size),
                      'value': a.lines * p.size / dt, 'seconds': dt, 'includes': 'host<->device copies',
                      'cpu_numpy': {'value': sub['nu'].size * p.size / dt_cpu, 'lines': int(sub['nu'].size),
                                    'seconds': dt_cpu, 'cores': 1},
                      'max_rel_err_vs_numpy': float(np.abs(got[nz] / ref[nz] - 1).max()),
                      'checksum': float(k.sum())}))


if __name__ == '__main__':
    main()
