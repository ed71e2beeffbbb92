"""Shared helpers for the test-suite: golden-case and table loading."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, 'golden')
TABLES = os.path.join(GOLDEN, 'tables')

CASES = ['co2_nz50_earth', 'co2_nz100_iso', 'gray_nz60_200b', 'single_line_nz30', 'two_mol_nz41']


def load_table(name):
    d = np.load(os.path.join(TABLES, name + '.npz'))
    return {k: d[k] for k in d.files}


def load_case(name):
    d = np.load(os.path.join(GOLDEN, name + '.npz'))
    c = {k: d[k] for k in d.files}
    c['bands'] = {'start': c['band_start'], 'len': c['band_len'], 'centre': c['band_centre'],
                  'delta': c['band_delta'], 'sw': c['band_sw']}
    c['tables'] = [load_table(str(m)) for m in c['molecule_names']]
    for k in ('T_g', 'albedo', 'T_star', 'R_star', 'dist'):
        c[k] = float(c[k])
    return c


def relmax(a, b):
    """Norm-wise relative error max|a-b| / max|b| (the parity metric of SURVEY.md 7.3-3)."""
    a = np.asarray(a, dtype=float)
    b = np.asarray(b, dtype=float)
    scale = float(np.abs(b).max())
    diff = float(np.abs(a - b).max())
    if scale == 0.0:            # all-zero reference: exact agreement is 0, anything else is infinite
        return 0.0 if diff == 0.0 else float('inf')
    return diff / scale
