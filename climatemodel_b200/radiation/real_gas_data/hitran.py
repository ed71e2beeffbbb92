"""Lookup-table location and loading -- the part of ``real_gas_data/hitran.py`` the hot path
touches (hitran.py:23-24 ``LookupTableFolder``; table format written by ``make_table``,
hitran.py:315-356: a pickled dict {'p', 'T', 'nu', 'absorption_coef'} in a ``.npy`` file).

``LookupTableFolder`` may be reassigned (module attribute, like the reference) to point at any
folder holding tables in the reference's ``.npy`` format or as ``.npz`` with the same keys.
"""
import os

import numpy as np

LookupTableFolder = os.environ.get(
    'CLIMATEMODEL_LOOKUP_TABLES',
    os.path.join(os.path.dirname(os.path.abspath(__file__)), 'HitranData', 'LookupTables')) + os.sep

# grids of the reference's own tables (hitran.py:35-37)
table_p_values = np.logspace(np.log10(1.0132e5), np.log10(20), 200)   # p_surface_earth .. p_toa_earth
table_T_values = np.arange(250, 350 + 10, 20)
table_dnu = 10

_cache = {}


def load_table(molecule_name, folder=None):
    """Return the table dict of ``molecule_name``.  Unlike the reference, which re-reads the file
    on every ``load_absorption_coef`` call (real_gas.py:77), tables are cached per path+mtime."""
    folder = LookupTableFolder if folder is None else folder
    for ext in ('.npy', '.npz'):
        path = os.path.join(folder, molecule_name + ext)
        if os.path.exists(path):
            key = (path, os.path.getmtime(path))
            if key not in _cache:
                if ext == '.npy':
                    d = np.load(path, allow_pickle=True).item()
                else:
                    z = np.load(path)
                    d = {k: z[k] for k in z.files}
                _cache[key] = {k: np.asarray(d[k]) for k in ('p', 'T', 'nu', 'absorption_coef')}
            return _cache[key]
    raise FileNotFoundError('no lookup table %s(.npy|.npz) in %s' % (molecule_name, folder))
