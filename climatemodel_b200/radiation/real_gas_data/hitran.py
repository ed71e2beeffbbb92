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
    on every ``load_absorption_coef`` call (real_gas.py:77), tables are cached per path, mtime and size."""
    folder = LookupTableFolder if folder is None else folder
    for ext in ('.npy', '.npz'):
        path = os.path.join(folder, molecule_name + ext)
        if os.path.exists(path):
            st = os.stat(path)
            key = (path, st.st_mtime_ns, st.st_size)
            if key not in _cache:
                if ext == '.npy':
                    d = np.load(path, allow_pickle=True).item()
                else:
                    z = np.load(path)
                    d = {k: z[k] for k in z.files}
                _cache[key] = {k: np.asarray(d[k]) for k in ('p', 'T', 'nu', 'absorption_coef')}
            return _cache[key]
    raise FileNotFoundError('no lookup table %s(.npy|.npz) in %s -- point hitran.LookupTableFolder (or the environment '
                            'variable CLIMATEMODEL_LOOKUP_TABLES) at a folder of tables in the reference\'s format; the '
                            'reference\'s three shipped tables are kept as test fixtures under tests/golden/tables'
                            % (molecule_name, folder))


# ---- building tables: line-by-line absorption coefficient (hitran.py:122-247, 315-356) -----------------
# Reference conditions of the HITRAN line parameters (hitran.py:29-30).
from ...constants import Avogadro, p_one_atmosphere, speed_of_light, h_planck, k_boltzmann   # noqa: E402

p_reference = p_one_atmosphere
T_reference = 296
data_folder = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'HitranData') + os.sep
required_fields = ['molec_id', 'local_iso_id', 'nu', 'sw', 'elower', 'gamma_air', 'n_air']


def s_conversion(s, M):
    """Line intensity from cm^-1/(molec cm^-2) to cm^-1 m^2/kg for molar mass M (g/mol)."""
    return 0.1 * Avogadro / M * s


def gamma_extrapolate(p, T, gamma_reference, n):
    """Air-broadened half-width at (p, T) from its value at (p_reference, T_reference)."""
    return gamma_reference * (p / p_reference) * (T_reference / T) ** n


def s_extrapolate(T, s_reference, wave_number_line_center, n):
    """Line strength at T from its value at T_reference (no partition function, like the reference)."""
    freq = 100 * wave_number_line_center * speed_of_light
    return s_reference * (T / T_reference) ** n * np.exp(-(h_planck * freq / k_boltzmann) * (1 / T - 1 / T_reference))


def lorentzian_profile(wave_number_array, wave_number_line_center, gamma):
    return (1 / np.pi) * gamma / (gamma ** 2 + (wave_number_array - wave_number_line_center) ** 2)


def wavenumbers_near_line(wavenumber_array, array_spacing, wavenumber_line_center, line_width, n_line_widths):
    """Index range [i1, i2) of the grid within n_line_widths half-widths of the line centre."""
    centre = int(np.abs(wavenumber_array - wavenumber_line_center).argmin())
    reach = int(n_line_widths * line_width / array_spacing)
    return max(0, centre - reach), min(np.size(wavenumber_array) - 1, centre + reach) + 1


def load_molecule_data(molecule_name):
    """Main-isotopologue lines of a HITRAN text export in ``data_folder`` as a dict of arrays
    ('nu', 'sw' in m^2/kg units, 'elower', 'gamma_air', 'n_air')."""
    from .specific_humidity import molecules
    raw = np.genfromtxt(os.path.join(data_folder, molecule_name + '.txt'), names=True)
    keep = np.where(np.logical_and(raw['molec_id'] == molecules[molecule_name]['hitran_id'],
                                   raw['local_iso_id'] == 1))
    raw = raw[keep]
    data = {field: raw[field] for field in required_fields[2:]}
    data['sw'] = s_conversion(data['sw'], molecules[molecule_name]['M'])
    return data


def update_molecule_data(molecule_data, wavenumber_array):
    """Only the lines inside the wavenumber grid."""
    keep = np.where(np.logical_and(molecule_data['nu'] >= wavenumber_array.min(),
                                   molecule_data['nu'] <= wavenumber_array.max()))
    return {key: molecule_data[key][keep] for key in molecule_data}


def get_absorption_coefficient(p, T, wavenumber_array, molecule_name, molecule_data=None, n_line_widths=1000):
    """Absorption coefficient [n_p x n_nu] (m^2/kg): all lines of a molecule summed on the GPU
    (``cmb200_line_absorption_host``).  Same arguments as the reference; ``molecule_data`` is the
    dict of line arrays ('nu', 'sw', 'gamma_air', 'n_air'), loaded from file when None."""
    from ... import _lib
    if molecule_data is None:
        molecule_data = load_molecule_data(molecule_name)
    p, T, nu = _lib.as_f64(p).reshape(-1), _lib.as_f64(T).reshape(-1), _lib.as_f64(wavenumber_array).reshape(-1)
    if p.shape != T.shape:
        raise ValueError('p and T must have the same size')
    lines = [_lib.as_f64(molecule_data[k]).reshape(-1) for k in ('nu', 'sw', 'gamma_air', 'n_air')]
    if len({a.size for a in lines}) != 1:
        raise ValueError('line arrays must have the same length')
    out = np.zeros((p.size, nu.size))
    lib = _lib.load()
    cst = _lib.constants_struct()
    _lib.check(lib.cmb200_line_absorption_host(cst, float(p_reference), float(T_reference), p.size, _lib.dptr(p),
                                               _lib.dptr(T), nu.size, _lib.dptr(nu), lines[0].size,
                                               *[_lib.dptr(a) for a in lines], float(n_line_widths), _lib.dptr(out)))
    return out


def make_table(molecule_name, p_array=table_p_values, T_array=table_T_values, dwavenumber=table_dnu,
               n_line_widths=1000, wavenumber_array=None, save=True):
    """[n_p x n_T x n_nu] lookup table of a molecule (or of a dict of line arrays, saved as
    'custom'), written to ``LookupTableFolder`` in the reference's format.  The wavenumber grid
    must be given (the reference's automatic range finder and the O3 / CFC12 cross-section
    files are not part of this package)."""
    if isinstance(molecule_name, dict):
        molecule_data, molecule_name = molecule_name, 'custom'
    else:
        molecule_data = load_molecule_data(molecule_name)
    if wavenumber_array is None:
        raise ValueError('wavenumber_array must be specified')
    output_file = os.path.join(LookupTableFolder, molecule_name + '.npy')
    if save and os.path.isfile(output_file):
        raise ValueError('Lookuptable file already exists')
    if molecule_name != 'custom':
        molecule_data = update_molecule_data(molecule_data, wavenumber_array)
    grid = np.zeros((np.size(p_array), np.size(T_array), np.size(wavenumber_array)))
    for i in range(np.size(T_array)):
        grid[:, i, :] = get_absorption_coefficient(p_array, np.ones_like(p_array, dtype=float) * T_array[i],
                                                   wavenumber_array, molecule_name, molecule_data, n_line_widths)
    table = {'p': np.asarray(p_array), 'T': np.asarray(T_array), 'nu': np.asarray(wavenumber_array),
             'absorption_coef': grid}
    if save:
        np.save(output_file, table)
    return table
