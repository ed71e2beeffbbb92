"""TEST / BASELINE INFRASTRUCTURE ONLY -- never imported by the product package.

Stages the UNMODIFIED reference package for the CPU baseline: copies the Python files of
``/root/reference/Model`` that the real-gas and grey paths import (constants, radiation,
real_gas_data, the three shipped lookup tables) to ``oracle/_ref/Model``.  ``oracle/_ref/`` is
git-ignored (reference sources are never committed) but travels to the GPU box with the snapshot,
like a built ``.so``; there ``bench.py --impl reference`` and the ``cpu_baseline`` leg import it
through ``oracle/reference_shim.py`` (stub ``sympl`` constants and an empty ``matplotlib``) and time
the reference's own ``optical_depth`` + ``RealGas.get_flux`` per column (``oracle/ref_runner.py``).

    python oracle/make_ref.py            # build container only: needs /root/reference

Nothing is modified: the files are byte-for-byte copies (the recipe prints their count and a
SHA-256 over the concatenation so that a run can say which reference it timed).
"""
import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get('CLIMATEMODEL_REFERENCE_SRC', '/root/reference')
DST = os.path.join(HERE, '_ref')

WANTED = [
    'Model/__init__.py',
    'Model/constants.py',
    'Model/radiation/base.py',
    'Model/radiation/convective_adjustment.py',
    'Model/radiation/grey.py',
    'Model/radiation/grey_optical_depth.py',
    'Model/radiation/real_gas.py',
    'Model/radiation/animation.py',
    'Model/radiation/real_gas_data/hitran.py',
    'Model/radiation/real_gas_data/specific_humidity.py',
    'Model/radiation/real_gas_data/temperature_profiles.py',
    'Model/radiation/real_gas_data/HitranData/LookupTables/CO2_no_sw.npy',
    'Model/radiation/real_gas_data/HitranData/LookupTables/gray.npy',
    'Model/radiation/real_gas_data/HitranData/LookupTables/single_line.npy',
]


def stage(verbose=True):
    if not os.path.isdir(os.path.join(SRC, 'Model', 'radiation')):
        raise RuntimeError('reference tree not present at %s' % SRC)
    digest = hashlib.sha256()
    for rel in WANTED:
        src, dst = os.path.join(SRC, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        os.chmod(dst, 0o644)
        with open(src, 'rb') as f:
            digest.update(f.read())
    for pkg in ('Model/radiation', 'Model/radiation/real_gas_data', 'Model/dynamics'):
        init = os.path.join(SRC, pkg, '__init__.py')
        if os.path.exists(init):
            os.makedirs(os.path.join(DST, pkg), exist_ok=True)
            shutil.copyfile(init, os.path.join(DST, pkg, '__init__.py'))
    with open(os.path.join(DST, 'STAGED.txt'), 'w') as f:
        f.write('unmodified copy of %d files of %s\nsha256 %s\n' % (len(WANTED), SRC, digest.hexdigest()))
    if verbose:
        print('staged %d reference files under %s (sha256 %s)' % (len(WANTED), DST, digest.hexdigest()[:16]))
    return DST


if __name__ == '__main__':
    stage()
    sys.exit(0)
