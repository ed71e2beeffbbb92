"""Builds libclimatemodel_b200.so in-tree with nvcc for sm_100a (no torch dependency).

    python -m climatemodel_b200.csrc.build [--force] [--verbose]

The library is one translation unit (unity.cu); it is rebuilt when any source is newer than
the .so.  Called by ``__graft_entry__.build()``.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
LIB = os.path.join(PKG, 'libclimatemodel_b200.so')
SOURCES = ['unity.cu', 'plan.cu', 'tables.cu', 'flux.cu', 'transmission.cu', 'peaks.cu', 'lines.cu',
           'convect.cu', 'stepper.cu', 'grey.cu', 'noalloc.cu', 'common.cuh', 'internal.cuh', 'flux_kernels.cuh', 'tmem.cuh',
           os.path.join('..', '..', 'include', 'climatemodel_b200.h')]
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
              '--expt-relaxed-constexpr', '--expt-extended-lambda',
              '-Xcompiler', '-fPIC', '-shared']


def nvcc_path():
    for cand in (os.environ.get('NVCC'), shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError('nvcc not found')


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(HERE, s)) > t for s in SOURCES)


def build(force=False, verbose=False):
    if not force and not stale():
        return LIB
    extra = []
    experiments = [v for v in os.environ.get('CMB200_EXPERIMENTS', '').split(',') if v.strip()]
    if experiments:      # extra flux kernel variants for tools/ab_flux_variants.py (flux.cu)
        extra.append('-DCMB_FLUX_EXPERIMENTS(X)=' + ' '.join('X(%d)' % int(v) for v in experiments))
    if os.environ.get('CMB200_DEV'):      # kernel development: bench shapes only (flux.cu)
        extra.append('-DCMB_DEV_BUILD')
    cmd = [nvcc_path()] + NVCC_FLAGS + extra + (['-Xptxas', '-v'] if verbose else []) + \
          ['-o', LIB, os.path.join(HERE, 'unity.cu')]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError('nvcc failed building %s' % LIB)
    if verbose:
        sys.stderr.write(r.stdout + r.stderr)
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='--verbose' in sys.argv))
