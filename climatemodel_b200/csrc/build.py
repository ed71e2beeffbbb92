"""Builds libclimatemodel_b200.so in-tree with nvcc for sm_100a (no torch dependency).

    python -m climatemodel_b200.csrc.build [--force] [--verbose]

The library is unity.cu plus ten translation units of flux kernel instantiations (flux_slice.cu), compiled in parallel;
it is rebuilt unless the stamp next to the .so matches the sources (content hash), flags and build mode.  Called by ``__graft_entry__.build()``.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
LIB = os.path.join(PKG, 'libclimatemodel_b200.so')
STAMP = LIB + '.stamp'
SOURCES = ['unity.cu', 'plan.cu', 'tables.cu', 'flux.cu', 'flux_slice.cu', 'transmission.cu', 'peaks.cu', 'lines.cu',
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


def source_stamp():
    """Hash of what the library is built from: every source, the flags, the build mode."""
    import hashlib
    h = hashlib.sha256()
    for name in SOURCES:
        with open(os.path.join(HERE, name), 'rb') as f:
            h.update(name.encode() + b'\0' + f.read())
    h.update(repr((NVCC_FLAGS, os.environ.get('CMB200_EXPERIMENTS', ''), bool(os.environ.get('CMB200_DEV')))).encode())
    return h.hexdigest()


def stale():
    """The library is rebuilt unless the stamp written next to it says it was built from exactly these sources in
    this mode (so a development build, or one older than an edit or a checkout, is never mistaken for the library)."""
    if not os.path.exists(LIB) or not os.path.exists(STAMP):
        return True
    with open(STAMP) as f:
        return f.read().strip() != source_stamp()


def run_nvcc(cmd, verbose):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError('nvcc failed: %s' % ' '.join(cmd[-3:]))
    if verbose:
        sys.stderr.write(r.stdout + r.stderr)


def build(force=False, verbose=False):
    """unity.cu (everything but the flux kernel instantiations) and one flux_slice.cu per (precision, molecule count)
    are compiled in parallel to objects under csrc/_build/, then linked.  CMB200_DEV=1: the bench shapes only."""
    if not force and not stale():
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    extra = []
    experiments = [v for v in os.environ.get('CMB200_EXPERIMENTS', '').split(',') if v.strip()]
    if experiments:      # extra flux kernel variants for tools/ab_flux_variants.py (flux.cu)
        extra.append('-DCMB_FLUX_EXPERIMENTS(X)=' + ' '.join('X(%d)' % int(v) for v in experiments))
    dev = bool(os.environ.get('CMB200_DEV'))
    if dev:              # kernel development: bench shapes only (flux.cu)
        extra.append('-DCMB_DEV_BUILD')
    slices = [('double', 0), ('double', 2), ('float', 2)] if dev else [(r, m) for r in ('double', 'float') for m in range(5)]
    obj_dir = os.path.join(HERE, '_build')
    os.makedirs(obj_dir, exist_ok=True)
    flags = [f for f in NVCC_FLAGS if f != '-shared'] + extra + (['-Xptxas', '-v'] if verbose else [])
    jobs = [[nvcc_path()] + flags + ['-DCMB_FLUX_SLICES', '-c', '-o', os.path.join(obj_dir, 'unity.o'), os.path.join(HERE, 'unity.cu')]]
    for real, m in slices:
        name = 'launch_flux_slice_%s%d' % (real[0], m)
        jobs.append([nvcc_path()] + flags + ['-DCMB_SLICE_REAL=' + real, '-DCMB_SLICE_M=%d' % m, '-DCMB_SLICE_NAME=' + name,
                                            '-c', '-o', os.path.join(obj_dir, name + '.o'), os.path.join(HERE, 'flux_slice.cu')])
    # the big slices first
    jobs.sort(key=lambda j: 0 if 'M=2' in ' '.join(j) else 1)
    with ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 1))) as pool:
        list(pool.map(lambda j: run_nvcc(j, verbose), jobs))
    objs = [j[j.index('-o') + 1] for j in jobs]
    if os.path.exists(STAMP):
        os.remove(STAMP)
    run_nvcc([nvcc_path(), '-gencode', 'arch=compute_100a,code=sm_100a', '-shared', '-Xcompiler', '-fPIC', '-o', LIB] + objs, verbose)
    with open(STAMP, 'w') as f:
        f.write(source_stamp() + '\n')
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='--verbose' in sys.argv))
