"""Build libchroma_b200.so in-tree with nvcc for sm_100a.

Usage: python -m chroma_b200.csrc.build [--force]
The shared library is written next to the package (chroma_b200/libchroma_b200.so) so it
travels to the GPU box with the repository snapshot.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
LIB = os.path.join(PKG, 'libchroma_b200.so')
SOURCES = ['engine_core.cu', 'engine_photons.cu', 'engine_propagate.cu', 'engine_wavefront.cu', 'engine_wide.cu', 'engine_wide_ool.cu', 'engine_inst.cu', 'engine_inst_ool.cu', 'engine_daq.cu', 'engine_pdfeval.cu', 'engine_render.cu',
           'bvh_build.cu']
HEADERS = ['common.cuh', 'traverse.cuh', 'wide.cuh', 'inst.cuh', 'wide_prop_kernel.cuh', 'inst_prop_kernel.cuh', 'wide_prop2.cuh', 'physics.cuh', '../../include/chroma_b200.h']
NVCC = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
# --use_fast_math: same numerics switches as the reference build (chroma/gpu/tools.py:12)
FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '--use_fast_math', '-lineinfo',
         '-std=c++17', '-Xcompiler', '-fPIC', '-Xcompiler', '-fvisibility=default']


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, variant=None, extra_flags=()):
    """variant: build an A/B copy libchroma_b200_<variant>.so with `extra_flags` (-D switches) in its
    own object directory; select it at run time with CHROMA_B200_LIB."""
    objdir = os.path.join(HERE, 'build' + ('_' + variant if variant else ''))
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(HERE, h) for h in HEADERS] + [os.path.abspath(__file__)]
    extra = os.environ.get('CHROMA_NVCC_FLAGS', '').split() + list(extra_flags)
    LIB = os.path.join(PKG, 'libchroma_b200%s.so' % ('_' + variant if variant else ''))
    jobs = []
    for src in SOURCES:
        s = os.path.join(HERE, src)
        o = os.path.join(objdir, src.replace('.cu', '.o'))
        if force or _stale(o, [s] + headers):
            jobs.append([NVCC] + FLAGS + extra + ['-c', s, '-o', o])

    def run(cmd):
        if verbose:
            print(' '.join(cmd))
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError('nvcc failed:\n%s\n%s' % (' '.join(cmd), r.stderr))
        return r.stderr

    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 1)) as ex:
            for err in ex.map(run, jobs):
                if verbose and err:
                    print(err)
    objs = [os.path.join(objdir, s.replace('.cu', '.o')) for s in SOURCES]
    if force or jobs or _stale(LIB, objs):
        run([NVCC, '-shared', '-o', LIB] + objs + ['-lcudart'])
    return LIB


if __name__ == '__main__':
    # python -m chroma_b200.csrc.build [--force] [--variant NAME -DFLAG=1 ...]
    args = [a for a in sys.argv[1:] if a != '--force']
    variant, flags = None, []
    if args and args[0] == '--variant':
        variant, flags = args[1], args[2:]
    print(build(force='--force' in sys.argv, verbose=True, variant=variant, extra_flags=flags))
