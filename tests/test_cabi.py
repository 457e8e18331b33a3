"""CPU tests of the drop-in boundary: libchroma_b200.so loads, exports every symbol
include/chroma_b200.h declares, the ctypes binding covers the same set, and compute calls fail
loudly (no CPU fallback) when no CUDA device is present."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'chroma_b200.h')


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(chroma_[a-z0-9_]+)\s*\(', text)))


def test_header_cites_the_reference_for_every_group():
    text = open(HEADER).read()
    for cite in ('chroma/gpu/photon.py', 'chroma/cuda/propagate.cu', 'chroma/cuda/mesh.h', 'chroma/gpu/daq.py',
                 'chroma/cuda/daq.cu', 'chroma/gpu/pdf.py', 'chroma/cuda/pdf.cu', 'chroma/bvh/grid.py',
                 'chroma/cuda/bvh.cu', 'chroma/gpu/geometry.py', 'chroma/gpu/detector.py', 'chroma/gpu/tools.py'):
        assert cite in text, cite


def test_library_exports_every_declared_symbol():
    from chroma_b200 import _lib
    lib = C.CDLL(_lib.LIB_PATH)
    names = declared_functions()
    assert len(names) > 50
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_covers_the_header():
    from chroma_b200 import _lib
    assert sorted(_lib.EXPORTED_SYMBOLS) == declared_functions()
    lib = _lib.load()
    assert b'sm_100a' in lib.chroma_version()


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip('a CUDA device is present')
    from chroma_b200 import _lib, gpu
    with pytest.raises(_lib.ChromaError) as e:
        gpu.create_cuda_context(0)
    assert e.value.status == -1          # CHROMA_ERR_CUDA
    # null handles are rejected with CHROMA_ERR_ARG, not dereferenced
    lib = _lib.load()
    assert lib.chroma_photons_propagate(None, None, None, 0, 0, 10, 0, 0, None) == -2
    assert lib.chroma_last_error()


def test_product_never_imports_the_oracle():
    "the oracle is test infrastructure: nothing under chroma_b200/ may import or load it"
    pkg = os.path.join(ROOT, 'chroma_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                text = open(os.path.join(dirpath, f)).read()
                assert 'liboracle' not in text and 'from oracle' not in text and 'import oracle' not in text, f


def test_struct_mirrors_match_the_header_layout(tmp_path):
    "ctypes mirrors of chroma_source / chroma_geometry_desc have the size and field offsets the C compiler gives"
    import ctypes as C
    import subprocess
    from chroma_b200 import _lib
    src = tmp_path / 'layout.c'
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "chroma_b200.h"\n'
                   'int main(void) { printf("%zu %zu %zu %zu %zu %zu\\n", sizeof(chroma_source), '
                   'offsetof(chroma_source, rot), offsetof(chroma_source, cos_half_angle), '
                   'offsetof(chroma_source, evidx), offsetof(chroma_source, sequence), '
                   'sizeof(chroma_geometry_desc)); return 0; }\n')
    exe = tmp_path / 'layout'
    subprocess.run(['gcc', '-I', os.path.join(ROOT, 'include'), '-o', str(exe), str(src)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    S = _lib.SourceDesc
    want = [C.sizeof(S), S.rot.offset, S.cos_half_angle.offset, S.evidx.offset, S.sequence.offset,
            C.sizeof(_lib.GeometryDesc)]
    assert got == want
