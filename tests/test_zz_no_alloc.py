"""chroma_intersect_host_static (include/chroma_b200.h): the first-hit entry point for callers that may
not allocate.  It has no counterpart in the reference; it must return exactly what
chroma_intersect_host returns (same kernel, same tree: bit-exact triangles and distances, the tier of
tests/test_gpu_intersect.py), on two calls in a row (the static scratch keeps the last call's
contents), and refuse what does not fit.

Collected last on purpose: a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

import fixtures as F
from chroma_b200 import _lib, gpu

pytestmark = pytest.mark.gpu

MAX_RAYS = 262144            # CHROMA_STATIC_MAX_RAYS


def both(gg, pos, d, last=None):
    n = len(pos)
    pos, d = _lib.carray(pos, np.float32), _lib.carray(d, np.float32)
    last = _lib.carray(last, np.int32) if last is not None else None
    out = []
    for name in ('chroma_intersect_host', 'chroma_intersect_host_static'):
        tri = np.full(n, -7, np.int32)
        dist = np.full(n, np.nan, np.float32)
        _lib.call(name, gg.handle, n, _lib.ptr(pos), _lib.ptr(d), _lib.ptr(last), _lib.ptr(tri), _lib.ptr(dist))
        out.append((tri, dist))
    return out


def test_static_scratch_entry_equals_intersect_host(ctx, oracle_mod):
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    # call 1: a full-size batch, no excluded triangle
    pos, d = F.isotropic_rays(MAX_RAYS, seed=21)
    (t0, d0), (t1, d1) = both(gg, pos, d)
    assert np.array_equal(t0, t1) and np.array_equal(d0.view(np.uint32), d1.view(np.uint32))
    assert (t0 >= 0).mean() > 0.99
    # call 2, straight after, smaller and with the triangle just hit excluded (mesh.h:82): the head,
    # the status word and the staged rays of call 1 are still in the array
    ok = t0 >= 0
    n2 = 100_003
    p2 = (pos + d * d0[:, None])[ok][:n2].astype(np.float32)
    _, dir2 = F.isotropic_rays(n2, seed=22)
    (u0, e0), (u1, e1) = both(gg, p2, dir2, t0[ok][:n2])
    assert np.array_equal(u0, u1) and np.array_equal(e0.view(np.uint32), e1.view(np.uint32))
    assert (u1 != t0[ok][:n2]).all()
    # call 3: one ray; call 4: none
    (v0, f0), (v1, f1) = both(gg, pos[:1], d[:1])
    assert v0[0] == v1[0] == t0[0] and f0[0] == f1[0]
    _lib.call('chroma_intersect_host_static', gg.handle, 0, None, None, None, None, None)


def test_static_scratch_entry_refuses_what_does_not_fit(ctx, oracle_mod):
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    n = MAX_RAYS + 1
    pos, d = np.zeros((n, 3), np.float32), np.ones((n, 3), np.float32)
    tri, dist = np.empty(n, np.int32), np.empty(n, np.float32)
    with pytest.raises(_lib.ChromaError, match='at most 262144 rays'):
        _lib.call('chroma_intersect_host_static', gg.handle, n, _lib.ptr(pos), _lib.ptr(d), None, _lib.ptr(tri), _lib.ptr(dist))
    inst = gpu.GPUDetector(det, instanced=True)
    with pytest.raises(_lib.ChromaError, match='flat geometries'):
        _lib.call('chroma_intersect_host_static', inst.handle, 16, _lib.ptr(pos), _lib.ptr(d), None, _lib.ptr(tri), _lib.ptr(dist))
    # the context is still usable and the entry still works
    p, dd = F.isotropic_rays(1000, seed=3)
    (t0, d0), (t1, d1) = both(gg, p, dd)
    assert np.array_equal(t0, t1) and np.array_equal(d0.view(np.uint32), d1.view(np.uint32))
