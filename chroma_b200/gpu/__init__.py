"""Device layer of chroma_b200: same public names as ``chroma.gpu`` (chroma/gpu/__init__.py)."""
import ctypes as _C

import numpy as _np

from chroma_b200 import _lib
from chroma_b200.gpu.tools import *  # noqa: F401,F403
from chroma_b200.gpu.tools import DeviceArray, current_context
from chroma_b200.gpu.geometry import GPUGeometry
from chroma_b200.gpu.detector import GPUDetector
from chroma_b200.gpu.photon import GPUPhotons, GPUPhotonsSlice
from chroma_b200.gpu.daq import GPUDaq, GPUChannels
from chroma_b200.gpu.pdf import GPUPDF, GPUKernelPDF
from chroma_b200.gpu.render import GPURays


def intersect(gpu_geometry, pos, dir, last_hit_triangles=None, counts=False):
    """First-hit triangle index and distance for rays (the reference's ``distance_to_mesh``
    kernel, chroma/cuda/mesh.h:123-151, plus the triangle index it drops).

    `pos`, `dir`: (n,3) host arrays, or DeviceArray of float32 with 3n elements.  Returns
    (triangle int32[n], distance float32[n]) as numpy arrays for host input, DeviceArrays for
    device input.  With ``counts=True`` (device input) also returns the per-ray
    (node visits, triangle tests) of the reference traversal."""
    ctx = gpu_geometry.ctx
    if isinstance(pos, DeviceArray):
        n = pos.size // 3 if pos.dtype == _np.float32 else pos.size
        tri = DeviceArray(n, _np.int32, ctx=ctx)
        dist = DeviceArray(n, _np.float32, ctx=ctx)
        last = _C.c_void_p(last_hit_triangles.ptr) if last_hit_triangles is not None else None
        if counts:
            cnt = DeviceArray((n, 2), _np.uint32, ctx=ctx)
            _lib.call('chroma_intersect_count_device', gpu_geometry.handle, n, _C.c_void_p(pos.ptr),
                      _C.c_void_p(dir.ptr), last, _C.c_void_p(tri.ptr), _C.c_void_p(dist.ptr),
                      _C.c_void_p(cnt.ptr))
            return tri, dist, cnt
        _lib.call('chroma_intersect_device', gpu_geometry.handle, n, _C.c_void_p(pos.ptr),
                  _C.c_void_p(dir.ptr), last, _C.c_void_p(tri.ptr), _C.c_void_p(dist.ptr))
        return tri, dist
    pos = _lib.carray(pos, _np.float32)
    dir = _lib.carray(dir, _np.float32)
    n = len(pos)
    last = _lib.carray(last_hit_triangles, _np.int32) if last_hit_triangles is not None else None
    tri = _np.empty(n, dtype=_np.int32)
    dist = _np.empty(n, dtype=_np.float32)
    _lib.call('chroma_intersect_host', gpu_geometry.handle, n, _lib.ptr(pos), _lib.ptr(dir),
              _lib.ptr(last), _lib.ptr(tri), _lib.ptr(dist))
    return tri, dist
