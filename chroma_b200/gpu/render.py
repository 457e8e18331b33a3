"""GPURays: ray positions and directions on the device and the headless alpha-blended ray cast
of a geometry (API of ``chroma/gpu/render.py``).  The window / movie loop of the reference's
camera (pygame) is not part of this package: ``snapshot()`` returns the ARGB pixel array."""
import ctypes as C

import numpy as np

from chroma_b200 import _lib
from chroma_b200.gpu.tools import DeviceArray, current_context, float3, float4, empty


def _vec3(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.float32).reshape(3))


class GPURays(object):
    def __init__(self, pos, dir, max_alpha_depth=10, nblocks=64):
        self.ctx = current_context()
        pos = np.ascontiguousarray(np.asarray(pos, dtype=np.float32).reshape(-1, 3))
        dir = np.ascontiguousarray(np.asarray(dir, dtype=np.float32).reshape(-1, 3))
        if pos.shape != dir.shape:
            raise ValueError('pos and dir must have the same shape')
        n = len(pos)
        h = C.c_void_p()
        _lib.call('chroma_rays_create', self.ctx.handle, n, _lib.ptr(pos), _lib.ptr(dir),
                  int(max_alpha_depth), C.byref(h))
        self.handle = h
        self.max_alpha_depth = max_alpha_depth
        self.nblocks = nblocks
        p, d, dx, col, dxlen = (C.c_void_p() for _ in range(5))
        cnt = C.c_uint64()
        _lib.call('chroma_rays_arrays', self.handle, C.byref(p), C.byref(d), C.byref(dx),
                  C.byref(col), C.byref(dxlen), C.byref(cnt))
        self.pos = DeviceArray(n, float3, ctx=self.ctx, ptr=p.value, base=self)
        self.dir = DeviceArray(n, float3, ctx=self.ctx, ptr=d.value, base=self)
        self.dx = DeviceArray(n * max_alpha_depth, np.float32, ctx=self.ctx, ptr=dx.value, base=self)
        self.color = DeviceArray(n * max_alpha_depth, float4, ctx=self.ctx, ptr=col.value, base=self)
        self.dxlen = DeviceArray(n, np.uint32, ctx=self.ctx, ptr=dxlen.value, base=self)

    def rotate(self, phi, n):
        "Rotate by an angle phi around the axis `n`."
        _lib.call('chroma_rays_rotate', self.handle, float(phi), _lib.ptr(_vec3(n)))

    def rotate_around_point(self, phi, n, point):
        "Rotate by an angle phi around the axis `n` passing through the point `point`."
        _lib.call('chroma_rays_rotate_around_point', self.handle, float(phi), _lib.ptr(_vec3(n)),
                  _lib.ptr(_vec3(point)))

    def translate(self, v):
        "Translate the ray positions by the vector `v`."
        _lib.call('chroma_rays_translate', self.handle, _lib.ptr(_vec3(v)))

    def render(self, gpu_geometry, pixels, alpha_depth=10, keep_last_render=False,
               bg_color=0x00000000):
        "Render `gpu_geometry` and fill the device array `pixels` with pixel colors."
        if alpha_depth > self.max_alpha_depth:
            raise Exception('alpha_depth > max_alpha_depth')
        if not isinstance(pixels, DeviceArray):
            raise TypeError('`pixels` must be a %s instance.' % DeviceArray)
        if pixels.size != self.pos.size:
            raise ValueError('`pixels`.size != number of rays')
        _lib.call('chroma_rays_render', self.handle, gpu_geometry.handle, int(alpha_depth),
                  int(bool(keep_last_render)), int(bg_color) & 0xFFFFFFFF, C.c_void_p(pixels.ptr))

    def snapshot(self, gpu_geometry, alpha_depth=10):
        "Render `gpu_geometry` and return a numpy array of pixel colors."
        pixels = empty(self.pos.size, np.uint32, ctx=self.ctx)
        self.render(gpu_geometry, pixels, alpha_depth)
        return pixels.get()

    def __del__(self):
        try:
            if self.handle is not None and self.ctx.handle is not None:
                _lib.call('chroma_rays_destroy', self.handle)
            self.handle = None
        except Exception:
            pass
