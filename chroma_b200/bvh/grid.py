"""Recursive-grid BVH (API of ``chroma/bvh/grid.py:11-95``), built entirely on the device by
``chroma_bvh_build``: Morton sort, grouping, parent boxes, concatenation and chain collapse
never leave HBM.  Ties between equal Morton codes keep triangle order (stable radix sort);
the reference's numpy ``argsort`` leaves that order unspecified."""
import ctypes as C

import numpy as np

from chroma_b200 import _lib
from chroma_b200.bvh.bvh import BVH, WorldCoords, CHILD_BITS
from chroma_b200.gpu.tools import current_context, uint4

MAX_CHILD = 2 ** (32 - CHILD_BITS) - 1


class _DeviceBVH(object):
    "Owns a chroma_bvh handle so GPUGeometry can take the nodes without a host round trip."

    def __init__(self, ctx, handle):
        self.ctx, self.handle = ctx, handle

    def __del__(self):
        try:
            if self.handle is not None and self.ctx.handle is not None:
                _lib.call('chroma_bvh_destroy', self.handle)
                self.handle = None
        except Exception:
            pass


def _wrap(ctx, h, fetch_nodes, keep_on_device, verbose=False):
    dev = _DeviceBVH(ctx, h)
    info = (C.c_uint64 * 2)()
    origin = (C.c_float * 3)()
    scale = C.c_float()
    _lib.call('chroma_bvh_info', h, info, origin, C.byref(scale))
    nnodes, nlayers = info[0], info[1]
    offsets = np.empty(nlayers, dtype=np.uint64)
    nodes = np.empty(nnodes, dtype=uint4) if fetch_nodes else None
    _lib.call('chroma_bvh_get', h, nodes.ctypes.data_as(C.c_void_p) if fetch_nodes else None,
              _lib.ptr(offsets))
    if verbose:
        print('BVH: %d nodes in %d layers' % (nnodes, nlayers))
    bvh = BVH(WorldCoords(np.array(list(origin), dtype=np.float32), np.float32(scale.value)),
              nodes if fetch_nodes else np.empty(0, dtype=uint4), [int(x) for x in offsets])
    if not fetch_nodes:
        bvh.layer_bounds = [int(x) for x in offsets] + [int(nnodes)]
        bvh.nodes = None
    if keep_on_device:
        bvh.device_handle = dev.handle
        bvh._device_owner = dev
    return bvh


def make_recursive_grid_bvh_from_leaves(world_coords, leaf_nodes, morton_codes, target_degree=3,
                                        keep_on_device=False):
    """The part of ``chroma/bvh/grid.py:24-95`` after ``create_leaf_nodes``: Morton sort (stable),
    grouping, parents, concatenation and chain collapse on given leaves, on the device."""
    ctx = current_context()
    leaves = np.ascontiguousarray(leaf_nodes).view(np.uint32).reshape(-1, 4)
    codes = _lib.carray(morton_codes, np.uint64)
    origin = (C.c_float * 3)(*[float(x) for x in world_coords.world_origin])
    h = C.c_void_p()
    _lib.call('chroma_bvh_build_from_leaves', ctx.handle, _lib.ptr(leaves), _lib.ptr(codes), len(codes),
              origin, float(world_coords.world_scale), int(target_degree), C.byref(h))
    return _wrap(ctx, h, True, keep_on_device)


def make_recursive_grid_bvh(mesh, target_degree=3, verbose=False, keep_on_device=True,
                            fetch_nodes=True):
    """BVH of `mesh` with about `target_degree` children per node (at most 15).

    keep_on_device: keep the device copy alive on the returned BVH (``bvh.device_handle``) so
    GPUGeometry reuses it.  fetch_nodes=False skips the D2H copy of the node array (for very
    large meshes that only need the device copy); ``bvh.nodes`` is then None."""
    ctx = current_context()
    vertices = _lib.carray(mesh.vertices, np.float32)
    triangles = _lib.carray(mesh.triangles, np.uint32)
    h = C.c_void_p()
    _lib.call('chroma_bvh_build', ctx.handle, _lib.ptr(vertices), len(vertices),
              _lib.ptr(triangles), len(triangles), int(target_degree), C.byref(h))
    return _wrap(ctx, h, fetch_nodes, keep_on_device, verbose)
