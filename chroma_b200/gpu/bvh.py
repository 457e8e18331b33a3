"""BVH construction helpers on the device (API of ``chroma/gpu/bvh.py``)."""
import ctypes as C

import numpy as np

from chroma_b200 import _lib
from chroma_b200.gpu.tools import current_context, uint4


def world_coords_for(mesh):
    "WorldCoords spanning the mesh with 2**16-2 quanta on the longest axis (gpu/bvh.py:43-48)."
    from chroma_b200.bvh.bvh import WorldCoords
    vertices = np.asarray(mesh.vertices, dtype=np.float32)
    world_origin = vertices.min(axis=0)
    world_scale = np.float32(np.max(vertices.max(axis=0) - world_origin)) / np.float32(2 ** 16 - 2)
    return WorldCoords(world_origin=world_origin, world_scale=world_scale)


def create_leaf_nodes(mesh, morton_bits=16, round_to_multiple=1):
    """(world_coords, leaf nodes uint4[ntriangles], Morton codes uint64[ntriangles]) for a
    mesh: each triangle boxed by one leaf (chroma/cuda/bvh.cu:148-203)."""
    ctx = current_context()
    wc = world_coords_for(mesh)
    vertices = _lib.carray(mesh.vertices, np.float32)
    triangles = _lib.carray(mesh.triangles, np.uint32)
    nt = len(triangles)
    padded = -(-nt // round_to_multiple) * round_to_multiple
    nodes = np.zeros(padded, dtype=uint4)
    codes = np.empty(nt, dtype=np.uint64)
    origin = (C.c_float * 3)(*[float(x) for x in wc.world_origin])
    _lib.call('chroma_bvh_make_leaves', ctx.handle, _lib.ptr(vertices), len(vertices),
              _lib.ptr(triangles), nt, origin, float(wc.world_scale),
              nodes[:nt].ctypes.data_as(C.c_void_p), _lib.ptr(codes))
    return wc, nodes, codes >> np.uint64(16 - morton_bits)


def merge_nodes_detailed(nodes, first_child, nchild):
    "Parent boxes for explicit (first_child, nchild) groups (chroma/cuda/bvh.cu:269-308)."
    ctx = current_context()
    nodes = np.ascontiguousarray(nodes, dtype=uint4)
    fc = _lib.carray(first_child, np.int32)
    nc = _lib.carray(nchild, np.int32)
    parents = np.empty(len(fc), dtype=uint4)
    _lib.call('chroma_bvh_merge_nodes_detailed', ctx.handle, nodes.ctypes.data_as(C.c_void_p),
              len(nodes), _lib.ptr(fc), _lib.ptr(nc), len(fc), parents.ctypes.data_as(C.c_void_p))
    return parents


def merge_nodes(nodes, degree, max_ratio=None):
    "Fixed-degree parents: node i covers children [i*degree, (i+1)*degree)."
    nparent = -(-len(nodes) // degree)
    first = np.arange(nparent, dtype=np.int32) * degree
    nchild = np.minimum(degree, len(nodes) - first).astype(np.int32)
    return merge_nodes_detailed(nodes, first, nchild)


def concatenate_layers(layers):
    """Stack layers root-first and turn per-layer child indices into global ones
    (chroma/gpu/bvh.py:239-267).  Returns (nodes, layer_bounds)."""
    layer_bounds = np.insert(np.cumsum([len(x) for x in layers]), 0, 0)
    nodes = np.empty(int(layer_bounds[-1]), dtype=uint4)
    for start, end, layer in zip(layer_bounds[:-1], layer_bounds[1:], layers):
        out = np.array(layer, dtype=uint4, copy=True)
        if end != layer_bounds[-1]:  # leaf layer keeps its triangle ids
            nchild = out['w'] >> np.uint32(28)
            child = (out['w'] & np.uint32(0x0FFFFFFF)) + np.uint32(end)
            out['w'] = (nchild << np.uint32(28)) | child
        nodes[start:end] = out
    return nodes, layer_bounds


def collapse_chains(nodes, layer_bounds):
    "Replace every single-child node by its child, deepest layers first (bvh.cu:530-543)."
    ctx = current_context()
    nodes = np.array(nodes, dtype=uint4, copy=True)
    bounds = _lib.carray(layer_bounds, np.uint64)
    _lib.call('chroma_bvh_collapse_chains', ctx.handle, nodes.ctypes.data_as(C.c_void_p), len(nodes),
              _lib.ptr(bounds), len(bounds))
    return nodes
