"""Generate tests/golden/bvh_golden.npz by running the REFERENCE's own recursive-grid builder,
chroma/bvh/grid.py:11-95 (imported from /root/reference, build container only), on fixed leaves.

grid.py is plain numpy around four helpers of chroma/gpu/bvh.py, which are PyCUDA launches of
chroma/cuda/bvh.cu kernels.  Here (no GPU, no PyCUDA) the module `chroma.gpu.bvh` is a stand-in:

  create_leaf_nodes      returns leaves / Morton codes made beforehand (CPU restatement of
                         make_leaves, bvh.cu:148-203; the leaves are INPUT to what is pinned here
                         and are stored in the golden file)
  merge_nodes_detailed   integer numpy min / max per (first_child, nchild) group
                         (make_parents_detailed, bvh.cu:269-308)
  concatenate_layers     gpu/bvh.py:239-267 with copy_and_offset (bvh.cu:364-384) as numpy
  collapse_chains        gpu/bvh.py:114-130 with collapse_child (bvh.cu:530-543) as numpy

The three numpy helpers are themselves checked against the reference's bvh.cubin on the GPU box
(tests/test_gpu_bvh.py::test_reference_kernels_on_golden_layers): the per-layer
(first_child, nchild) arrays grid.py handed to merge_nodes_detailed are recorded here for that.

One documented choice: grid.py sorts with `morton_codes.argsort()` (numpy's default, unstable:
the order of equal codes is unspecified and differs between numpy versions).  The codes are
returned as an ndarray subclass whose argsort is stable, the order the device builder (stable
radix sort) produces; every other line of grid.py runs unmodified.

Usage: python tests/golden/make_bvh_golden.py
"""
import hashlib
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
REF = '/root/reference'
OUT = os.path.join(HERE, 'bvh_golden.npz')

UINT4 = np.dtype([('x', np.uint32), ('y', np.uint32), ('z', np.uint32), ('w', np.uint32)])
CHILD_BITS = 28
CHILD_MASK = np.uint32((1 << CHILD_BITS) - 1)


class StableCodes(np.ndarray):
    "uint64 Morton codes whose argsort() keeps the order of equal codes"

    def argsort(self, *a, **kw):
        kw['kind'] = 'stable'
        return np.asarray(self).argsort(*a, **kw)


# ---- numpy stand-ins of the three bvh.cu kernels grid.py drives ---------------------------------
def np_merge_nodes_detailed(nodes, first_child, nchild):
    "make_parents_detailed, bvh.cu:269-308: union of the children's 16-bit boxes"
    nodes = np.asarray(nodes)
    fc = np.asarray(first_child).astype(np.int64)
    nc = np.asarray(nchild).astype(np.int64)
    out = np.zeros(len(fc), dtype=UINT4)
    for axis in 'xyz':
        lo = (nodes[axis] & np.uint32(0xFFFF)).astype(np.uint32)
        hi = (nodes[axis] >> np.uint32(16)).astype(np.uint32)
        out[axis] = np.minimum.reduceat(lo, fc) | (np.maximum.reduceat(hi, fc) << np.uint32(16))
        # reduceat runs to the next first_child: identical to [first, first+nchild) because the
        # groups grid.py builds tile the layer
    assert np.array_equal(fc[1:], (fc + nc)[:-1]) and fc[-1] + nc[-1] == len(nodes)
    out['w'] = (nc.astype(np.uint32) << np.uint32(CHILD_BITS)) | fc.astype(np.uint32)
    return out


def np_concatenate_layers(layers):
    "gpu/bvh.py:239-267: root first; children of every layer but the last are offset by layer end"
    bounds = np.insert(np.cumsum([len(x) for x in layers]), 0, 0)
    nodes = np.empty(int(bounds[-1]), dtype=UINT4)
    for start, end, layer in zip(bounds[:-1], bounds[1:], layers):
        out = np.array(layer, dtype=UINT4, copy=True)
        if end != bounds[-1]:      # copy_and_offset (bvh.cu:364-384); the leaf layer is memcpy'd
            nchild = out['w'] >> np.uint32(CHILD_BITS)
            child = (out['w'] & CHILD_MASK) + np.uint32(end)
            out['w'] = (nchild << np.uint32(CHILD_BITS)) | child
        nodes[start:end] = out
    return nodes, bounds


def np_collapse_chains(nodes, layer_bounds):
    "gpu/bvh.py:114-130 + collapse_child (bvh.cu:530-543): deepest non-leaf layer first"
    nodes = np.array(nodes, dtype=UINT4, copy=True)
    bounds = list(zip(layer_bounds[:-1], layer_bounds[1:]))[:-1]
    bounds.reverse()
    for start, end in bounds:
        sl = nodes[start:end]
        single = (sl['w'] >> np.uint32(CHILD_BITS)) == 1
        child = (sl['w'] & CHILD_MASK)[single]
        sl[single] = nodes[child]
    return nodes


def import_reference_grid(record):
    "chroma.bvh.grid from /root/reference with chroma.gpu.bvh replaced by the stand-ins above"
    if not hasattr(np, 'asfarray'):
        np.asfarray = lambda a, dtype=np.float64: np.asarray(a, dtype=dtype)
    pkg = types.ModuleType('chroma')
    pkg.__path__ = [os.path.join(REF, 'chroma')]
    sys.modules['chroma'] = pkg
    pycuda = types.ModuleType('pycuda')
    ga = types.ModuleType('pycuda.gpuarray')
    ga.vec = types.SimpleNamespace(uint4=UINT4)
    pycuda.gpuarray = ga
    sys.modules['pycuda'], sys.modules['pycuda.gpuarray'] = pycuda, ga
    bpkg = types.ModuleType('chroma.bvh')
    bpkg.__path__ = [os.path.join(REF, 'chroma', 'bvh')]
    sys.modules['chroma.bvh'] = bpkg
    src = open(os.path.join(REF, 'chroma', 'bvh', 'bvh.py')).read()
    # numpy 2 rejects the overflowing literal numpy 1 wrapped (same value)
    src = src.replace('np.uint32(0xFFFF << CHILD_BITS)', 'np.uint32((0xFFFF << CHILD_BITS) & 0xFFFFFFFF)')
    bmod = types.ModuleType('chroma.bvh.bvh')
    exec(compile(src, os.path.join(REF, 'chroma', 'bvh', 'bvh.py'), 'exec'), bmod.__dict__)
    sys.modules['chroma.bvh.bvh'] = bmod
    gpkg = types.ModuleType('chroma.gpu')
    gpkg.__path__ = []
    sys.modules['chroma.gpu'] = gpkg
    gb = types.ModuleType('chroma.gpu.bvh')

    def create_leaf_nodes(mesh):
        return mesh['world_coords'], mesh['leaves'].copy(), mesh['codes'].copy().view(StableCodes)

    def merge_nodes_detailed(nodes, first_child, nchild):
        record.append((np.asarray(first_child).astype(np.int32), np.asarray(nchild).astype(np.int32)))
        return np_merge_nodes_detailed(nodes, first_child, nchild)

    gb.create_leaf_nodes = create_leaf_nodes
    gb.merge_nodes_detailed = merge_nodes_detailed
    gb.concatenate_layers = np_concatenate_layers
    gb.collapse_chains = np_collapse_chains
    sys.modules['chroma.gpu.bvh'] = gb
    gsrc = open(os.path.join(REF, 'chroma', 'bvh', 'grid.py')).read()
    gmod = types.ModuleType('chroma.bvh.grid')
    exec(compile(gsrc, os.path.join(REF, 'chroma', 'bvh', 'grid.py'), 'exec'), gmod.__dict__)

    # numpy 2 refuses `np.ediff1d(uint64 array, to_begin=1)` (python int -> uint64 is no longer a
    # same-kind cast); numpy 1, which the reference targets, cast the scalar to the array's dtype
    class NumpyProxy(object):
        def __getattr__(self, k):
            return getattr(np, k)

        @staticmethod
        def ediff1d(ary, to_end=None, to_begin=None):
            ary = np.asarray(ary)
            cast = lambda v: None if v is None else np.asarray(v).astype(ary.dtype)
            return np.ediff1d(ary, to_end=cast(to_end), to_begin=cast(to_begin))
    gmod.np = NumpyProxy()
    return gmod, bmod


def golden_meshes():
    "name -> Mesh; the cases of tests/test_gpu_bvh.py plus PMT-sized ones"
    import fixtures as F
    from chroma_b200 import make
    from chroma_b200.geometry import Mesh
    det = F.small_detector()
    det.flatten()
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [100, 100, 100]], dtype=np.float32)
    return [('cube', make.cube(1000.0)),
            ('sphere', make.sphere(20000.0, nsteps=64)),
            ('detector', det.mesh),
            # many coincident centroids: groups of more than 15 children are split (grid.py:48-71)
            ('duplicates', Mesh(v, np.array([[0, 1, 2]] * 40 + [[1, 2, 3]]), remove_null_triangles=False)),
            ('torus', make.torus(50.0, 400.0, nsteps=24, circle_steps=12))]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    from oracle import oracle as O
    from chroma_b200.gpu.bvh import world_coords_for
    record = []
    grid, bmod = import_reference_grid(record)
    out = {}
    names = []
    for name, mesh in golden_meshes():
        wc = world_coords_for(mesh)
        lv, codes = O.make_leaves(mesh, wc.world_origin, wc.world_scale)
        leaves = np.ascontiguousarray(lv, dtype=np.uint32).view(UINT4)[:, 0]
        ref_wc = bmod.WorldCoords(np.asarray(wc.world_origin), wc.world_scale)
        small = len(leaves) <= 20000
        names.append(name)
        out[name + '/world_origin'] = np.asarray(wc.world_origin, np.float32)
        out[name + '/world_scale'] = np.float32(wc.world_scale)
        # inputs: stored in full when small; big meshes are rebuilt by the test from the same
        # fixture and checked against the digest before use
        out[name + '/leaves_sha256'] = np.array(digest(lv))
        out[name + '/codes_sha256'] = np.array(digest(codes))
        if small:
            out[name + '/leaves'] = np.ascontiguousarray(lv, np.uint32)
            out[name + '/codes'] = np.asarray(codes, np.uint64)
        for degree in (2, 3):
            del record[:]
            bvh = grid.make_recursive_grid_bvh(dict(world_coords=ref_wc, leaves=leaves, codes=np.asarray(codes, np.uint64)),
                                               target_degree=degree)
            nodes = np.ascontiguousarray(bvh.nodes).view(np.uint32).reshape(-1, 4)
            key = '%s/d%d' % (name, degree)
            out[key + '/layer_offsets'] = np.asarray(bvh.layer_offsets, np.int64)
            out[key + '/nnodes'] = np.int64(len(nodes))
            out[key + '/nodes_sha256'] = np.array(digest(nodes))
            if small:
                out[key + '/nodes'] = nodes
                # bottom-up: grouping of layer k's children as grid.py passed it to the kernel
                out[key + '/nlayers_merged'] = np.int64(len(record))
                for k, (fc, nc) in enumerate(record):
                    out['%s/first_child%d' % (key, k)] = fc
                    out['%s/nchild%d' % (key, k)] = nc
            print('%-12s degree %d: %8d leaves -> %8d nodes, %2d layers, %s' %
                  (name, degree, len(leaves), len(nodes), len(bvh.layer_offsets), 'full' if small else 'digest'))
    out['names'] = np.array(names)
    np.savez_compressed(OUT, **out)
    print('wrote', OUT, '%d arrays, %d bytes' % (len(out), os.path.getsize(OUT)))


if __name__ == '__main__':
    main()
