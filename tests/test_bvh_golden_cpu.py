"""V2 pin, CPU half: the oracle's restatement of the recursive-grid builder against the REFERENCE's
own chroma/bvh/grid.py:11-95, run in the build container on fixed leaves and committed as
tests/golden/bvh_golden.npz (generator: tests/golden/make_bvh_golden.py).  The GPU half
(tests/test_gpu_bvh.py) compares chroma_bvh_build_from_leaves and the reference's bvh.cubin
kernels with the same file."""
import hashlib
import os

import numpy as np
import pytest

import fixtures as F
from chroma_b200 import make
from chroma_b200.bvh.bvh import WorldCoords
from chroma_b200.geometry import Mesh

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'bvh_golden.npz')


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def golden_mesh(name):
    if name == 'cube':
        return make.cube(1000.0)
    if name == 'sphere':
        return make.sphere(20000.0, nsteps=64)
    if name == 'detector':
        det = F.small_detector()
        det.flatten()
        return det.mesh
    if name == 'duplicates':
        v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [100, 100, 100]], dtype=np.float32)
        return Mesh(v, np.array([[0, 1, 2]] * 40 + [[1, 2, 3]]), remove_null_triangles=False)
    if name == 'torus':
        return make.torus(50.0, 400.0, nsteps=24, circle_steps=12)
    raise KeyError(name)


def golden_inputs(G, name, oracle_mod):
    """(world_coords, leaves uint32[n][4], codes uint64[n]) the golden was made from: stored in full
    for small meshes, rebuilt by the oracle's make_leaves and checked against the digest otherwise."""
    wc = WorldCoords(G[name + '/world_origin'], G[name + '/world_scale'])
    if name + '/leaves' in G.files:
        leaves, codes = G[name + '/leaves'], G[name + '/codes']
    else:
        leaves, codes = oracle_mod.make_leaves(golden_mesh(name), wc.world_origin, wc.world_scale)
    assert digest(np.ascontiguousarray(leaves, np.uint32)) == str(G[name + '/leaves_sha256']), name
    assert digest(np.asarray(codes, np.uint64)) == str(G[name + '/codes_sha256']), name
    return wc, np.ascontiguousarray(leaves, np.uint32), np.asarray(codes, np.uint64)


def check_against_golden(G, name, degree, nodes, layer_offsets):
    key = '%s/d%d' % (name, degree)
    nodes = np.ascontiguousarray(nodes).view(np.uint32).reshape(-1, 4)
    assert list(layer_offsets) == list(G[key + '/layer_offsets']), key
    assert len(nodes) == int(G[key + '/nnodes']), key
    if key + '/nodes' in G.files:
        assert np.array_equal(nodes, G[key + '/nodes']), key
    assert digest(nodes) == str(G[key + '/nodes_sha256']), key


@pytest.mark.parametrize('degree', [2, 3])
def test_oracle_builder_equals_reference_grid_py(oracle_mod, degree):
    G = np.load(GOLDEN)
    for name in G['names']:
        name = str(name)
        wc, leaves, codes = golden_inputs(G, name, oracle_mod)
        bvh = oracle_mod.build_bvh(golden_mesh(name), target_degree=degree, leaves=leaves, codes=codes,
                                   world_coords=wc)
        check_against_golden(G, name, degree, bvh.nodes, bvh.layer_offsets)


def test_golden_covers_the_split_of_oversized_groups():
    "grid.py:48-71: the duplicates mesh has 40 coincident centroids -> a group of > 15 children is split"
    G = np.load(GOLDEN)
    nchild = G['duplicates/d3/nchild0']
    assert nchild.max() == 15 and (nchild == 15).sum() >= 2
