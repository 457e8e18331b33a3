"""GPU parity: device BVH construction vs the CPU restatement of bvh/grid.py (bit-exact).

The reference makes its leaves on the GPU under --use_fast_math (chroma/cuda/bvh.cu:148-203,
div.approx), so the engine's leaves are pinned bit for bit against the reference cubin
(tests/test_reference_gpu.py::test_make_leaves_bit_exact); the CPU restatement divides in
IEEE and may land one quantum away where a quotient sits within an ulp of an integer.  The
host part of the builder (sort, grouping, parents, concatenation, chain collapse) is exact
integer work and is compared bit for bit on the same leaves."""
import numpy as np
import pytest

import fixtures as F
from chroma_b200 import make
from chroma_b200.bvh import make_recursive_grid_bvh, make_simple_bvh, unpack_nodes
from chroma_b200.gpu import bvh as gpubvh

pytestmark = pytest.mark.gpu


def meshes():
    det = F.small_detector()
    det.flatten()
    yield 'cube', make.cube(1000.0)
    yield 'sphere', make.sphere(20000.0, nsteps=64)
    yield 'detector', det.mesh
    from chroma_b200.geometry import Mesh
    yield 'one-triangle', Mesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    # many coincident centroids: forces groups of more than 15 children to be split
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [100, 100, 100]], dtype=np.float32)
    yield 'duplicates', Mesh(v, np.array([[0, 1, 2]] * 40 + [[1, 2, 3]]), remove_null_triangles=False)


@pytest.mark.parametrize('degree', [2, 3])
def test_device_build_equals_oracle(ctx, oracle_mod, degree):
    for name, mesh in meshes():
        wc, leaves, codes = gpubvh.create_leaf_nodes(mesh)
        ref = oracle_mod.build_bvh(mesh, target_degree=degree, leaves=leaves, codes=codes, world_coords=wc)
        cpu = oracle_mod.build_bvh(mesh, target_degree=degree)
        got = make_recursive_grid_bvh(mesh, target_degree=degree)
        assert np.array_equal(got.world_coords.world_origin, cpu.world_coords.world_origin), name
        assert got.world_coords.world_scale == cpu.world_coords.world_scale, name
        assert got.layer_offsets == ref.layer_offsets, name
        assert np.array_equal(got.nodes.view(np.uint32), ref.nodes.view(np.uint32)), name
        u = unpack_nodes(got.nodes)
        assert u['nchild'].max() <= 15


def test_helper_api(ctx, oracle_mod):
    "create_leaf_nodes / merge_nodes_detailed / concatenate_layers / collapse_chains"
    mesh = make.sphere(100.0, nsteps=16)
    wc, leaves, codes = gpubvh.create_leaf_nodes(mesh)
    ref_nodes, ref_codes = oracle_mod.make_leaves(mesh, wc.world_origin, wc.world_scale)
    # IEEE (CPU) vs div.approx (GPU, as the reference): at most one quantum apart, rarely
    got = unpack_nodes(leaves)
    ref = unpack_nodes(np.ascontiguousarray(ref_nodes).view(leaves.dtype)[:, 0])
    for f in ('xlo', 'xhi', 'ylo', 'yhi', 'zlo', 'zhi'):
        d = np.abs(got[f].astype(np.int64) - ref[f].astype(np.int64))
        assert d.max() <= 1 and (d != 0).mean() < 0.02, f
    assert np.array_equal(got['child'], ref['child'])
    assert (codes == ref_codes).mean() > 0.95
    first = np.arange(0, len(leaves), 4, dtype=np.int32)
    nchild = np.minimum(4, len(leaves) - first).astype(np.int32)
    parents = gpubvh.merge_nodes_detailed(leaves, first, nchild)
    u, up = unpack_nodes(leaves), unpack_nodes(parents)
    for p in (0, len(first) // 2, len(first) - 1):
        sl = slice(first[p], first[p] + nchild[p])
        assert up['xlo'][p] == u['xlo'][sl].min() and up['zhi'][p] == u['zhi'][sl].max()
        assert up['child'][p] == first[p] and up['nchild'][p] == nchild[p]
    bvh = make_simple_bvh(mesh, 3)
    assert bvh.layer_count() >= 3 and unpack_nodes(bvh.nodes[:1])['nchild'][0] >= 1


# ---- V2 pin: the reference's own grid.py (golden, made in the build container) and its own
# ---- bvh.cubin kernels (run here) -----------------------------------------------------------
@pytest.mark.parametrize('degree', [2, 3])
def test_device_builder_equals_reference_grid_py(ctx, oracle_mod, degree):
    """chroma_bvh_build_from_leaves on the golden leaves == what chroma/bvh/grid.py:11-95 itself made of
    them (tests/golden/make_bvh_golden.py), node for node, for every golden mesh."""
    import test_bvh_golden_cpu as T
    from chroma_b200.bvh.grid import make_recursive_grid_bvh_from_leaves
    G = np.load(T.GOLDEN)
    for name in G['names']:
        name = str(name)
        wc, leaves, codes = T.golden_inputs(G, name, oracle_mod)
        got = make_recursive_grid_bvh_from_leaves(wc, leaves, codes, target_degree=degree)
        T.check_against_golden(G, name, degree, got.nodes, got.layer_offsets)


def test_reference_kernels_on_golden_layers(ctx, oracle_mod):
    """The numpy stand-ins the golden was made with == the reference's bvh.cubin kernels
    (make_parents_detailed bvh.cu:269-308, copy_and_offset :364-384, collapse_child :530-543), and
    == the engine's helper API, layer by layer with the groupings grid.py chose."""
    import test_bvh_golden_cpu as T
    from oracle import ref_gpu
    if not ref_gpu.available():
        pytest.skip('oracle/_ref/*.cubin not built')
    G = np.load(T.GOLDEN)
    rc = ref_gpu.RefContext(0)
    try:
        for name in ('cube', 'sphere', 'duplicates', 'torus'):
            wc, leaves, codes = T.golden_inputs(G, name, oracle_mod)
            order = np.argsort(codes, kind='stable')
            for degree in (2, 3):
                key = '%s/d%d' % (name, degree)
                layers = [leaves[order]]
                for k in range(int(G[key + '/nlayers_merged'])):
                    fc, nc = G['%s/first_child%d' % (key, k)], G['%s/nchild%d' % (key, k)]
                    ref = ref_gpu.merge_nodes_detailed(rc, layers[0], fc, nc)
                    eng = gpubvh.merge_nodes_detailed(layers[0].view(gpubvh.uint4)[:, 0], fc, nc)
                    assert np.array_equal(ref, np.ascontiguousarray(eng).view(np.uint32).reshape(-1, 4)), (key, k)
                    layers = [ref] + layers
                ref_nodes, ref_bounds = ref_gpu.concatenate_layers(rc, layers)
                eng_nodes, eng_bounds = gpubvh.concatenate_layers([l.view(gpubvh.uint4)[:, 0] for l in layers])
                assert np.array_equal(ref_bounds, eng_bounds)
                assert np.array_equal(ref_nodes, np.ascontiguousarray(eng_nodes).view(np.uint32).reshape(-1, 4)), key
                ref_final = ref_gpu.collapse_chains(rc, ref_nodes, ref_bounds)
                eng_final = gpubvh.collapse_chains(eng_nodes, eng_bounds)
                assert np.array_equal(ref_final, np.ascontiguousarray(eng_final).view(np.uint32).reshape(-1, 4)), key
                T.check_against_golden(G, name, degree, ref_final, ref_bounds[:-1])
    finally:
        rc.close()
