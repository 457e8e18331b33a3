"""Generate tests/golden/host_golden.npz by importing the Python reference from
/root/reference (build container only; the GPU box has no reference checkout).

What is pinned: the reference's HOST-side numpy code on the hot path's boundary
(chroma/bvh/bvh.py node packing and WorldCoords, chroma/make.py meshes, chroma/geometry.py
Mesh/Geometry.flatten material-code packing, chroma/tools.py argsort_direction/from_film,
chroma/sample.py, chroma/transform.py, chroma/detector.py CDFs).  The CUDA kernels of the
reference cannot run here (no GPU, no PyCUDA); their pin is tests/test_reference_gpu.py, which
runs the reference's own kernels (oracle/_ref/*.cubin) next to the engine on the GPU box.

Shims (reference bugs / missing packages): `pycuda.gpuarray.vec.uint4` is replaced
by the equivalent numpy record dtype; `np.asfarray` (removed in numpy 2) is restored; the
numpy-2 overflow of the NCHILD_MASK literal in chroma/bvh/bvh.py:9 is masked to 32 bits; the
numpy-2 `inverse` shape in Mesh.remove_duplicate_vertices is flattened.

Usage: python tests/golden/make_golden.py
"""
import os
import sys
import types

import numpy as np

REF = '/root/reference'
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'host_golden.npz')


def import_reference():
    if not hasattr(np, 'asfarray'):
        np.asfarray = lambda a, dtype=np.float64: np.asarray(a, dtype=dtype)
    pkg = types.ModuleType('chroma')
    pkg.__path__ = [os.path.join(REF, 'chroma')]
    sys.modules['chroma'] = pkg
    # pycuda stand-in: only the uint4 record dtype is used by chroma/bvh/bvh.py
    pycuda = types.ModuleType('pycuda')
    ga = types.ModuleType('pycuda.gpuarray')
    ga.vec = types.SimpleNamespace(uint4=np.dtype([('x', np.uint32), ('y', np.uint32), ('z', np.uint32),
                                                   ('w', np.uint32)]))
    pycuda.gpuarray = ga
    sys.modules['pycuda'] = pycuda
    sys.modules['pycuda.gpuarray'] = ga
    import importlib
    mods = {}
    # chroma/bvh/__init__.py pulls in the PyCUDA builders: register a bare package instead and
    # load bvh.py alone; numpy 2 rejects `np.uint32(0xFFFF << 28)` (numpy 1 wrapped it), so that
    # one literal is masked to 32 bits -- same value the reference gets under numpy 1
    bpkg = types.ModuleType('chroma.bvh')
    bpkg.__path__ = [os.path.join(REF, 'chroma', 'bvh')]
    sys.modules['chroma.bvh'] = bpkg
    src = open(os.path.join(REF, 'chroma', 'bvh', 'bvh.py')).read()
    src = src.replace('np.uint32(0xFFFF << CHILD_BITS)', 'np.uint32((0xFFFF << CHILD_BITS) & 0xFFFFFFFF)')
    bmod = types.ModuleType('chroma.bvh.bvh')
    exec(compile(src, os.path.join(REF, 'chroma', 'bvh', 'bvh.py'), 'exec'), bmod.__dict__)
    sys.modules['chroma.bvh.bvh'] = bmod
    mods['bvh.bvh'] = bmod
    for name in ('make', 'geometry', 'tools', 'sample', 'transform', 'detector'):
        mods[name] = importlib.import_module('chroma.' + name)

    # numpy 2 returns `inverse` with the shape of the (N,1) structured input in
    # Mesh.remove_duplicate_vertices (chroma/geometry.py:63-67); numpy 1 returned it flat
    class NumpyProxy(object):
        def __getattr__(self, k):
            return getattr(np, k)

        @staticmethod
        def unique(a, *args, **kw):
            r = np.unique(a, *args, **kw)
            if kw.get('return_inverse') and isinstance(r, tuple):
                r = (r[0], r[1].ravel()) + tuple(r[2:])
            return r
    mods['geometry'].np = NumpyProxy()
    return mods


def main():
    m = import_reference()
    out = {}
    B = m['bvh.bvh']
    # --- test/test_bvh.py:54-121: the hand-built 3-layer binary tree -------------------------
    nodes = np.zeros(7, dtype=B.uint4)
    nodes['x'][3:7] = [0x00010000, 0x00020001, 0x00010000, 0x00010000]
    nodes['y'][3:7] = [0x00010000, 0x00010000, 0x00020001, 0x00010000]
    nodes['z'][3:7] = [0x00010000, 0x00010000, 0x00010000, 0x00020001]
    nodes['w'][3:7] = 0x80000000
    nodes['x'][1:3] = [0x00020000, 0x00010000]
    nodes['y'][1:3] = [0x00010000, 0x00020000]
    nodes['z'][1:3] = [0x00010000, 0x00020000]
    nodes['w'][1:3] = [0x00000003, 0x00000005]
    nodes['x'][0:1], nodes['y'][0:1], nodes['z'][0:1], nodes['w'][0:1] = 0x00020000, 0x00020000, 0x00020000, 1
    wc = B.WorldCoords(np.array([-1.0, -1.0, -1.0]), 0.1)
    bvh = B.BVH(world_coords=wc, nodes=nodes, layer_offsets=[0, 1, 3])
    out['bvh_nodes'] = nodes.view(np.uint32).reshape(-1, 4)
    u = B.unpack_nodes(nodes)
    for f in u.dtype.names:
        out['bvh_unpack_' + f] = u[f]
    out['bvh_node_areas'] = B.node_areas(nodes)
    for i in range(3):
        layer = bvh.get_layer(i)
        lo, hi = layer.get_bounds()
        out['bvh_layer%d_lo' % i], out['bvh_layer%d_hi' % i] = np.atleast_2d(lo), np.atleast_2d(hi)
        out['bvh_layer%d_area_fixed' % i] = np.float64(layer.area_fixed())
        out['bvh_layer%d_area' % i] = np.float64(layer.area())
    # --- WorldCoords (test/test_bvh.py:13-52) ---------------------------------------------------
    rs = np.random.RandomState(5)
    w = rs.uniform(-1, 5000, (64, 3))
    out['wc_world'] = w
    out['wc_fixed'] = wc.world_to_fixed(w)
    out['wc_back'] = wc.fixed_to_world(out['wc_fixed'])
    # --- meshes (chroma/make.py) ----------------------------------------------------------------
    mk = m['make']
    for name, mesh in (('cube', mk.cube(1000.0)), ('box', mk.box(100.0, 200.0, 300.0, center=(1, 2, 3))),
                       ('sphere', mk.sphere(20000.0, nsteps=8)), ('cyl', mk.cylinder(10.0, 30.0, nsteps=6)),
                       ('torus', mk.torus(5.0, 20.0, nsteps=6, circle_steps=5))):
        out['mesh_%s_v' % name] = np.asarray(mesh.vertices, np.float64)
        out['mesh_%s_t' % name] = np.asarray(mesh.triangles, np.int64)
    # --- tools / sample / transform -------------------------------------------------------------
    T, S, X = m['tools'], m['sample'], m['transform']
    np.random.seed(11)
    d = S.uniform_sphere(1000)
    out['sample_uniform_sphere_seed11'] = d
    out['argsort_direction'] = T.argsort_direction(d)
    pos, dirs = T.from_film(size=(8, 6))
    out['from_film_pos'], out['from_film_dir'] = pos, dirs
    np.random.seed(12)
    out['flashlight_seed12'] = S.flashlight(phi=0.3, direction=(0, 1, 1), size=50)
    v = rs.normal(size=(20, 3))
    out['rot_in'] = v
    out['rot_out'] = X.rotate(v, 0.7, X.normalize(np.array([1.0, 2.0, 3.0])))
    out['rot_matrix'] = X.make_rotation_matrix(0.7, X.normalize(np.array([1.0, 2.0, 3.0])))
    # --- Geometry.flatten: material / surface indices, solid ids, displaced+rotated solids -------
    G = m['geometry']
    water, glass = G.Material('water'), G.Material('glass')
    surf = G.Surface('s')
    geo = G.Geometry(water)
    geo.add_solid(G.Solid(mk.cube(10.0), glass, water, surface=surf), displacement=(5, 0, 0),
                  rotation=X.make_rotation_matrix(0.3, (0, 0, 1)))
    geo.add_solid(G.Solid(mk.sphere(30.0, nsteps=6), water, water))
    geo.flatten()
    out['flat_vertices'] = np.asarray(geo.mesh.vertices, np.float64)
    out['flat_triangles'] = np.asarray(geo.mesh.triangles, np.int64)
    # index values are labels into unique_materials / unique_surfaces whose order is an artefact of
    # set iteration in the reference (geometry.py silly_unique): pin the objects they resolve to
    mname = lambda i: geo.unique_materials[i].name
    sname = lambda i: 'None' if i < 0 or geo.unique_surfaces[i] is None else geo.unique_surfaces[i].name
    out['flat_inner'] = np.array([mname(i) for i in geo.inner_material_index])
    out['flat_outer'] = np.array([mname(i) for i in geo.outer_material_index])
    out['flat_surface'] = np.array([sname(i) for i in geo.surface_index])
    out['flat_solid_id'] = np.asarray(geo.solid_id, np.int64)
    # --- Detector time / charge CDFs (chroma/detector.py) ---------------------------------------
    D = m['detector']
    det = D.Detector(water)
    det.set_time_dist_gaussian(1.5, -7.5, 7.5)
    det.set_charge_dist_gaussian(1.0, 0.1, 0.0, 1.5)
    out['time_cdf_x'], out['time_cdf_y'] = np.asarray(det.time_cdf[0]), np.asarray(det.time_cdf[1])
    out['charge_cdf_x'], out['charge_cdf_y'] = np.asarray(det.charge_cdf[0]), np.asarray(det.charge_cdf[1])
    np.savez_compressed(OUT, **out)
    print('wrote', OUT, '%d arrays, %d bytes' % (len(out), os.path.getsize(OUT)))


if __name__ == '__main__':
    main()
