"""GPU tests of the engine's own search tree (chroma_b200/csrc/wide.cuh, engine_wide.cu) and of the
streamed photon upload.  The wide tree must return, for every ray, the triangle and the distance
bits of the reference-order search (mesh.h:41-118) — the claim argued in DESIGN.md section 4."""
import os

import numpy as np
import pytest

import fixtures as F
from chroma_b200 import _lib, event, gpu
from chroma_b200.geometry import Geometry, Solid, Mesh, vacuum

pytestmark = pytest.mark.gpu


class env(object):
    "temporarily set environment switches of the library"

    def __init__(self, **kw):
        self.kw = {k: str(v) for k, v in kw.items()}

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kw}
        os.environ.update(self.kw)
        gpu.current_context().reload_knobs()     # the library reads its knobs once per context

    def __exit__(self, *a):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
        gpu.current_context().reload_knobs()


# the reference-order search the wide tree is compared with: the warp-scheduled kernel of round 1 in an
# A/B build (python -m chroma_b200.csrc.build --variant ab -DCHROMA_AB_KERNELS=1, CHROMA_B200_LIB=...),
# else the plain one-ray-per-lane kernel the counting entry uses (same order, same result)
REF_ORDER_V = 2 if _lib.has_ab_kernels() else 1


def both_searches(gg, pos, dir, last=None):
    with env(CHROMA_ISECT_V=REF_ORDER_V):
        tri_r, dist_r = gpu.intersect(gg, pos, dir, last)      # reference-order search, reference tree
    with env(CHROMA_ISECT_V=3):
        tri_w, dist_w = gpu.intersect(gg, pos, dir, last)      # near-first search, wide tree
    bad = np.flatnonzero((tri_r != tri_w) | (dist_r.view(np.uint32) != dist_w.view(np.uint32)))
    assert len(bad) == 0, '%d / %d rays differ; first: ray %d reference-order (%d, %r) wide (%d, %r)' % (
        len(bad), len(pos), bad[0], tri_r[bad[0]], dist_r[bad[0]], tri_w[bad[0]], dist_w[bad[0]])
    return tri_w, dist_w


@pytest.mark.parametrize('leaf', [1, 3, 8])
def test_wide_equals_reference_order_on_detector(ctx, oracle_mod, leaf):
    with env(CHROMA_WIDE_LEAF=leaf):
        det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
        gg = gpu.GPUGeometry(det)
    info = gg.info()
    assert info['wide_groups'] > 0 and 1 <= info['wide_levels'] <= 64
    pos, dir = F.isotropic_rays(300000, seed=5)
    tri, dist = both_searches(gg, pos, dir)
    assert (tri >= 0).mean() > 0.99
    pos, dir = F.isotropic_rays(300000, seed=6, radius=1800.0)
    tri, dist = both_searches(gg, pos, dir)
    ok = tri >= 0
    hitpos = (pos + dir * dist[:, None]).astype(np.float32)[ok]
    _, dir2 = F.isotropic_rays(int(ok.sum()), seed=7)
    both_searches(gg, hitpos, dir2, last=tri[ok])              # starts on a surface, mesh.h:82
    # the wide tree does much less work per ray than the reference-order search
    dp, dd = gpu.to_gpu(pos.ravel()), gpu.to_gpu(dir.ravel())
    _, _, c_ref = gpu.intersect(gg, dp, dd, counts=True)
    with env(CHROMA_COUNT_WIDE=1):
        _, _, c_w = gpu.intersect(gg, dp, dd, counts=True)
    assert c_w.get()[:, 0].mean() * 2 < c_ref.get()[:, 0].mean()


def test_wide_degenerate_rays(ctx, oracle_mod):
    "axis-parallel, unnormalised, zero-component and missing rays; rays along box faces"
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUGeometry(det)
    dirs = [[0, 0, 1], [0, 0, -1], [1, 0, 0], [0, -1, 0], [0, 3, 4], [1e-30, 1, 0], [-1e-30, 0, 1],
            [1, 1, 0], [0.3, -0.2, 0.1], [1e-20, 1e-20, 1], [5, 0, 0]]
    pos = np.zeros((len(dirs), 3), np.float32)
    both_searches(gg, pos, np.array(dirs, np.float32))
    both_searches(gg, np.full((3, 3), 1e5, np.float32), np.array([[1, 1, 1], [-1, -1, -1], [0, 0, 1]], np.float32))
    g = F.attach_bvh(F.cube_geometry(100.0), oracle_mod.build_bvh)
    gc = gpu.GPUGeometry(g)
    # along a face plane and through an edge / a corner of the cube
    pos = np.array([[50, 0, -200], [50, 50, -200], [0, 0, 0], [0, 0, 0]], np.float32)
    dir = np.array([[0, 0, 1], [0, 0, 1], [1, 1, 0], [1, 1, 1]], np.float32)
    both_searches(gc, pos, dir)


def test_wide_ties_follow_the_reference_rank(ctx, oracle_mod):
    """coincident triangles: equal distances are resolved by the reference's visiting order
    (mesh.h:88), whatever the order in which the wide tree meets them"""
    rs = np.random.RandomState(3)
    base = F.small_detector()
    base.flatten()
    m = base.mesh
    # the whole mesh twice: every hit is a tie between a triangle and its copy
    tris = np.vstack([m.triangles, m.triangles + len(m.vertices)])
    tris = tris[rs.permutation(len(tris))]              # either copy may come first in triangle order
    mesh = Mesh(np.vstack([m.vertices, m.vertices]), tris, remove_duplicate_vertices=False)
    g = Geometry(vacuum)
    g.add_solid(Solid(mesh, vacuum, vacuum))
    F.attach_bvh(g, oracle_mod.build_bvh)
    gg = gpu.GPUGeometry(g)
    pos, dir = F.isotropic_rays(200000, seed=8, radius=1500.0)
    tri, dist = both_searches(gg, pos, dir)
    tri_o, dist_o = oracle_mod.Oracle(g).intersect(pos, dir)
    assert (tri == tri_o).mean() > 1 - 1e-4            # CPU restatement picks the same copy


@pytest.mark.parametrize('ntri', [1, 2, 5, 12])
def test_wide_tiny_meshes(ctx, oracle_mod, ntri):
    "fewer triangles than one group; a single leaf at the root is never tested by the reference"
    rs = np.random.RandomState(ntri)
    v = rs.uniform(-100, 100, (ntri * 3, 3))
    v[:, 2] = np.repeat(np.arange(1, ntri + 1) * 50.0, 3)      # stacked sheets along z
    mesh = Mesh(v, np.arange(ntri * 3).reshape(-1, 3), remove_duplicate_vertices=False)
    g = Geometry(vacuum)
    g.add_solid(Solid(mesh, vacuum, vacuum))
    F.attach_bvh(g, oracle_mod.build_bvh)
    gg = gpu.GPUGeometry(g)
    n = 20000
    pos = np.zeros((n, 3), np.float32)
    dir = np.column_stack([rs.uniform(-1, 1, n), rs.uniform(-1, 1, n), np.ones(n)]).astype(np.float32)
    tri, dist = both_searches(gg, pos, dir)
    tri_o, _ = oracle_mod.Oracle(g).intersect(pos, dir)
    assert np.array_equal(tri, tri_o)


@pytest.mark.skipif(not _lib.has_ab_kernels(), reason="round 1's kernels on the reference-format tree are compiled "
                    "into A/B builds only (-DCHROMA_AB_KERNELS=1)")
def test_propagate_same_on_both_trees(ctx, oracle_mod):
    "counter-based RNG + identical first hits => photon histories are identical, kernel by kernel"
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    ph = F.isotropic_photons(200000, seed=21, wavelength_range=(400, 800), t0=100.0)
    outs = {}
    for v in (2, 3, 4, 5):
        with env(CHROMA_PROP_V=v):
            rng = gpu.get_rng_states(1, seed=9)
            gp = gpu.GPUPhotons(ph)
            gp.propagate(gg, rng, max_steps=10)
            outs[v] = gp.get()
    for v in (2, 3, 5):
        assert np.array_equal(outs[v].flags, outs[4].flags)
        assert np.array_equal(outs[v].last_hit_triangles, outs[4].last_hit_triangles)
        assert np.array_equal(outs[v].pos.view(np.uint32), outs[4].pos.view(np.uint32))
        assert np.array_equal(outs[v].t.view(np.uint32), outs[4].t.view(np.uint32))
        assert np.array_equal(outs[v].dir.view(np.uint32), outs[4].dir.view(np.uint32))
        assert np.array_equal(outs[v].pol.view(np.uint32), outs[4].pol.view(np.uint32))


def test_streamed_upload_and_chunked_propagate(ctx, oracle_mod):
    "chunk plan of chroma_photons_upload: same photons, same histories, pinned or pageable"
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    n = 700000
    ph = F.isotropic_photons(n, seed=22, wavelength_range=(400, 800), t0=100.0)
    fields = ('pos', 'dir', 'pol', 'wavelengths', 't', 'last_hit_triangles', 'flags', 'weights', 'evidx')
    pinned = event.Photons(*[gpu.pinned_copy(getattr(ph, f)) for f in fields])
    ref = None
    for chunks, src in ((1, ph), (0, ph), (5, ph), (0, pinned), (3, pinned)):
        with env(CHROMA_UPLOAD_CHUNKS=chunks):
            gp = gpu.GPUPhotons(src)
        back = gp.get()                                       # joins the pending chunks
        for f in fields:
            assert np.array_equal(getattr(back, f), getattr(ph, f)), f
        with env(CHROMA_UPLOAD_CHUNKS=chunks):
            gp = gpu.GPUPhotons(src)
            rng = gpu.get_rng_states(1, seed=4)
            gp.propagate(gg, rng, max_steps=10)               # starts per chunk
        out = gp.get()
        if ref is None:
            ref = out
        else:
            assert np.array_equal(out.flags, ref.flags)
            assert np.array_equal(out.pos.view(np.uint32), ref.pos.view(np.uint32))
    # a propagate over a sub-range while an upload is pending
    with env(CHROMA_UPLOAD_CHUNKS=4):
        gp = gpu.GPUPhotons(pinned)
        rng = gpu.get_rng_states(1, seed=4)
        for sl in gp.iterate_copies() if False else [gp]:
            sl.propagate(gg, rng, max_steps=10)
    assert np.array_equal(gp.get().flags, ref.flags)


def test_rays_the_grid_cannot_hold_end_without_a_hit(ctx, oracle_mod):
    """wide.cuh wide_ray_unsearchable: infinite / huge / NaN origins and non-finite directions (a photon
    uploaded with such coordinates, or garbage read after the upload watchdog fired) must end as "no hit",
    on both geometries, and must not follow an empty record's index out of the arrays (the two-pass group
    step does not test for empty records: their inverted box rejects every ray the grid can hold)."""
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    inf, nan = np.inf, np.nan
    pos = np.array([[inf, 0, 0], [1e30, 0, 0], [nan, 0, 0], [-inf, inf, 1e38], [3e38, -3e38, 3e38], [1e17, 0, 0],
                    [0, 0, 0], [0, 0, 0], [0, 0, 0], [0, 0, 0]], np.float32)
    d = np.array([[1, 0, 0], [1, 0, 0], [1, 0, 0], [1, 1, 1], [1, 0, 0], [-1, 0, 0],
                  [inf, inf, inf], [nan, 1, 0], [1e30, 1e30, 1e30], [inf, 0, 0]], np.float32)
    reps = 2000
    pos, d = np.tile(pos, (reps, 1)), np.tile(d, (reps, 1))       # every warp of a few CTAs sees them
    for gg in (gpu.GPUDetector(det), gpu.GPUDetector(det, instanced=True)):
        tri, dist = gpu.intersect(gg, pos, d)
        assert (tri.reshape(reps, -1)[:, :6] == -1).all()
        # the same as photons: no crash, every photon ends
        ph = event.Photons(pos, d, np.tile(np.array([[0, 1, 0]], np.float32), (len(pos), 1)),
                           np.full(len(pos), 500.0, np.float32))
        gp = gpu.GPUPhotons(ph)
        gp.propagate(gg, gpu.get_rng_states(1, seed=2), max_steps=10)
        fl = gp.get().flags
        terminal = event.NO_HIT | event.SURFACE_DETECT | event.SURFACE_ABSORB | event.BULK_ABSORB | event.NAN_ABORT
        assert ((fl.reshape(reps, -1)[:, :6] & (event.NO_HIT | event.NAN_ABORT)) != 0).all()
        assert (fl & terminal != 0).mean() > 0.5
        # and the context is intact: ordinary rays still hit
        p, dd = F.isotropic_rays(20000, seed=1)
        t2, _ = gpu.intersect(gg, p, dd)
        assert (t2 >= 0).mean() > 0.99
