"""GPU parity: first-hit intersection, engine (through the C-ABI) vs the CPU oracle.

Bar: triangle index identical; distance identical except where the GPU's approximate
reciprocal/sqrt (div.approx, sqrt.approx of --use_fast_math, which the CPU cannot emulate)
perturbs the normalised direction: |d_gpu - d_cpu| <= 4 ulp.  The bit-exact tier against the
reference's own CUDA kernels is tests/test_reference_gpu.py."""
import numpy as np
import pytest

import fixtures as F
from chroma_b200 import gpu
from chroma_b200.tools import from_film, argsort_direction

pytestmark = pytest.mark.gpu


def ulp_diff(a, b):
    ai = a.view(np.int32).astype(np.int64)
    bi = b.view(np.int32).astype(np.int64)
    return np.abs(ai - bi)


def check(ctx, oracle_mod, geometry, pos, dir, last=None, min_hit=0.0):
    F.attach_bvh(geometry, oracle_mod.build_bvh)
    orc = oracle_mod.Oracle(geometry)
    gg = gpu.GPUGeometry(geometry)
    tri_o, dist_o, cnt_o = orc.intersect(pos, dir, last, counts=True)
    tri_g, dist_g = gpu.intersect(gg, pos, dir, last)
    assert (tri_g >= 0).mean() >= min_hit
    same = tri_g == tri_o
    # rays grazing an edge can legitimately flip between the two triangles sharing it when
    # the direction differs in the last bit; allow 1e-4 of the rays, none may differ in distance
    assert same.mean() >= 1 - 1e-4, 'triangle mismatch fraction %g' % (1 - same.mean())
    hit = same & (tri_o >= 0)
    assert ulp_diff(dist_g[hit], dist_o[hit]).max() <= 4
    assert np.all(dist_g[tri_g < 0] == -1.0)
    # device-resident path + visit counters
    dpos, ddir = gpu.to_gpu(pos.astype(np.float32).ravel()), gpu.to_gpu(dir.astype(np.float32).ravel())
    dlast = gpu.to_gpu(last.astype(np.int32)) if last is not None else None
    t2, d2, c2 = gpu.intersect(gg, dpos, ddir, dlast, counts=True)
    assert np.array_equal(t2.get(), tri_g)
    assert np.array_equal(d2.get().view(np.uint32), dist_g.view(np.uint32))
    cnt_g = c2.get()
    assert (cnt_g == cnt_o).all(axis=1).mean() > 0.99
    assert abs(cnt_g[:, 0].mean() / cnt_o[:, 0].mean() - 1) < 1e-3
    return tri_g, dist_g, cnt_g


def test_cube_film_rays(ctx, oracle_mod):
    "setup of the reference's test/test_ray_intersection.py: 480k film rays vs a 1000 mm cube"
    pos, dir = from_film()
    tri, dist, _ = check(ctx, oracle_mod, F.cube_geometry(), pos.astype(np.float32),
                         dir.astype(np.float32), min_hit=1.0)
    assert dist.min() >= 500.0 - 1e-3 and dist.max() < 870.0


def test_detector_isotropic_sorted(ctx, oracle_mod):
    "config C2 shape on a small demo detector: rays from the origin, direction-sorted"
    pos, dir = F.isotropic_rays(200000, seed=0)
    dir = dir[argsort_direction(dir)]
    check(ctx, oracle_mod, F.small_detector(), pos, dir, min_hit=1.0)


def test_detector_incoherent_rays_and_last_hit(ctx, oracle_mod):
    "random origins inside the detector; second bounce excludes the triangle just hit"
    det = F.small_detector()
    pos, dir = F.isotropic_rays(100000, seed=1, radius=1500.0)
    tri, dist, _ = check(ctx, oracle_mod, det, pos, dir, min_hit=1.0)
    hitpos = pos + dir * dist[:, None]
    _, dir2 = F.isotropic_rays(100000, seed=2)
    det2 = F.small_detector()
    check(ctx, oracle_mod, det2, hitpos.astype(np.float32), dir2, last=tri)


def test_miss_and_edge_cases(ctx, oracle_mod):
    "rays starting outside pointing away miss; axis-parallel rays (infinite inv_dir) work"
    g = F.cube_geometry(100.0)
    pos = np.array([[0, 0, 500], [0, 0, 0], [10, 20, 30], [0, 0, 0], [500, 500, 500]], dtype=np.float32)
    dir = np.array([[0, 0, 1], [1, 0, 0], [0, -1, 0], [0, 0, 1], [1, 1, 1]], dtype=np.float32)
    tri, dist, _ = check(ctx, oracle_mod, g, pos, dir)
    assert tri[0] == -1 and tri[4] == -1
    assert np.allclose(dist[1:4], [50, 70, 50])
    # empty input is a no-op
    F.attach_bvh(g, oracle_mod.build_bvh)
    t, d = gpu.intersect(gpu.GPUGeometry(g), np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32))
    assert len(t) == 0 and len(d) == 0
