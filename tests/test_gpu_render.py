"""GPU parity of the headless ray-cast render: GPURays.render / snapshot and the
ray transforms (csrc/engine_render.cu) against the reference's own render.cu / transform.cu kernels
(oracle/_ref/render.cubin, transform.cubin) and the CPU oracle.  Bar: pixels, list lengths and list
distances identical to the reference kernels, ray for ray; colours and transformed rays within
1e-6 relative (float expressions compiled in another translation unit); against the CPU oracle
(libm division / sqrt instead of the device's fast-math forms) list lengths and distances identical,
pixel channels within one unit."""
import numpy as np
import pytest

import fixtures as F
from chroma_b200 import gpu, make
from chroma_b200.geometry import Geometry, Solid
from chroma_b200.tools import from_film

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def rctx(ctx):
    from oracle import ref_gpu
    if not ref_gpu.available():
        pytest.fail('oracle/_ref/*.cubin missing (built by __graft_entry__.build() where the reference is mounted)')
    r = ref_gpu.RefContext(0)
    yield r
    r.close()


def film(position, size=(160, 120), axis1=(0, 0, 1), axis2=(1, 0, 0)):
    pos, dir = from_film(position=position, size=size, axis1=axis1, axis2=axis2)
    return pos.astype(np.float32), dir.astype(np.float32)


def channels(px):
    return np.stack([(px >> s) & 0xff for s in (24, 16, 8, 0)], axis=1).astype(np.int64)


def compare(rctx, oracle_mod, geo, pos, dir, alpha_depth, bg_color, max_alpha_depth=10):
    from oracle import ref_gpu
    gg = gpu.GPUGeometry(geo)
    rg = ref_gpu.RefGeometry(rctx, geo)
    rays = gpu.GPURays(pos, dir, max_alpha_depth=max_alpha_depth)
    rref = ref_gpu.RefRays(rctx, pos, dir, max_alpha_depth=max_alpha_depth)
    pixels = gpu.empty(len(pos), np.uint32)
    rays.render(gg, pixels, alpha_depth=alpha_depth, bg_color=bg_color)
    px = pixels.get()
    px_ref = rref.render(rg, alpha_depth=alpha_depth, bg_color=bg_color)
    n_ref, dx_ref, col_ref = rref.lists(alpha_depth)
    n_got = rays.dxlen.get()
    dx_got = rays.dx.get()[:len(pos) * alpha_depth].reshape(len(pos), alpha_depth)
    col_got = rays.color.get().view(np.float32).reshape(-1, 4)[:len(pos) * alpha_depth].reshape(len(pos), alpha_depth, 4)
    assert np.array_equal(n_got, n_ref)
    used = np.arange(alpha_depth)[None, :] < n_got[:, None]
    assert np.array_equal(dx_got.view(np.uint32)[used], dx_ref.view(np.uint32)[used])
    np.testing.assert_allclose(col_got[used], col_ref[used], rtol=1e-6, atol=1e-6)
    bad = np.flatnonzero(px != px_ref)
    assert len(bad) == 0, '%d of %d pixels differ from the reference kernel; first %d: %08x vs %08x' % (
        len(bad), len(px), bad[0], px[bad[0]], px_ref[bad[0]])
    # CPU oracle
    px_o, (dx_o, n_o, _) = oracle_mod.Oracle(geo).render(pos, dir, alpha_depth=alpha_depth, bg_color=bg_color)
    assert np.array_equal(n_got, n_o)
    assert np.array_equal(dx_got.view(np.uint32)[used], dx_o.view(np.uint32)[used])
    assert np.abs(channels(px) - channels(px_o)).max() <= 1
    return px, n_got, (gg, rg, rays, rref)


def test_cube_inside_and_outside(ctx, rctx, oracle_mod):
    geo = F.attach_bvh(F.cube_geometry(1000.0), oracle_mod.build_bvh)
    geo.colors[:] = np.random.RandomState(0).randint(0, 2 ** 32, len(geo.colors), dtype=np.uint64).astype(np.uint32)
    pos, dir = film((0, 0, 0))
    px, n, _ = compare(rctx, oracle_mod, geo, pos, dir, 10, 0)
    assert (n >= 1).all()                       # from inside every ray leaves through a face
    pos, dir = film((0, -3000.0, 0), axis1=(0, 0, 1), axis2=(1, 0, 0))
    dir = -dir if (dir[:, 1] < 0).mean() > 0.5 else dir
    px, n, _ = compare(rctx, oracle_mod, geo, pos, dir, 10, 0x80102030)
    assert (n == 2).any() and (n == 0).any()    # through two faces, or past the cube
    assert (px[n == 0] == 0x80102030).all()


@pytest.mark.parametrize('alpha_depth,bg', [(1, 0), (3, 0xFF000000), (10, 0x40FFFFFF)])
def test_detector_vs_reference(ctx, rctx, oracle_mod, alpha_depth, bg):
    "a demo detector seen from outside: up to dozens of surfaces along a ray, coincident PMT layers"
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    pos, dir = film((0, -6000.0, 0), size=(200, 150))
    dir = -dir if (dir[:, 1] < 0).mean() > 0.5 else dir
    px, n, _ = compare(rctx, oracle_mod, det, pos, dir, alpha_depth, bg)
    assert n.max() == alpha_depth and len(np.unique(px)) > 50


def test_keep_last_render_and_second_geometry(ctx, rctx, oracle_mod):
    "gpu/render.py:46-47: with keep_last_render the lists of the previous pass are merged into"
    from oracle import ref_gpu
    box = Geometry(F.vacuum)
    box.add_solid(Solid(make.cube(400.0), F.vacuum, F.vacuum, color=0x330000FF), displacement=(100, 0, 50))
    box = F.attach_bvh(box, oracle_mod.build_bvh)
    cube = F.attach_bvh(F.cube_geometry(1000.0), oracle_mod.build_bvh)
    pos, dir = film((0, -3000.0, 0))
    dir = -dir if (dir[:, 1] < 0).mean() > 0.5 else dir
    px, n, (gg, rg, rays, rref) = compare(rctx, oracle_mod, cube, pos, dir, 4, 0)
    gbox, rbox = gpu.GPUGeometry(box), ref_gpu.RefGeometry(rctx, box)
    pixels = gpu.empty(len(pos), np.uint32)
    for _ in range(2):        # the second pass re-offers hits equal to entries already in the list
        rays.render(gbox, pixels, alpha_depth=4, keep_last_render=True, bg_color=0)
        px_ref = rref.render(rbox, alpha_depth=4, keep_last_render=True, bg_color=0)
        n_ref, dx_ref, _ = rref.lists(4)
        assert np.array_equal(rays.dxlen.get(), n_ref)
        used = np.arange(4)[None, :] < n_ref[:, None]
        assert np.array_equal(rays.dx.get()[:len(pos) * 4].reshape(-1, 4).view(np.uint32)[used], dx_ref.view(np.uint32)[used])
        assert np.array_equal(pixels.get(), px_ref)
    assert n_ref.max() == 4
    # without keep_last_render the lists start empty again
    rays.render(gbox, pixels, alpha_depth=4, bg_color=0)
    fresh = gpu.GPURays(pos, dir, max_alpha_depth=10)
    assert np.array_equal(fresh.snapshot(gbox, alpha_depth=4), pixels.get())
    assert np.array_equal(rays.dxlen.get(), fresh.dxlen.get()) and rays.dxlen.get().max() < 4


def test_transforms_vs_reference(ctx, rctx):
    from oracle import ref_gpu
    pos, dir = film((100.0, -200.0, 300.0), size=(64, 48))
    rays, rref = gpu.GPURays(pos, dir), ref_gpu.RefRays(rctx, pos, dir)
    axis = np.array([1.0, 2.0, -0.5]) / np.linalg.norm([1.0, 2.0, -0.5])
    for step in (lambda r: r.rotate(0.37, axis), lambda r: r.translate((10.0, -20.0, 5.5)),
                 lambda r: r.rotate_around_point(-1.2, axis, (50.0, 60.0, -70.0))):
        step(rays); step(rref)
        p_ref, d_ref = rref.get()
        np.testing.assert_allclose(rays.pos.get().view(np.float32).reshape(-1, 3), p_ref, rtol=2e-6, atol=1e-4)
        np.testing.assert_allclose(rays.dir.get().view(np.float32).reshape(-1, 3), d_ref, rtol=2e-6, atol=1e-6)
    # and against the rotation written in numpy (chroma/transform.py rotate)
    from chroma_b200.transform import rotate
    r2 = gpu.GPURays(pos, dir)
    r2.rotate(0.37, axis)
    np.testing.assert_allclose(r2.dir.get().view(np.float32).reshape(-1, 3), rotate(dir, 0.37, axis), atol=2e-6)


def test_render_api_errors_and_snapshot(ctx, oracle_mod):
    geo = F.attach_bvh(F.cube_geometry(1000.0), oracle_mod.build_bvh)
    gg = gpu.GPUGeometry(geo)
    pos, dir = film((0, 0, 0), size=(32, 24))
    rays = gpu.GPURays(pos, dir, max_alpha_depth=4)
    with pytest.raises(Exception, match='alpha_depth > max_alpha_depth'):
        rays.render(gg, gpu.empty(len(pos), np.uint32), alpha_depth=5)
    with pytest.raises(TypeError):
        rays.render(gg, np.empty(len(pos), np.uint32), alpha_depth=4)
    with pytest.raises(ValueError):
        rays.render(gg, gpu.empty(len(pos) - 1, np.uint32), alpha_depth=4)
    px = rays.snapshot(gg, alpha_depth=4)
    assert px.dtype == np.uint32 and px.shape == (len(pos),) and (px >> 24).min() > 0
