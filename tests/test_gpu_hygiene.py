"""GPU tests of the launcher state the round-1 review flagged:
per-context / per-geometry launch state instead of process-wide statics, the staged-upload watchdog,
status reporting for callers that do not ask for it, the bound on material / surface counts, and
decorrelated streams for ranks that are dealt events."""
import ctypes as C
import os

import numpy as np
import pytest

import fixtures as F
from chroma_b200 import _lib, dist, event, gpu, make
from chroma_b200.geometry import Geometry, Material, Solid

pytestmark = pytest.mark.gpu


def run(gg, ph, seed=3, max_steps=10):
    gp = gpu.GPUPhotons(ph)
    gp.propagate(gg, gpu.get_rng_states(1, seed=seed), max_steps=max_steps)
    return gp.get()


def same(a, b):
    return np.array_equal(a.flags, b.flags) and np.array_equal(a.pos.view(np.uint32), b.pos.view(np.uint32)) and \
        np.array_equal(a.last_hit_triangles, b.last_hit_triangles)


def test_two_geometries_and_two_contexts_in_one_process(ctx, oracle_mod):
    """The L2 access-policy window and the dynamic-shared-memory opt-in belong to (stream, geometry) and to the
    context: alternating between geometries, destroying one, and a second context give the results of a
    fresh process (engine_wide.cu kept both in function-local statics in round 1)."""
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    sph = F.attach_bvh(F.water_sphere(2000.0, nsteps=32), oracle_mod.build_bvh)
    ph = F.isotropic_photons(200000, seed=5, wavelength_range=(400, 800))
    g1, g2 = gpu.GPUDetector(det), gpu.GPUGeometry(sph)
    a1, b1 = run(g1, ph), run(g2, ph)
    a2, b2 = run(g1, ph), run(g2, ph)
    assert same(a1, a2) and same(b1, b2) and not same(a1, b1)
    del g1                                   # the window pointed into g1's groups: withdrawn with it
    import gc
    gc.collect()
    assert same(run(g2, ph), b1)
    g3 = gpu.GPUDetector(det)                # same detector again, new device arrays
    assert same(run(g3, ph), a1)
    # a second context on the same device: its own stream, scratch and launch state
    c2 = gpu.create_cuda_context(0)
    try:
        h = gpu.GPUDetector(det)
        assert h.ctx is c2 and same(run(h, ph), a1)
        del h
        gc.collect()
    finally:
        c2.pop()
    assert gpu.current_context() is ctx and same(run(g3, ph), a1)


def test_staged_upload_watchdog_and_sticky_status(ctx, oracle_mod):
    """engine_wide.cu: a propagate kernel that waits for staged photons longer than the watchdog gives up,
    raises status bit 1 and the call fails; a caller that passes no status pointer gets the failure from
    the next synchronising call; the context stays usable."""
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    n = 6_000_000                            # 360 MB over PCIe: milliseconds, the watchdog is set to ~2 us
    ph = F.isotropic_photons(n, seed=6, wavelength_range=(400, 800))
    fields = ('pos', 'dir', 'pol', 'wavelengths', 't', 'last_hit_triangles', 'flags', 'weights', 'evidx')
    pinned = event.Photons(*[gpu.pinned_copy(getattr(ph, f)) for f in fields])
    os.environ['CHROMA_UPLOAD_WATCHDOG'] = '2'
    ctx.reload_knobs()
    try:
        gp = gpu.GPUPhotons(pinned)
        with pytest.raises(_lib.ChromaError, match='upload stalled'):
            gp.propagate(gg, gpu.get_rng_states(1, seed=1), max_steps=10)
        # the same through the C-ABI without a status pointer: success now, error at the next synchronise
        gp = gpu.GPUPhotons(pinned)
        rng = gpu.get_rng_states(1, seed=1)
        _lib.call('chroma_photons_propagate', gp.handle, gg.handle, rng.handle, 0, n, 10, 0, 0, None)
        with pytest.raises(_lib.ChromaError, match='upload stalled'):
            ctx.synchronize()
        ctx.synchronize()                    # reported once
    finally:
        os.environ.pop('CHROMA_UPLOAD_WATCHDOG')
        ctx.reload_knobs()
    small = F.isotropic_photons(100000, seed=7, wavelength_range=(400, 800))
    out = run(gg, small)
    assert ((out.flags & (event.NO_HIT | event.SURFACE_DETECT | event.SURFACE_ABSORB | event.BULK_ABSORB)) != 0).mean() > 0.9


def test_pageable_and_by_value_uploads_equal_pinned(ctx, oracle_mod):
    "the page-locked ring for pageable sources and fields passed by value give the photons of a plain upload"
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    n = 1_500_000
    ph = F.isotropic_photons(n, seed=8, wavelength_range=(400, 800), t0=100.0, pos=(10, -20, 30))
    fields = ('pos', 'dir', 'pol', 'wavelengths', 't', 'last_hit_triangles', 'flags', 'weights', 'evidx')
    pinned = event.Photons(*[gpu.pinned_copy(getattr(ph, f)) for f in fields])
    slim = event.Photons(np.broadcast_to(np.array([10, -20, 30], np.float32), (n, 3)), ph.dir, ph.pol, ph.wavelengths,
                         t=np.broadcast_to(np.float32(100.0), (n,)))
    back = gpu.GPUPhotons(slim).get()
    for f in fields:
        assert np.array_equal(getattr(back, f), getattr(ph, f)), f
    ref = None
    for src, kw in ((pinned, {}), (ph, {}), (slim, {}), (slim, dict(copy_weights=False, copy_triangles=False))):
        for threads in (1, 8):
            os.environ['CHROMA_UPLOAD_THREADS'] = str(threads)
            ctx.reload_knobs()
            gp = gpu.GPUPhotons(src, **kw)
            gp.propagate(gg, gpu.get_rng_states(1, seed=4), max_steps=10)     # reads the stage as it lands
            out = gp.get()
            if ref is None:
                ref = out
            assert same(out, ref)
    os.environ.pop('CHROMA_UPLOAD_THREADS')
    ctx.reload_knobs()
    # pageable arrays LENT to the engine (borrow=True): staged by a worker thread while propagate already runs
    for src in (ph, slim):
        gp = gpu.GPUPhotons(src, borrow=True)
        gp.propagate(gg, gpu.get_rng_states(1, seed=4), max_steps=10)
        assert same(gp.get(), ref)
    back = gpu.GPUPhotons(ph, borrow=True).get()              # consumed by a download instead (photons_join)
    for f in fields:
        assert np.array_equal(getattr(back, f), getattr(ph, f)), f
    a, b = gpu.GPUPhotons(ph, borrow=True), gpu.GPUPhotons(ph, borrow=True)   # the second upload waits for the first worker
    ctx.synchronize()
    b.propagate(gg, gpu.get_rng_states(1, seed=4), max_steps=10)
    a.propagate(gg, gpu.get_rng_states(1, seed=4), max_steps=10)
    assert same(a.get(), ref) and same(b.get(), ref)
    # a pageable source may be overwritten as soon as the constructor returns
    scratch = event.Photons(*[np.array(getattr(ph, f)) for f in fields])
    gp = gpu.GPUPhotons(scratch)
    scratch.dir[:] = 0
    scratch.wavelengths[:] = 0
    gp.propagate(gg, gpu.get_rng_states(1, seed=4), max_steps=10)
    assert same(gp.get(), ref)


def test_more_than_128_materials_are_refused(ctx, oracle_mod):
    "8-bit signed material index of the triangle's code word (photon.h:68-75): 129 materials cannot be encoded"
    mesh = make.sphere(100.0, nsteps=24)
    mats = []
    for i in range(129):
        m = Material('m%d' % i)
        m.set('refractive_index', 1.0 + 0.001 * i)
        m.set('absorption_length', 1e6)
        m.set('scattering_length', 1e6)
        mats.append(m)
    inner = np.array([mats[i % 129] for i in range(len(mesh.triangles))], dtype=object)
    g = Geometry(mats[0])
    g.add_solid(Solid(mesh, inner, mats[0]))
    F.attach_bvh(g, oracle_mod.build_bvh)
    with pytest.raises(_lib.ChromaError, match='at most 128'):
        gpu.GPUGeometry(g)


def test_ranks_fed_identical_photons_draw_different_streams(ctx, oracle_mod):
    "ADVICE r1: with one seed on every rank, photon i of launch k got the same uniforms on all ranks"
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    ph = F.isotropic_photons(100000, seed=9, wavelength_range=(400, 800))
    outs = [run(gg, ph, seed=dist.seed_for_rank(1, r)) for r in range(3)]
    assert same(outs[0], run(gg, ph, seed=1))                      # rank 0 keeps the seed
    assert (outs[0].flags != outs[1].flags).mean() > 0.2 and (outs[1].flags != outs[2].flags).mean() > 0.2
    # events keyed by their global number: the same event gives the same photons whatever ran before it
    rng = gpu.get_rng_states(1, seed=1)

    def event_out(g):
        rng.epoch = dist.event_epoch(g)
        gp = gpu.GPUPhotons(ph)
        gp.propagate(gg, rng, max_steps=10)
        return gp.get()
    a = event_out(7)
    b = event_out(3)
    assert same(event_out(7), a) and not same(a, b)


def test_photon_fields_as_device_arrays(ctx):
    "chroma/gpu/photon.py:29-46 exposes GPUArrays: get / set / fill / slices and a contiguous device copy"
    n = 5000
    ph = F.isotropic_photons(n, seed=1, wavelength_range=(400, 800), t0=5.0, pos=(1, 2, 3))
    gp = gpu.GPUPhotons(ph)
    for name in ('pos', 'dir', 'pol', 'wavelengths', 't', 'last_hit_triangles', 'flags', 'weights', 'evidx'):
        dev = getattr(gp, name).to_device()
        assert isinstance(dev, gpu.DeviceArray) and np.array_equal(dev.get(), getattr(ph, name)), name
        assert getattr(gp, name).gpudata != 0
    gp.t.fill(7.0)
    gp.pos.set(ph.pos * 2)
    gp.flags[10:20].set(np.full(10, event.CHERENKOV, np.uint32))
    scaled = gp.wavelengths.to_device()                       # a caller's own device-side edit ...
    scaled.set(scaled.get() + 1.0)
    gp.wavelengths.set(scaled)                                # ... written back from the device array
    back = gp.get()
    assert (back.t == 7.0).all() and np.array_equal(back.pos, ph.pos * 2)
    assert (back.flags[10:20] == event.CHERENKOV).all() and (back.flags[:10] == 0).all() and (back.flags[20:] == 0).all()
    assert np.array_equal(back.wavelengths, ph.wavelengths + np.float32(1.0))
    assert np.array_equal(back.dir, ph.dir) and np.array_equal(back.pol, ph.pol)
    with pytest.raises(ValueError):
        gp.t.set(np.zeros(3, np.float32))


def test_flat_geometry_attributes_are_made_on_demand(ctx, oracle_mod):
    """GPUGeometry.nodes / .vertices / .triangles / .material_codes / .solid_id_map (chroma/gpu/geometry.py:243-253):
    the engine does not keep device copies of them (8.9 GB at 29,023 PMTs); reading one makes it"""
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    before = gg.info()['device_bytes']
    nodes = gg.nodes.get()                                      # rebuilt on the device from the engine's copy
    assert np.array_equal(nodes.view(np.uint32).reshape(-1, 4), np.ascontiguousarray(det.bvh.nodes).view(np.uint32).reshape(-1, 4))
    assert gg.info()['device_bytes'] == before + nodes.nbytes
    assert np.array_equal(gg.vertices.get().view(np.float32).reshape(-1, 3), np.asarray(det.mesh.vertices, np.float32))
    assert np.array_equal(gg.triangles.get().view(np.uint32).reshape(-1, 3), np.asarray(det.mesh.triangles, np.uint32))
    assert np.array_equal(gg.solid_id_map.get(), np.asarray(det.solid_id, np.uint32))
    assert np.array_equal(gg.colors.get(), np.asarray(det.colors, np.uint32))
    assert len(gg.material_codes.get()) == len(det.mesh.triangles)
    # the released arrays are not handed out by the library any more
    p, n = C.c_void_p(), C.c_uint64()
    with pytest.raises(_lib.ChromaError, match='released'):
        _lib.call('chroma_geometry_array', gg.handle, 1, C.byref(p), C.byref(n))
