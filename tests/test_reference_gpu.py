"""GPU parity against the REFERENCE's own CUDA kernels (oracle/_ref/*.cubin, built from the
untouched sources under /root/reference by `make -C oracle ref` and launched through the CUDA
driver API by oracle/ref_gpu.py — the reference's PyCUDA launcher is not installed).

Deterministic tier (BASELINE.json north_star): first-hit triangle index AND the bits of the
distance must be identical for every ray — no tolerance.  Same for make_leaves (BVH leaf boxes
+ Morton codes) and bin_hits.  Stochastic tier: the reference draws from per-thread XORWOW
states (chroma/cuda/propagate.cu:241), the engine from per-photon Philox, so agreement is
statistical: history-bit fractions within 4 sigma, detection efficiency within 0.5 % relative
(+ 3 sigma of its binomial error at the test's sample size), per-channel hit-count chi-square
and hit-time KS p > 0.01."""
import numpy as np
import pytest

import fixtures as F
import stats_util as S
from chroma_b200 import event, gpu
from chroma_b200.tools import from_film, argsort_direction

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def rctx(ctx):
    from oracle import ref_gpu
    if not ref_gpu.available():
        pytest.fail('oracle/_ref/*.cubin missing: run `python -c "import __graft_entry__ as g; g.build()"` '
                    'where /root/reference is mounted (the cubins travel with the snapshot)')
    r = ref_gpu.RefContext(0)
    yield r
    r.close()


def exact(ref_gpu_geom, gg, pos, dir, last=None):
    tri_r, dist_r = ref_gpu_geom.intersect(pos, dir, last)
    tri_g, dist_g = gpu.intersect(gg, pos, dir, last)
    bad = np.flatnonzero((tri_r != tri_g) | (dist_r.view(np.uint32) != dist_g.view(np.uint32)))
    assert len(bad) == 0, '%d / %d rays differ; first: ray %d ref (%d, %r) engine (%d, %r)' % (
        len(bad), len(pos), bad[0], tri_r[bad[0]], dist_r[bad[0]], tri_g[bad[0]], dist_g[bad[0]])
    return tri_g, dist_g


def test_intersect_bit_exact_cube_film(ctx, rctx, oracle_mod):
    "the reference's test/test_ray_intersection.py setup: 480,000 film rays vs a 1000 mm cube"
    from oracle import ref_gpu
    geo = F.attach_bvh(F.cube_geometry(), oracle_mod.build_bvh)
    pos, dir = from_film()
    tri, dist = exact(ref_gpu.RefGeometry(rctx, geo), gpu.GPUGeometry(geo), pos.astype(np.float32),
                      dir.astype(np.float32))
    assert (tri >= 0).all()


def test_intersect_bit_exact_detector(ctx, rctx, oracle_mod):
    """config C2 shape (chroma/benchmark.py:22-46) on a small demo detector: 1e6 rays from the
    origin, direction-sorted; then incoherent rays from random origins; then second-bounce rays
    that exclude the triangle just hit (mesh.h:82)."""
    from oracle import ref_gpu
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    rg, gg = ref_gpu.RefGeometry(rctx, det), gpu.GPUGeometry(det)
    pos, dir = F.isotropic_rays(1000000, seed=0)
    dir = dir[argsort_direction(dir)]
    tri, dist = exact(rg, gg, pos, dir)
    assert (tri >= 0).mean() > 0.99
    pos, dir = F.isotropic_rays(300000, seed=1, radius=1500.0)
    tri, dist = exact(rg, gg, pos, dir)
    hitpos = (pos + dir * dist[:, None]).astype(np.float32)
    _, dir2 = F.isotropic_rays(300000, seed=2)
    exact(rg, gg, hitpos, dir2, last=tri)
    # unnormalised, axis-parallel and missing rays
    pos = np.array([[0, 0, 0], [0, 0, 0], [1e5, 1e5, 1e5], [0, 0, 0]], np.float32)
    dir = np.array([[0, 0, 2], [3, 0, 0], [1, 1, 1], [0.3, -0.2, 0.1]], np.float32)
    exact(rg, gg, pos, dir)


def test_make_leaves_bit_exact(ctx, rctx):
    "chroma/cuda/bvh.cu:148-203 vs chroma_bvh_make_leaves: leaf boxes and 48-bit Morton codes"
    from oracle import ref_gpu
    from chroma_b200.gpu.bvh import create_leaf_nodes
    det = F.small_detector()
    det.flatten()
    wc, nodes, codes = create_leaf_nodes(det.mesh)
    rn, rc = ref_gpu.make_leaves(rctx, det.mesh, wc.world_origin, wc.world_scale)
    assert np.array_equal(np.ascontiguousarray(nodes).view(np.uint32).reshape(-1, 4), rn)
    assert np.array_equal(codes, rc)


def propagate_both(rctx, oracle_mod, det, ph, max_steps, seed=5):
    from oracle import ref_gpu
    rg = ref_gpu.RefGeometry(rctx, det)
    rrng = ref_gpu.RefRng(rctx, 64 * 1024, seed=seed)
    rp = ref_gpu.RefPhotons(rctx, ph)
    rp.propagate(rg, rrng, max_steps=max_steps)
    out_r = rp.get()
    gg = gpu.GPUDetector(det)
    rng = gpu.get_rng_states(64 * 1024, seed=seed)
    gp = gpu.GPUPhotons(ph)
    gp.propagate(gg, rng, max_steps=max_steps)
    return out_r, gp.get(), rg, rrng, rp, gg, rng, gp


def check_statistics(out_r, out_g, n):
    for bit, fr, fg, ok in S.fractions_agree(out_r.flags, out_g.flags, nsigma=4.0):
        assert ok, 'history bit %d: reference %.5f engine %.5f' % (bit, fr, fg)
    det_r = ((out_r.flags & event.SURFACE_DETECT) != 0).mean()
    det_g = ((out_g.flags & event.SURFACE_DETECT) != 0).mean()
    tol = 0.005 * det_r + 3 * np.sqrt(2) * S.binomial_sigma(det_r, n)
    assert abs(det_r - det_g) <= tol, 'detection efficiency: reference %.5f engine %.5f' % (det_r, det_g)
    assert ((out_g.flags & event.NAN_ABORT) != 0).sum() == 0
    return det_r, det_g


@S.two_stage
def test_propagate_statistics_demo_detector(ctx, rctx, oracle_mod):
    "config C3 shape on a small demo detector: 4e5-photon bomb, U(400,800) nm, max_steps=10 / 100"
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    n = 400000
    ph = F.isotropic_photons(n, seed=11, wavelength_range=(400, 800), t0=100.0)
    for max_steps in (10, 100):
        out_r, out_g, rg, rrng, rp, gg, rng, gp = propagate_both(rctx, oracle_mod, det, ph, max_steps)
        check_statistics(out_r, out_g, n)
        # per-channel hit counts and hit times of detected photons
        s2c = det.solid_id_to_channel_index
        def hits(out):
            m = ((out.flags & event.SURFACE_DETECT) != 0) & (out.last_hit_triangles >= 0)
            ch = s2c[det.solid_id[out.last_hit_triangles[m]]]
            return ch[ch >= 0], out.t[m][ch >= 0]
        ch_r, t_r = hits(out_r)
        ch_g, t_g = hits(out_g)
        nch = det.num_channels()
        assert S.chi2_counts_p(np.bincount(ch_r, minlength=nch), np.bincount(ch_g, minlength=nch)) > 0.01
        assert S.ks_p(t_r, t_g) > 0.01
        assert S.ks_p(out_r.wavelengths, out_g.wavelengths) > 0.01
        if max_steps == 10:
            # run_daq: the reference kernel and the engine on the SAME (engine) photons.  With
            # weight 1 the set of hit channels and their history bits are deterministic; times
            # and charges are drawn from different generators -> distributions agree.
            from oracle import ref_gpu
            rdaq = ref_gpu.RefDaq(rctx, rg)
            rp2 = ref_gpu.RefPhotons(rctx, out_g)
            rdaq.begin_acquire(); rdaq.acquire(rp2, rrng); rdaq.end_acquire()
            ch_ref = rdaq.get()
            daq = gpu.GPUDaq(gg)
            daq.begin_acquire(); daq.acquire(gp, rng)
            ch_eng = daq.end_acquire().get()
            assert np.array_equal(ch_ref.hit, ch_eng.hit)
            assert np.array_equal(ch_ref.flags, ch_eng.flags)
            assert S.ks_p(ch_ref.t[ch_ref.hit], ch_eng.t[ch_eng.hit]) > 0.01
            assert S.ks_p(ch_ref.q[ch_ref.hit], ch_eng.q[ch_eng.hit]) > 0.01
            rp2.free()
        rp.free()


def test_propagate_statistics_scintillator_wls_dichroic(ctx, rctx, oracle_mod):
    """config C4 shape: bulk reemission, WLS, dichroic and thin-film (complex) surfaces.
    The reference kernel's photon queue is filled through atomics, so which XORWOW state meets which photon
    differs from run to run (and with whatever the session's reference context ran before): each p > 0.01
    assertion fails one time in a hundred by construction.  A first failure is therefore repeated once on a
    fresh photon sample; a real discrepancy fails both (the pair of runs has a false-alarm rate of 1e-4)."""
    det = F.attach_bvh(F.scintillator_detector(), oracle_mod.build_bvh)
    n = 300000

    def compare(attempt):
        rs = np.random.RandomState(3 + attempt)
        ph = F.isotropic_photons(n, seed=21 + attempt, wavelength_range=(350, 500))
        ph.pos[:] = (rs.normal(size=(n, 3)) * 200.0).astype(np.float32)
        out_r, out_g, *_ = propagate_both(rctx, oracle_mod, det, ph, 30)
        check_statistics(out_r, out_g, n)
        for bit in (event.BULK_REEMIT, event.SURFACE_REEMIT, event.SURFACE_TRANSMIT):
            assert ((out_g.flags & bit) != 0).mean() > 0.01   # the fixture exercises these models
        dr = (out_r.flags & event.SURFACE_DETECT) != 0
        dg = (out_g.flags & event.SURFACE_DETECT) != 0
        assert S.ks_p(out_r.wavelengths[dr], out_g.wavelengths[dg]) > 0.01
        assert S.ks_p(out_r.t[dr], out_g.t[dg]) > 0.01
    try:
        compare(0)
    except AssertionError as e:
        print('first sample failed (%s): repeating on a fresh sample' % str(e).split('\n')[0])
        compare(1)


@S.two_stage
def test_propagate_statistics_c4_detector(ctx, rctx, oracle_mod):
    """config C4 at BASELINE.md's size: the 907-module scintillator / WLS-plate / dichroic-disc detector
    (chroma_b200/demo/scintillator.py), bombs at random vertices of the scintillator, U(350,500) nm,
    max_steps=100: the flat engine AND the two-level engine against the reference's propagate.cubin."""
    from chroma_b200.demo import scintillator
    det = F.attach_bvh(scintillator.detector(), oracle_mod.build_bvh)
    assert det.num_channels() == 907
    n = 400000
    parts = [F.isotropic_photons(n // 4, seed=60 + k, wavelength_range=(350, 500), pos=scintillator.event_vertex(k))
             for k in range(4)]
    ph = event.Photons.join(parts)
    out_r, out_g, rg, rrng, rp, gg, rng, gp = propagate_both(rctx, oracle_mod, det, ph, 100)
    rp.free()
    gi = gpu.GPUDetector(det, instanced=True)
    gp2 = gpu.GPUPhotons(ph)
    gp2.propagate(gi, gpu.get_rng_states(1, seed=5), max_steps=100)
    out_i = gp2.get()
    s2c = det.solid_id_to_channel_index
    nch = det.num_channels()

    def hits(out):
        m = ((out.flags & event.SURFACE_DETECT) != 0) & (out.last_hit_triangles >= 0)
        ch = s2c[det.solid_id[out.last_hit_triangles[m]]]
        return ch[ch >= 0], out.t[m][ch >= 0], out.wavelengths[m][ch >= 0]
    ch_r, t_r, w_r = hits(out_r)
    for name, out in (('flat', out_g), ('two-level', out_i)):
        check_statistics(out_r, out, n)
        for bit in (event.BULK_REEMIT, event.SURFACE_REEMIT, event.SURFACE_TRANSMIT, event.REFLECT_SPECULAR):
            assert ((out.flags & bit) != 0).mean() > 0.005, (name, bit)     # the detector exercises these models
        ch_g, t_g, w_g = hits(out)
        assert len(ch_g) > 5000, name
        assert S.chi2_counts_p(np.bincount(ch_r, minlength=nch), np.bincount(ch_g, minlength=nch)) > 0.01, name
        assert S.ks_p(t_r, t_g) > 0.01 and S.ks_p(w_r, w_g) > 0.01, name   # hit times, detected-wavelength spectrum
    # same Philox streams: the two engines give the same history for all but forked trajectories
    assert (out_g.flags == out_i.flags).mean() > 0.98


def test_propagate_vacuum_box_no_abort(ctx, rctx, oracle_mod):
    "test/test_propagation.py:11-58 testAbort: axis-aligned photons in a 100 mm vacuum box"
    geo = F.attach_bvh(F.vacuum_box(100.0), oracle_mod.build_bvh)
    n = 10000
    pos = np.tile([0, 0, 0], (n, 1)).astype(np.float32)
    dir = np.tile([0, 0, 1], (n, 1)).astype(np.float32)
    pol = np.zeros_like(pos)
    phi = np.random.RandomState(0).uniform(0, 2 * np.pi, n).astype(np.float32)
    pol[:, 0], pol[:, 1] = np.cos(phi), np.sin(phi)
    ph = event.Photons(pos, dir, pol, np.full(n, 400.0, np.float32))
    from oracle import ref_gpu
    rg = ref_gpu.RefGeometry(rctx, geo)
    rp = ref_gpu.RefPhotons(rctx, ph)
    rp.propagate(rg, ref_gpu.RefRng(rctx, 64 * 1024, seed=1), max_steps=10)
    out_r = rp.get()
    gp = gpu.GPUPhotons(ph)
    gp.propagate(gpu.GPUGeometry(geo), gpu.get_rng_states(1, seed=1), max_steps=10)
    out_g = gp.get()
    assert ((out_r.flags & event.NAN_ABORT) != 0).sum() == 0
    assert ((out_g.flags & event.NAN_ABORT) != 0).sum() == 0
    assert not np.isnan(out_g.pos).any()
    # nearly all leave the box through the +z face into the void (vacuum absorption/scattering
    # lengths are 1e6 mm, so a handful in 1e4 may interact on the 50 mm path)
    assert (out_r.flags == event.NO_HIT).mean() > 0.998
    assert (out_g.flags == event.NO_HIT).mean() > 0.998


def test_bin_hits_exact(ctx, rctx):
    "chroma/cuda/pdf.cu:9-32 vs GPUPDF.add_hits_to_pdf on the same channel arrays"
    nch, tbins, qbins = 500, 100, 10
    rs = np.random.RandomState(0)
    hc_ref = np.zeros(nch, np.uint32)
    pdf_ref = np.zeros(nch * tbins * qbins, np.uint32)
    d_hc, d_pdf = rctx.to_gpu(hc_ref), rctx.to_gpu(pdf_ref)
    pdf = gpu.GPUPDF()
    pdf.setup_pdf(nch, tbins, (-0.5, 999.5), qbins, (-0.5, 9.5))
    for ev in range(20):
        t = np.where(rs.uniform(size=nch) < 0.6, rs.uniform(-5, 1100, nch), 1e9).astype(np.float32)
        q = rs.uniform(0, 12, nch).astype(np.float32)
        dt, dq = rctx.to_gpu(t), rctx.to_gpu(q)
        rctx.launch('pdf', 'bin_hits', nch // 64 + 1, 64, np.int32(nch), dq, dt, d_hc, np.int32(tbins),
                    np.float32(-0.5), np.float32(999.5), np.int32(qbins), np.float32(-0.5),
                    np.float32(9.5), d_pdf)
        rctx.synchronize()
        dt.free(); dq.free()
        pdf.add_hits_to_pdf(gpu.GPUChannels(gpu.to_gpu(t), gpu.to_gpu(q), gpu.to_gpu(np.zeros(nch, np.uint32))))
    hitcount, hist = pdf.get_pdfs()
    assert np.array_equal(hitcount, rctx.get(d_hc, np.uint32, nch))
    assert np.array_equal(hist.ravel(), rctx.get(d_pdf, np.uint32, nch * tbins * qbins))
    assert np.array_equal(hitcount, hist.reshape(nch, -1).sum(axis=1))   # test/test_pdf.py:51-53
