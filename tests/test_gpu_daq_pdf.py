"""GPU parity: run_daq, hit compaction, bin_hits and the GPUPhotons/Simulation API."""
import numpy as np
import pytest

import fixtures as F
import stats_util as S
from chroma_b200 import _lib, event, gpu
from chroma_b200.sim import Simulation

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def setup(ctx, oracle_mod):
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    orc = oracle_mod.Oracle(det)
    ph = F.isotropic_photons(100000, seed=5, wavelength_range=(400, 800), t0=100.0)
    rng = gpu.get_rng_states(1, seed=12)
    gp = gpu.GPUPhotons(ph)
    gp.propagate(gg, rng, max_steps=100)   # epoch 0
    return det, gg, orc, gp, rng


def test_run_daq_vs_oracle(setup, oracle_mod):
    det, gg, orc, gp, rng = setup
    end = gp.get()
    daq = gpu.GPUDaq(gg)
    daq.begin_acquire()
    daq.acquire(gp, rng)                    # epoch 1
    ch = daq.end_acquire().get()
    ref = orc.run_daq(end, seed=12, epoch=1)
    assert np.array_equal(ch.hit, ref.hit)
    assert ch.hit.sum() > 0.9 * det.num_channels()
    assert np.array_equal(ch.flags, ref.flags)
    # same streams: times agree to float rounding of the fast-math interpolation, charges to
    # one quantum (charge_unit = cdf_x[-1]/2**16) per contributing photon
    assert np.abs(ch.t[ch.hit] - ref.t[ref.hit]).max() < 1e-3
    assert np.abs(ch.q - ref.q).max() < 1e-3 * max(1.0, ref.q.max())
    assert (ch.t[~ch.hit] == np.float32(1e9)).all()
    # weight = 0.5 halves the accepted photons
    daq.begin_acquire()
    daq.acquire(gp, rng, weight=0.5)
    half = daq.end_acquire().get()
    assert abs(half.q.sum() / ch.q.sum() - 0.5) < 0.02
    # windowed acquisition only sees its photons
    daq.begin_acquire()
    daq.acquire(gp, rng, start_photon=0, nphotons=10)
    few = daq.end_acquire().get()
    assert few.hit.sum() <= 10


@pytest.mark.parametrize('ndaq', [2, 8, 33, 128])
def test_run_daq_many_vs_oracle(setup, oracle_mod, ndaq):
    """daq.cu:88-150 with ndaq replicas: same Philox streams as the oracle port (which equals the
    reference's own run_daq_many source, tests/test_oracle_vs_ref_cpu.py::test_run_daq_many)."""
    det, gg, orc, gp, rng = setup
    end = gp.get()
    daq = gpu.GPUDaq(gg, ndaq=ndaq)
    for weight in (1.0, 0.4):
        daq.begin_acquire()
        epoch = rng.epoch
        daq.acquire(gp, rng, weight=weight)
        chans = daq.end_acquire()
        assert chans.ndaq == ndaq and chans.stride == det.num_channels()
        ch = chans.get()
        t, q, h = [a.reshape(ndaq, -1) for a in (ch.t, ch.q, ch.flags)]
        to, qo, ho = orc.run_daq_many(end, seed=12, ndaq=ndaq, epoch=epoch, weight=weight)
        assert np.array_equal(t < 1e8, to < 1e8) and np.array_equal(h, ho)
        hit = to < 1e8
        assert hit.sum() > 0.5 * hit.size * weight
        # fast-math logf / sinf in the Box-Muller term and in the CDF interpolation
        assert np.abs(t[hit] - to[hit]).max() < 2e-3
        assert np.abs(q - qo).max() < 1e-3 * max(1.0, qo.max())
        if weight == 1.0:
            assert all(np.array_equal(hit[0], hit[k]) for k in range(1, ndaq))
            # the replicas smear independently: time differences of two replicas ~ N(0, 2 + 2 sigma_cdf^2)
            d = (t[0] - t[1])[hit[0]]
            assert 1.0 < d.std() < 4.0 and abs(d.mean()) < 5 * d.std() / np.sqrt(len(d))
    # iterate_copies yields the replicas as GPUChannels views (chroma/gpu/daq.py:22-26)
    copies = list(chans.iterate_copies())
    assert len(copies) == ndaq and np.array_equal(copies[ndaq - 1].get().t, t[ndaq - 1])


def test_run_daq_many_vs_reference_kernel(setup, ctx):
    """The engine's replicas against the reference's own run_daq_many (oracle/_ref/daq.cubin) on the same
    detected photons: hit pattern and history words identical per replica at weight 1 (every replica
    accepts every detected photon), times / charges from different generators -> KS; at weight 0.4 the
    accepted fraction agrees."""
    from oracle import ref_gpu
    if not ref_gpu.available():
        pytest.fail('oracle/_ref/*.cubin missing')
    det, gg, orc, gp, rng = setup
    end = gp.get()
    ndaq = 16
    rc = ref_gpu.RefContext(0)
    try:
        rg = ref_gpu.RefGeometry(rc, det)
        rrng = ref_gpu.RefRng(rc, 64 * 1024, seed=3)
        rp = ref_gpu.RefPhotons(rc, end)
        rdaq = ref_gpu.RefDaq(rc, rg, ndaq=ndaq)
        daq = gpu.GPUDaq(gg, ndaq=ndaq)
        unit = np.float32(det.charge_cdf[0][-1] / 2 ** 16)
        for weight in (1.0, 0.4):
            rdaq.begin_acquire(); rdaq.acquire(rp, rrng, weight=weight); rdaq.end_acquire()
            tr, qr, qir, hr = rdaq.get_many()
            daq.begin_acquire(); daq.acquire(gp, rng, weight=weight)
            ch = daq.end_acquire().get()
            t, q, h = [a.reshape(ndaq, -1) for a in (ch.t, ch.q, ch.flags)]
            # reference quirk: only replica 0's charges are converted (daq.cu:161-172); compare the ints
            assert (qr[1:] == 0).all() and np.array_equal(qr[0], qir[0] * unit)
            q_ref = qir * unit
            if weight == 1.0:
                assert np.array_equal(t < 1e8, tr < 1e8) and np.array_equal(h, hr)
            else:
                nr, ne = (tr < 1e8).sum(), (t < 1e8).sum()
                assert abs(nr - ne) < 5 * np.sqrt(nr + ne)
                assert abs(q_ref.sum() / q.sum() - 1.0) < 0.02
            hit_r, hit_e = tr < 1e8, t < 1e8
            assert S.ks_p(tr[hit_r], t[hit_e]) > 0.01
            assert S.ks_p(q_ref[hit_r], q[hit_e]) > 0.01
            # spread between replicas of one channel (the curand_normal term + the time CDF)
            both = hit_r[0] & hit_r[1] & hit_e[0] & hit_e[1]
            assert S.ks_p((tr[0] - tr[1])[both], (t[0] - t[1])[both]) > 0.01
    finally:
        rc.close()


def test_daq_time_charge_distributions(ctx, oracle_mod):
    "test/test_detector.py: hit-time std and charge mean/std follow the detector CDFs"
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    det.set_time_dist_gaussian(1.2, -6.0, 6.0)
    det.set_charge_dist_gaussian(1.0, 0.1, 0.5, 1.5)
    gg = gpu.GPUDetector(det)
    rng = gpu.get_rng_states(1, seed=1)
    # photons already "detected" on photocathode triangles, all at t = 0
    cathode = np.flatnonzero((det.surface_index >= 0) & (det.solid_id > 0))
    n = 20000
    tris = np.random.RandomState(0).choice(cathode, n)
    ph = event.Photons(np.zeros((n, 3)), np.tile([0, 0, 1.0], (n, 1)), np.tile([1.0, 0, 0], (n, 1)),
                       np.full(n, 400.0), t=np.full(n, 100.0), last_hit_triangles=tris,
                       flags=np.full(n, event.SURFACE_DETECT, np.uint32))  # t0 > 0: daq.cu:5-12 quirk
    gp = gpu.GPUPhotons(ph)
    daq = gpu.GPUDaq(gg)
    ts, qs = [], []
    for k in range(0, n, 1):
        if k >= 400:
            break
        daq.begin_acquire()
        daq.acquire(gp, rng, start_photon=k, nphotons=1)
        ch = daq.end_acquire().get()
        assert ch.hit.sum() == 1
        ts.append(ch.t[ch.hit][0]); qs.append(ch.q[ch.hit][0])
    assert abs(np.std(ts) - 1.2) < 0.15 and abs(np.mean(ts) - 100.0) < 0.3
    assert abs(np.mean(qs) - 1.0) < 0.03 and abs(np.std(qs) - 0.1) < 0.03


def test_flat_hits_select_gather(setup):
    det, gg, orc, gp, rng = setup
    end = gp.get()
    det_mask = (end.flags & event.SURFACE_DETECT) != 0
    solid = det.solid_id[np.maximum(end.last_hit_triangles, 0)]
    chan = np.where(end.last_hit_triangles >= 0, det.solid_id_to_channel_index[solid], -1)
    expect = np.flatnonzero(det_mask & (chan >= 0))
    hits = gp.get_flat_hits(gg)
    assert len(hits) == len(expect)
    assert np.array_equal(hits.pos, end.pos[expect])          # ascending photon order
    assert np.array_equal(hits.channel, chan[expect].astype(np.uint32))
    assert np.array_equal(hits.t, end.t[expect])
    hm = gp.get_hits(gg)
    assert sum(len(v) for v in hm.values()) == len(expect)
    sel = gp.select(event.RAYLEIGH_SCATTER)
    want = np.flatnonzero((end.flags & event.RAYLEIGH_SCATTER) != 0)
    assert len(sel) == len(want)
    assert np.array_equal(sel.get().wavelengths, end.wavelengths[want])
    sub = gp.select(event.SURFACE_DETECT, start_photon=1000, nphotons=5000)
    assert len(sub) == det_mask[1000:6000].sum()
    ids = np.array([5, 3, 99999, 0], dtype=np.uint32)
    q = gp.copy_queue(ids, len(ids)).get()
    assert np.array_equal(q.pos, end.pos[ids]) and np.array_equal(q.flags, end.flags[ids])
    assert len(gp.select(1 << 20)) == 0


def test_ncopies_and_views(ctx, oracle_mod):
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    ph = F.isotropic_photons(1000, seed=6, wavelength=420.0)
    ph.flags[:] = event.CHERENKOV
    gp = gpu.GPUPhotons(ph, ncopies=3)
    assert len(gp) == 3000 and gp.true_nphotons == 1000
    start = gp.get()
    assert np.array_equal(start.pos[1000:2000], start.pos[:1000])
    assert (start.flags == event.CHERENKOV).all()
    gp.propagate(gg, gpu.get_rng_states(1, seed=3), max_steps=20)
    copies = [c.get() for c in gp.iterate_copies()]
    assert len(copies) == 3 and all(len(c) == 1000 for c in copies)
    assert not np.array_equal(copies[0].flags, copies[1].flags)   # independent random streams
    assert ((copies[0].flags & event.CHERENKOV) != 0).all()       # generator bits survive
    assert len(gp.pos[10:20]) == 10 and gp.t[2000:].get().shape == (1000,)
    g2 = gpu.GPUPhotons(ph, copy_flags=False, copy_triangles=False, copy_weights=False).get()
    assert (g2.flags == 0).all() and (g2.last_hit_triangles == -1).all() and (g2.weights == 1).all()


def test_tracking(ctx, oracle_mod):
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    ph = F.isotropic_photons(2000, seed=7, wavelength=500.0)
    gp = gpu.GPUPhotons(ph)
    ids, steps = gp.propagate(gg, gpu.get_rng_states(1, seed=3), max_steps=5, track=True)
    assert len(ids) == len(steps) and len(ids[0]) == 2000
    assert all(len(i) == len(s) for i, s in zip(ids, steps))
    assert len(ids[-1]) <= len(ids[1])
    assert np.array_equal(steps[0].pos, ph.pos)


def test_pdf_accumulation_invariants(ctx, oracle_mod):
    "test/test_pdf.py: hitcount[i] == pdf[i].sum(); bin_hits equals the oracle bin by bin"
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    sim = Simulation(det, seed=3)
    nch = det.num_channels()
    events = [F.isotropic_photons(20000, seed=100 + k, wavelength_range=(400, 800), t0=20.0) for k in range(5)]
    trange, qrange = (-0.5, 199.5), (-0.5, 9.5)
    hitcount, pdf = sim.create_pdf(events, 100, trange, 10, qrange)
    assert pdf.shape == (nch, 100, 10) and hitcount.sum() > 0
    assert np.array_equal(hitcount, pdf.reshape(nch, -1).sum(axis=1))
    assert hitcount.max() <= 5
    # replay the same channel arrays through the oracle's bin_hits
    hc2, pdf2 = np.zeros(nch, np.uint32), np.zeros((nch, 100, 10), np.uint32)
    sim.gpu_pdf.clear_pdf()
    rng = sim.rng_states
    for ph in events:
        gp = gpu.GPUPhotons(ph)
        gp.propagate(sim.gpu_geometry, rng)
        sim.gpu_daq.begin_acquire()
        sim.gpu_daq.acquire(gp, rng)
        chans = sim.gpu_daq.end_acquire()
        sim.gpu_pdf.add_hits_to_pdf(chans)
        oracle_mod.bin_hits(chans.get(), hc2, pdf2, trange, qrange)
    hc3, pdf3 = sim.gpu_pdf.get_pdfs()
    assert np.array_equal(hc3, hc2) and np.array_equal(pdf3, pdf2)


def test_simulation_end_to_end(ctx, oracle_mod):
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    sim = Simulation(det, seed=9)
    bombs = [F.isotropic_photons(30000, seed=k, wavelength_range=(400, 800), t0=50.0) for k in range(3)]
    evs = list(sim.simulate(bombs, keep_photons_beg=True, keep_photons_end=True, run_daq=True,
                            max_steps=100, photons_per_batch=50000))
    assert [ev.id for ev in evs] == [0, 1, 2]
    for ev in evs:
        assert len(ev.photons_end) == 30000 and ev.channels is not None
        ndet = ((ev.photons_end.flags & event.SURFACE_DETECT) != 0).sum()
        assert len(ev.flat_hits) == ndet
        assert sum(len(h) for h in ev.hits.values()) == ndet
        hit_ch = set(np.unique(ev.flat_hits.channel).tolist())
        assert set(np.flatnonzero(ev.channels.hit).tolist()) <= hit_ch
        assert ev.channels.hit.sum() > 0.5 * len(hit_ch)
        assert (ev.channels.t[ev.channels.hit] > 40).all()
    single = list(sim.simulate(bombs[0], run_daq=True, max_steps=10))
    assert len(single) == 1 and single[0].channels.hit.any()


def test_daq_pack_and_merge_gathered_equal_one_daq(ctx, oracle_mod):
    """chroma_daq_pack / chroma_daq_merge_gathered (one event over several GPUs): three "ranks" run
    run_daq on a third of the photons each; merging their packed accumulators gives, bit for bit, the
    channels of one DAQ that saw all the photons (the streams are keyed by photon id and call epoch)."""
    import ctypes as C
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    n = 300000
    ph = F.isotropic_photons(n, seed=31, wavelength_range=(400, 800))
    gp = gpu.GPUPhotons(ph)
    rng = gpu.get_rng_states(1, seed=9)
    gp.propagate(gg, rng, max_steps=10)
    epoch = 77

    def acquire(d, start, count):
        rng.epoch = epoch                       # every share draws from the streams of the one event
        d.begin_acquire()
        d.acquire(gp, rng, start_photon=start, nphotons=count)
    whole = gpu.GPUDaq(gg)
    acquire(whole, 0, n)
    want = whole.end_acquire().get()
    nranks, nch = 3, det.num_channels()
    bounds = [0, 90001, 200003, n]
    gathered = gpu.DeviceArray((nranks, 3, nch), np.uint32, ctx=ctx)
    parts = []
    for r in range(nranks):
        d = gpu.GPUDaq(gg)
        acquire(d, bounds[r], bounds[r + 1] - bounds[r])
        _lib.call('chroma_daq_pack', d.handle, C.c_void_p(gathered.ptr + r * 3 * nch * 4))
        parts.append(d)
    # the packed layout is (time bits, charge ints, history words)
    g = gathered.get()
    assert np.array_equal(g[1, 0], parts[1].earliest_time_int_gpu.get())
    assert np.array_equal(g[2, 2], parts[2].channel_history_gpu.get())
    merged = parts[0]
    _lib.call('chroma_daq_merge_gathered', merged.handle, C.c_void_p(gathered.ptr), nranks)
    got = merged.end_acquire().get()
    assert np.array_equal(got.hit, want.hit) and want.hit.sum() > 10
    assert np.array_equal(got.flags, want.flags)
    assert np.array_equal(got.t.view(np.uint32), want.t.view(np.uint32))
    assert np.array_equal(got.q.view(np.uint32), want.q.view(np.uint32))
