"""GPU parity of the likelihood PDF-evaluation kernels: the fused
variable-bin accumulation and the moment / kernel-estimate kernels of csrc/engine_pdfeval.cu
against (1) the oracle's two-pass restatement and (2) the reference's own pdf.cu kernels
(oracle/_ref/pdf.cubin) on the same arrays.  Integer counters, the nearest-neighbour tables and
the float moment sums must be identical bit for bit; the Gaussian kernel sums agree within
2e-6 relative with the reference cubin (same fast-math build) and 1e-4 with the double oracle."""
import numpy as np
import pytest

import fixtures as F
from chroma_b200 import event, gpu
from chroma_b200.likelihood import Likelihood
from chroma_b200.sim import Simulation

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def rctx(ctx):
    from oracle import ref_gpu
    if not ref_gpu.available():
        pytest.fail('oracle/_ref/*.cubin missing (built by __graft_entry__.build() where the reference is mounted)')
    r = ref_gpu.RefContext(0)
    yield r
    r.close()


def synthetic_daq(rs, nch, ndaq, hit_fraction=0.7, sigma=3.0):
    t = rs.normal(20.0, sigma, size=(ndaq, nch)).astype(np.float32)
    t[rs.uniform(size=(ndaq, nch)) > hit_fraction] = 1e9
    return t


def channels_of(t, q=None, ndaq=1):
    t = np.ascontiguousarray(t, np.float32)
    nch = t.size // ndaq
    q = np.zeros(t.size, np.float32) if q is None else np.ascontiguousarray(q, np.float32)
    return gpu.GPUChannels(gpu.to_gpu(t.ravel()), gpu.to_gpu(q.ravel()), gpu.to_gpu(np.zeros(t.size, np.uint32)),
                           ndaq=ndaq, stride=nch)


def reference_pdf_eval(rctx, event_hit, event_time, batches, min_twidth, trange, k):
    "chroma/gpu/pdf.py:266-328 on the reference kernels: bincount pass, sync, one block per hit"
    nch, nhit = len(event_hit), int(event_hit.sum())
    hit_to_ch = np.where(event_hit)[0].astype(np.uint32)
    ch_to_hit = np.maximum(0, event_hit.cumsum() - 1).astype(np.uint32)
    d_hit, d_time = rctx.to_gpu(event_hit.astype(np.uint32)), rctx.to_gpu(event_time.astype(np.float32))
    d_h2c, d_c2h = rctx.to_gpu(hit_to_ch), rctx.to_gpu(ch_to_hit)
    d_hc, d_bc = rctx.to_gpu(np.zeros(nch, np.uint32)), rctx.to_gpu(np.zeros(nch, np.uint32))
    d_nn = rctx.to_gpu(np.full(nhit * k, 1e9, np.float32))
    for mc in batches:
        ndaq = mc.shape[0]
        d_mc = rctx.to_gpu(mc.ravel())
        d_q = rctx.to_gpu(np.ones(nhit * (ndaq + 1), np.uint32))
        rctx.launch('pdf', 'accumulate_bincount', nch // 64 + 1, 64, np.int32(nch), np.int32(ndaq), d_hit,
                    d_time, d_mc, d_hc, d_bc, np.float32(min_twidth), np.float32(trange[0]),
                    np.float32(trange[1]), np.int32(k), d_c2h, d_q)
        rctx.synchronize()
        rctx.launch('pdf', 'accumulate_nearest_neighbor_block', nhit, 64, np.int32(nhit), np.int32(ndaq),
                    d_h2c, d_q, d_time, d_mc, d_nn, np.int32(k))
        rctx.synchronize()
        d_mc.free(); d_q.free()
    return (rctx.get(d_hc, np.uint32, nch), rctx.get(d_bc, np.uint32, nch),
            rctx.get(d_nn, np.float32, nhit * k).reshape(nhit, k))


@pytest.mark.parametrize('k,ndaq,nch', [(1, 1, 300), (20, 128, 1000), (320, 64, 500), (7, 33, 65)])
def test_pdf_eval_exact_vs_oracle_and_reference(ctx, rctx, oracle_mod, k, ndaq, nch):
    rs = np.random.RandomState(k)
    event_hit = rs.uniform(size=nch) < 0.6
    event_hit[0], event_hit[-1] = False, True
    event_time = rs.normal(20.0, 3.0, nch).astype(np.float32)
    event_time[5] = 500.0          # observed far outside the Monte Carlo: never fills the bin
    trange, min_twidth = (10.0, 28.0), 0.4
    batches = [synthetic_daq(rs, nch, ndaq) for _ in range(6)]
    batches[2][:] = 1e9            # a DAQ pass with no hit at all
    pdf = gpu.GPUPDF()
    pdf.setup_pdf_eval(event_hit, event_time, np.zeros(nch), min_twidth, trange, 1.0, (0, 10), min_bin_content=k)
    hc_o, bc_o = np.zeros(nch, np.uint32), np.zeros(nch, np.uint32)
    nn_o = np.full((int(event_hit.sum()), k), 1e9, np.float32)
    for mc in batches:
        pdf.accumulate_pdf_eval(channels_of(mc, ndaq=ndaq))
        oracle_mod.accumulate_pdf_eval(event_hit, event_time, mc, hc_o, bc_o, nn_o, min_twidth, trange, k)
    hc, bc, nn = pdf.get_pdf_eval_arrays()
    assert np.array_equal(hc, hc_o) and np.array_equal(bc, bc_o)
    assert np.array_equal(nn.view(np.uint32), nn_o.view(np.uint32))
    if k + ndaq <= 1000:           # the reference's shared table holds 1000 floats (pdf.cu:160)
        hc_r, bc_r, nn_r = reference_pdf_eval(rctx, event_hit, event_time, batches, min_twidth, trange, k)
        assert np.array_equal(hc, hc_r) and np.array_equal(bc, bc_r)
        assert np.array_equal(nn.view(np.uint32), nn_r.view(np.uint32))
    assert hc.max() > 0 and (np.diff(nn, axis=1) >= 0).all()
    # host tail: values positive exactly where the event was hit and Monte Carlo arrived
    hitcount, value, uncert = pdf.get_pdf_eval()
    assert np.array_equal(hitcount, hc)
    assert (value[~event_hit] == 0).all() and (value[event_hit & (hc > 0)] > 0).all()
    assert np.isfinite(value).all() and np.isfinite(uncert).all()
    # clear_pdf_eval resets everything
    pdf.clear_pdf_eval()
    hc, bc, nn = pdf.get_pdf_eval_arrays()
    assert hc.sum() == 0 and bc.sum() == 0 and (nn == np.float32(1e9)).all()


def test_pdf_eval_estimates_the_density(ctx):
    "with enough Monte Carlo both regimes reproduce the N(20,3) density at the observed time"
    rs = np.random.RandomState(3)
    nch, ndaq = 64, 128
    event_hit = np.ones(nch, bool)
    event_time = np.linspace(14.0, 26.0, nch).astype(np.float32)
    pdf = gpu.GPUPDF()
    # fixed 0.5 ns bin that fills up / a bin too narrow to fill: k-th nearest distance decides
    for width, k, nbatch in ((0.5, 50, 40), (0.01, 30, 4)):
        pdf.setup_pdf_eval(event_hit, event_time, np.zeros(nch), width, (0.0, 40.0), 1.0, (0, 10), min_bin_content=k)
        for _ in range(nbatch):
            pdf.accumulate_pdf_eval(channels_of(synthetic_daq(rs, nch, ndaq, hit_fraction=1.0), ndaq=ndaq))
        hitcount, value, uncert = pdf.get_pdf_eval()
        assert (hitcount == nbatch * ndaq).all()
        expected = np.exp(-0.5 * ((event_time - 20.0) / 3.0) ** 2) / (3.0 * np.sqrt(2 * np.pi))
        pull = (value - expected) / np.maximum(uncert, 1e-12)
        assert (value > 0).all() and np.abs(pull).max() < 5.0 and abs(pull.mean()) < 1.0, (k, pull)


def test_pdf_eval_no_hit_event_and_argument_errors(ctx):
    nch = 40
    pdf = gpu.GPUPDF()
    pdf.setup_pdf_eval(np.zeros(nch, bool), np.zeros(nch), np.zeros(nch), 0.5, (0.0, 40.0), 1.0, (0, 10))
    pdf.accumulate_pdf_eval(channels_of(np.full((4, nch), 20.0, np.float32), ndaq=4))
    hitcount, value, uncert = pdf.get_pdf_eval()
    assert (hitcount == 4).all() and (value == 0).all()
    with pytest.raises(ValueError):
        pdf.accumulate_pdf_eval(channels_of(np.zeros((2, nch + 1), np.float32), ndaq=2))


@pytest.mark.parametrize('time_only', [True, False])
def test_kernel_pdf_vs_oracle_and_reference(ctx, rctx, oracle_mod, time_only):
    rs = np.random.RandomState(11)
    nch, nmc = 777, 30
    trange, qrange = (10.0, 28.0), (0.4, 1.8)
    event_hit = (rs.uniform(size=nch) < 0.7)
    event_time = rs.normal(20.0, 3.0, nch).astype(np.float32)
    event_charge = rs.normal(1.0, 0.3, nch).astype(np.float32)
    kp = gpu.GPUKernelPDF()
    kp.setup_moments(nch, trange, qrange, time_only=time_only)
    mom_o = [np.zeros(nch, np.uint32)] + [np.zeros(nch, np.float32) for _ in range(4)]
    d_mom = [rctx.to_gpu(a) for a in mom_o]
    draws = [(synthetic_daq(rs, nch, 1)[0], rs.normal(1.0, 0.3, nch).astype(np.float32)) for _ in range(nmc)]
    for t, q in draws:
        kp.accumulate_moments(channels_of(t, q))
        oracle_mod.accumulate_moments(time_only, t, q, trange, qrange, *mom_o)
        dt, dq = rctx.to_gpu(t), rctx.to_gpu(q)
        rctx.launch('pdf', 'accumulate_moments', nch // 64 + 1, 64, np.int32(time_only), np.int32(nch), dt, dq,
                    np.float32(trange[0]), np.float32(trange[1]), np.float32(qrange[0]), np.float32(qrange[1]),
                    *d_mom)
        rctx.synchronize()
        dt.free(); dq.free()
    got = kp.get_moments()
    for g, o, d in zip(got, mom_o, d_mom):
        assert np.array_equal(g.view(np.uint32), o.view(np.uint32))
        assert np.array_equal(g.view(np.uint32), rctx.get(d, g.dtype, nch).view(np.uint32))
    assert got[0].max() > 0
    kp.clear_moments()
    assert kp.get_moments()[0].sum() == 0
    for t, q in draws:
        kp.accumulate_moments(channels_of(t, q))

    kp.compute_bandwidth(event_hit, event_time, event_charge, scale_factor=1.0)
    itb, iqb = kp.inv_time_bandwidths, kp.inv_charge_bandwidths
    assert np.isfinite(itb).all() and (itb >= 0).all() and (itb > 0).any()
    iqb = np.nan_to_num(iqb, nan=0.0, posinf=0.0, neginf=0.0).astype(np.float32)
    kp.inv_charge_bandwidths = iqb
    kp.setup_kernel(event_hit, event_time, event_charge)
    hc_o, tp_o, qp_o = np.zeros(nch, np.uint32), np.zeros(nch, np.float32), np.zeros(nch, np.float32)
    d_hc, d_tp, d_qp = rctx.to_gpu(hc_o), rctx.to_gpu(tp_o), rctx.to_gpu(qp_o)
    d_ev = [rctx.to_gpu(event_hit.astype(np.uint32)), rctx.to_gpu(event_time), rctx.to_gpu(event_charge)]
    d_bw = [rctx.to_gpu(itb), rctx.to_gpu(iqb)]
    for t, q in draws:
        kp.accumulate_kernel(channels_of(t, q))
        oracle_mod.accumulate_kernel_eval(time_only, event_hit, event_time, event_charge, t, q, trange, qrange,
                                          itb, iqb, hc_o, tp_o, qp_o)
        dt, dq = rctx.to_gpu(t), rctx.to_gpu(q)
        rctx.launch('pdf', 'accumulate_kernel_eval', nch // 64 + 1, 64, np.int32(time_only), np.int32(nch),
                    d_ev[0], d_ev[1], d_ev[2], dt, dq, np.float32(trange[0]), np.float32(trange[1]),
                    np.float32(qrange[0]), np.float32(qrange[1]), d_bw[0], d_bw[1], d_hc, d_tp, d_qp)
        rctx.synchronize()
        dt.free(); dq.free()
    hc, tp, qp = kp.get_kernel_sums()
    assert np.array_equal(hc, hc_o) and np.array_equal(hc, rctx.get(d_hc, np.uint32, nch))
    tp_r, qp_r = rctx.get(d_tp, np.float32, nch), rctx.get(d_qp, np.float32, nch)
    np.testing.assert_allclose(tp, tp_r, rtol=2e-6, atol=0)
    np.testing.assert_allclose(qp, qp_r, rtol=2e-6, atol=0)
    np.testing.assert_allclose(tp, tp_o, rtol=1e-4, atol=1e-30)
    np.testing.assert_allclose(qp, qp_o, rtol=1e-4, atol=1e-30)
    assert (tp[~event_hit] == 0).all() and (tp[event_hit & (hc > 0)] > 0).any()
    if time_only:
        assert (qp == 0).all()
    hitcount, value, zeros = kp.get_kernel_eval()
    assert np.array_equal(hitcount, hc) and (zeros == 0).all() and np.isfinite(value).all()
    kp.clear_kernel()
    assert all(a.sum() == 0 for a in kp.get_kernel_sums())


def test_kernel_pdf_requires_setup(ctx):
    kp = gpu.GPUKernelPDF()
    kp.setup_moments(10, (0, 1), (0, 1))
    with pytest.raises(RuntimeError):
        kp.setup_kernel(np.ones(10), np.zeros(10), np.zeros(10))
    with pytest.raises(RuntimeError):     # C-ABI state check
        kp.accumulate_kernel(channels_of(np.zeros(10, np.float32)))


@pytest.fixture(scope='module')
def sim_and_event(ctx, oracle_mod):
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    sim = Simulation(det, seed=3)
    truth = F.isotropic_photons(20000, seed=1, wavelength_range=(400, 600), t0=0.0)
    ev = next(sim.simulate([truth], run_daq=True, keep_photons_beg=True, max_steps=100))
    assert ev.channels.hit.sum() > 10
    return det, sim, ev


def bombs(n, pos, count, seed):
    return (F.isotropic_photons(n, seed=seed + i, wavelength_range=(400, 600), pos=pos, t0=0.0)
            for i in range(count))


def test_sim_eval_pdf_and_likelihood(sim_and_event):
    "Simulation.eval_pdf (sim.py:219-283) and Likelihood.eval: the true vertex beats a displaced one"
    det, sim, ev = sim_and_event
    hit = ev.channels.hit
    hitcount, value, uncert = sim.eval_pdf(ev.channels, bombs(20000, (0, 0, 0), 2, 100), 0.5, (-10.0, 100.0), 1,
                                           (-0.5, 20.0), min_bin_content=20, nreps=2, ndaq=128, nscatter=2)
    assert hitcount.shape == (det.num_channels(),) and hitcount.sum() > 0
    assert np.isfinite(value).all() and (value[hit & (hitcount > 0)] > 0).all() and (value[~hit] == 0).all()
    assert (uncert >= 0).all()
    like = Likelihood(sim, ev, trange=(-10.0, 100.0))
    nll_true = like.eval(bombs(20000, (0, 0, 0), 2, 200), 2, nreps=2, ndaq=128)
    nll_far = like.eval(bombs(20000, (1200.0, 0, 0), 2, 300), 2, nreps=2, ndaq=128)
    assert np.isfinite(nll_true.nominal_value) and np.isfinite(nll_far.nominal_value)
    assert nll_true.nominal_value < nll_far.nominal_value


def test_sim_kernel_likelihood(sim_and_event):
    "Simulation.setup_kernel / eval_kernel (sim.py:285-343) through Likelihood"
    det, sim, ev = sim_and_event
    like = Likelihood(sim, ev, trange=(-10.0, 100.0))
    like.setup_kernel(bombs(20000, (0, 0, 0), 4, 400), 2, nreps=1, ndaq=4, oversample_factor=2)
    assert (sim.gpu_pdf_kernel.inv_time_bandwidths > 0).any()
    nll_true = like.eval_kernel(bombs(20000, (0, 0, 0), 6, 500), 2, nreps=1, ndaq=4, navg=3)
    assert np.isfinite(nll_true.nominal_value) and nll_true.std_dev() >= 0
    hitcount, value, zeros = sim.eval_kernel(ev.channels, bombs(20000, (0, 0, 0), 2, 600), (-10.0, 100.0),
                                             (-0.5, 49.5), nreps=1, ndaq=4)
    assert hitcount.sum() > 0 and (value[ev.channels.hit & (hitcount > 0)] >= 0).all()
