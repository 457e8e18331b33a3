"""CPU tests of the oracle's likelihood-evaluation restatement (pdf.cu:34-368) and of the host
tail of GPUPDF.get_pdf_eval / GPUKernelPDF.compute_bandwidth, against numpy written from the
reference's documented behaviour (chroma/gpu/pdf.py:229-372)."""
import numpy as np
import pytest

from chroma_b200.gpu.pdf import pdf_eval_values


@pytest.fixture(scope='module')
def O():
    from oracle import oracle
    oracle.load()
    return oracle


def synthetic_daq(rng, nch, ndaq, hit_fraction=0.7):
    t = rng.normal(20.0, 3.0, size=(ndaq, nch)).astype(np.float32)
    t[rng.random((ndaq, nch)) > hit_fraction] = 1e9
    return t


def brute_force(event_hit, event_time, batches, min_twidth, trange, k):
    "Per channel, sequentially: what the running-bincount rule keeps, then the k smallest."
    nch = len(event_hit)
    hitcount, bincount = np.zeros(nch, np.uint32), np.zeros(nch, np.uint32)
    kept = [[] for _ in range(nch)]
    for mc in batches:
        for ch in range(nch):
            hc, bc = np.float32(hitcount[ch]), np.float32(bincount[ch])
            new = []
            for t in mc[:, ch]:
                if t >= 1e8 or t < trange[0] or t > trange[1]:
                    continue
                hc += 1
                if not event_hit[ch]:
                    continue
                d = np.abs(np.float32(t) - np.float32(event_time[ch]))
                if d < min_twidth / 2.0:
                    bc += 1
                if bc < k:
                    new.append(d)
            hitcount[ch], bincount[ch] = hc, bc
            kept[ch] = sorted(kept[ch] + new)[:k]
    nearest = np.full((int(np.count_nonzero(event_hit)), k), 1e9, np.float32)
    for row, ch in enumerate(np.where(event_hit)[0]):
        nearest[row, :len(kept[ch])] = kept[ch]
    return hitcount, bincount, nearest


@pytest.mark.parametrize('k,ndaq', [(1, 1), (5, 8), (10, 64), (50, 3)])
def test_oracle_pdf_eval_matches_brute_force(O, k, ndaq):
    rng = np.random.default_rng(100 + k)
    nch = 97
    event_hit = rng.random(nch) < 0.6
    event_hit[0], event_hit[-1] = False, True
    event_time = rng.normal(20.0, 3.0, nch).astype(np.float32)
    trange, min_twidth = (10.0, 28.0), 0.5
    batches = [synthetic_daq(rng, nch, ndaq) for _ in range(4)]
    hitcount, bincount = np.zeros(nch, np.uint32), np.zeros(nch, np.uint32)
    nearest = np.full((int(event_hit.sum()), k), 1e9, np.float32)
    for mc in batches:
        O.accumulate_pdf_eval(event_hit, event_time, mc, hitcount, bincount, nearest, min_twidth, trange, k)
    hc, bc, nn = brute_force(event_hit, event_time, batches, min_twidth, trange, k)
    np.testing.assert_array_equal(hitcount, hc)
    np.testing.assert_array_equal(bincount, bc)
    np.testing.assert_array_equal(nearest, nn)
    assert (np.diff(nearest, axis=1) >= 0).all()
    assert (bincount[~event_hit] == 0).all()


def test_pdf_eval_no_hits_and_empty_daq(O):
    nch, k = 16, 4
    event_hit = np.zeros(nch, bool)
    hitcount, bincount = np.zeros(nch, np.uint32), np.zeros(nch, np.uint32)
    nearest = np.full((0, k), 1e9, np.float32)
    mc = np.full((2, nch), 1e9, np.float32)
    mc[0, 3] = 5.0
    O.accumulate_pdf_eval(event_hit, np.zeros(nch, np.float32), mc, hitcount, bincount, nearest, 1.0, (0, 10), k)
    assert hitcount.sum() == 1 and hitcount[3] == 1 and bincount.sum() == 0
    _, value, uncert = pdf_eval_values(event_hit, hitcount, bincount, nearest, np.zeros(0, np.uint32), 1.0, k)
    assert (value == 0).all() and (uncert == 0).all()


def test_pdf_eval_values_high_and_low_stats():
    "get_pdf_eval tail (pdf.py:336-372): fixed bin when full, k-NN distance otherwise"
    k, w = 4, 0.5
    evhit = np.array([True, True, True, False])
    hitcount = np.array([100, 50, 0, 7], np.uint32)
    bincount = np.array([8, 2, 0, 0], np.uint32)
    nearest = np.array([[0.01, 0.02, 0.03, 0.04], [0.1, 0.4, 1e9, 1e9], [1e9] * 4], np.float32)
    hc, value, uncert = pdf_eval_values(evhit, hitcount, bincount, nearest, np.array([0, 1, 2], np.uint32), w, k)
    assert value[0] == 8 / 100 / w and uncert[0] == pytest.approx(value[0] / np.sqrt(8))
    assert value[1] == pytest.approx(2 / 50 / float(np.float32(0.4)) / 2.0)
    assert uncert[1] == pytest.approx(value[1] / np.sqrt(2))
    assert value[2] == 0 and value[3] == 0


def test_oracle_moments(O):
    rng = np.random.default_rng(5)
    nch = 200
    mom0 = np.zeros(nch, np.uint32)
    m = [np.zeros(nch, np.float32) for _ in range(4)]
    ts, qs = [], []
    for _ in range(6):
        t = synthetic_daq(rng, nch, 1)[0]
        q = rng.normal(1.0, 0.3, nch).astype(np.float32)
        ts.append(t); qs.append(q)
        O.accumulate_moments(False, t, q, (10.0, 28.0), (0.5, 1.5), mom0, *m)
    t, q = np.array(ts, np.float64), np.array(qs, np.float64)
    ok = (t >= 10.0) & (t <= 28.0) & (q >= 0.5) & (q <= 1.5)
    np.testing.assert_array_equal(mom0, ok.sum(axis=0))
    np.testing.assert_allclose(m[0], (t * ok).sum(axis=0), rtol=1e-5)
    np.testing.assert_allclose(m[1], (t * t * ok).sum(axis=0), rtol=1e-5)
    np.testing.assert_allclose(m[2], (q * ok).sum(axis=0), rtol=1e-5)
    np.testing.assert_allclose(m[3], (q * q * ok).sum(axis=0), rtol=1e-5)


def test_oracle_kernel_eval_is_a_normalised_density(O):
    """Summing the truncated-Gaussian kernel over many MC draws estimates the density: for MC
    times drawn from N(20, 3) and a fixed bandwidth the value at t0 approaches the N(20, sqrt(9+h^2))
    density renormalised to the window."""
    from math import erf, sqrt, pi, exp
    rng = np.random.default_rng(9)
    nch, nmc, h = 8, 4000, 0.5
    event_hit = np.ones(nch, np.uint32); event_hit[-1] = 0
    event_time = np.linspace(15.0, 25.0, nch).astype(np.float32)
    inv_bw = np.full(nch, 1.0 / h, np.float32)
    hitcount = np.zeros(nch, np.uint32)
    tp, qp = np.zeros(nch, np.float32), np.zeros(nch, np.float32)
    for _ in range(nmc):
        t = rng.normal(20.0, 3.0, nch).astype(np.float32)
        O.accumulate_kernel_eval(True, event_hit, event_time, np.zeros(nch, np.float32), t, np.zeros(nch, np.float32),
                                 (0.0, 40.0), (0.0, 1.0), inv_bw, np.zeros(nch, np.float32), hitcount, tp, qp)
    assert (hitcount == nmc).all() and tp[-1] == 0 and (qp == 0).all()
    # kernel term = exp(-arg^2/2)/h / (sqrt(pi/2) * [erf-erf]) -> 1/(h sqrt(2 pi)) exp(..) for a wide window
    s = sqrt(9.0 + h * h)
    expected = np.array([exp(-0.5 * ((x - 20.0) / s) ** 2) / (s * sqrt(2 * pi)) for x in event_time[:-1]])
    np.testing.assert_allclose(tp[:-1] / nmc, expected, rtol=0.12)


def test_compute_bandwidth_rule_matches_reference_expression():
    "compute_bandwidth (pdf.py:61-113) on fixed moments, no device involved"
    from chroma_b200.gpu.pdf import GPUKernelPDF
    k = GPUKernelPDF.__new__(GPUKernelPDF)
    k.time_only = True
    hitcount = np.array([100, 0, 4], np.uint32)
    tm1 = np.array([2000.0, 0.0, 80.0], np.float32)
    tm2 = np.array([40900.0, 0.0, 1600.0], np.float32)   # var 9, -, 0
    k.get_moments = lambda: (hitcount, tm1, tm2, np.zeros(3, np.float32), np.zeros(3, np.float32))
    k.compute_bandwidth(np.ones(3, bool), np.array([20.0, 1.0, 20.0]), np.zeros(3))
    # channel 0: n=100, rms 3 -> (4/3/100)^(-1/5) / min(1/3, exp(0)/sqrt(2pi)/3)
    dens = min(1 / 3.0, 1 / np.sqrt(2 * np.pi) / 3.0)
    bw = ((4.0 / 3) / 100) ** (-0.2) / dens
    assert k.inv_time_bandwidths[0] == pytest.approx(1 / bw, rel=1e-6)
    assert k.inv_time_bandwidths.dtype == np.float32
    assert (k.inv_charge_bandwidths == 0).all()
    # zero-variance channels: density is inf or nan -> bandwidth 0 or nan -> inverse stays 0
    assert k.inv_time_bandwidths[1] == 0 and k.inv_time_bandwidths[2] == 0
