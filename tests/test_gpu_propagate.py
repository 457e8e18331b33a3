"""GPU parity: the propagate step loop (engine through the C-ABI) vs the CPU oracle and vs
analytic expectations.  Engine and oracle consume the same Philox stream per photon in the
reference's draw order, so most photons agree exactly; the remainder differ only through
fast-math rounding at decision boundaries, which the statistical checks bound
(tolerances from BASELINE.json: detection efficiency within 0.5 %, chi-square / KS p > 0.01)."""
import numpy as np
import pytest

import fixtures as F
import stats_util as S
from chroma_b200 import event, gpu

pytestmark = pytest.mark.gpu

TERMINAL = event.NO_HIT | event.BULK_ABSORB | event.SURFACE_DETECT | event.SURFACE_ABSORB | event.NAN_ABORT


def run_both(oracle_mod, geometry, photons, seed, max_steps, detector=False, **kw):
    F.attach_bvh(geometry, oracle_mod.build_bvh)
    orc = oracle_mod.Oracle(geometry)
    gg = gpu.GPUDetector(geometry) if detector else gpu.GPUGeometry(geometry)
    rng = gpu.get_rng_states(64 * 1024, seed=seed)
    gp = gpu.GPUPhotons(photons)
    gp.propagate(gg, rng, max_steps=max_steps, **kw)
    out_g = gp.get()
    out_o = orc.propagate(photons, seed=seed, epoch=0, max_steps=max_steps, **kw)
    return out_g, out_o, gg, gp, rng, orc


def test_c1_water_sphere_analytic(ctx, oracle_mod):
    """config C1: 1e5 isotropic 400 nm photons at the centre of a 20 m water sphere, one step.
    P(BULK_ABSORB)=0.1079, P(RAYLEIGH)=0.2375, P(reach wall)=0.6546 from n=1.34419,
    L_abs=151139 mm, L_scat=68646 mm; unscattered arrival at R*n/c."""
    n = 100000
    ph = F.isotropic_photons(n, seed=0, wavelength=400.0)
    out_g, out_o, *_ = run_both(oracle_mod, F.water_sphere(), ph, seed=1, max_steps=1)
    for bit, p in ((event.BULK_ABSORB, 0.1079), (event.RAYLEIGH_SCATTER, 0.2375),
                   (event.SURFACE_ABSORB, 0.6546)):
        f = ((out_g.flags & bit) != 0).mean()
        assert abs(f - p) < 4 * S.binomial_sigma(p, n) + 2e-3, (bit, f, p)  # 2e-3: polygonal sphere
    assert ((out_g.flags & event.NO_HIT) != 0).sum() == 0
    wall = (out_g.flags & event.SURFACE_ABSORB) != 0
    assert abs(out_g.t[wall].mean() - 89.67) < 0.15
    # same stream, same draw order: nearly all photons end identically
    same = out_g.flags == out_o.flags
    assert same.mean() > 0.999
    assert np.allclose(out_g.pos[same], out_o.pos[same], rtol=0, atol=0.5)
    assert np.allclose(out_g.t[same], out_o.t[same], rtol=0, atol=1e-2)
    # run to termination: everything ends absorbed in bulk or on the wall
    gp = gpu.GPUPhotons(ph)
    gg = gpu.GPUGeometry(F.attach_bvh(F.water_sphere(), oracle_mod.build_bvh))
    gp.propagate(gg, gpu.get_rng_states(1, seed=7), max_steps=100)
    fl = gp.flags.get()
    assert ((fl & TERMINAL) != 0).mean() > 0.9999
    frac_abs = ((fl & event.BULK_ABSORB) != 0).mean()
    # with scattering the mean path grows; absorbed fraction must exceed the one-step value
    assert 0.12 < frac_abs < 0.30


def test_rayleigh_angular_distribution(ctx, oracle_mod):
    "test/test_rayleigh.py: 100 mm water box, photons along +z pol +x; (1+cos^2)sin distribution"
    n = 200000
    ph = event.Photons(np.zeros((n, 3)), np.tile([0, 0, 1.0], (n, 1)), np.tile([1.0, 0, 0], (n, 1)),
                       np.full(n, 400.0))
    geo = F.attach_bvh(F.water_box(100000.0), oracle_mod.build_bvh)  # big box: many scatters in one step
    gp = gpu.GPUPhotons(ph)
    gp.propagate(gpu.GPUGeometry(geo), gpu.get_rng_states(1, seed=3), max_steps=1)
    out = gp.get()
    sc = (out.flags & event.RAYLEIGH_SCATTER) != 0
    assert sc.sum() > 20000
    # scattering angle relative to the incoming polarisation axis (+x): p(c) ~ 1 - c^2
    c = out.dir[sc, 0]
    hist, edges = np.histogram(c, bins=50, range=(-1, 1))
    mid = 0.5 * (edges[1:] + edges[:-1])
    expect = (1 - mid ** 2)
    expect *= hist.sum() / expect.sum()
    from scipy import stats
    chi2 = ((hist - expect) ** 2 / np.maximum(expect, 1)).sum()
    assert stats.chi2.sf(chi2, 49) > 1e-3
    # polarisation stays perpendicular to the new direction and normalised
    assert np.abs(np.einsum('ij,ij->i', out.dir[sc], out.pol[sc])).max() < 1e-3
    assert np.abs(np.linalg.norm(out.pol[sc], axis=1) - 1).max() < 1e-4


def test_vacuum_box_no_abort(ctx, oracle_mod):
    "test/test_propagation.py testAbort: axis-aligned photons at normal incidence never NaN"
    n = 10000
    geo = F.attach_bvh(F.vacuum_box(100.0), oracle_mod.build_bvh)
    gg = gpu.GPUGeometry(geo)
    rng = gpu.get_rng_states(1, seed=1)
    for d in ([1, 0, 0], [0, -1, 0], [0, 0, 1]):
        ph = event.Photons(np.zeros((n, 3)), np.tile(np.array(d, float), (n, 1)),
                           np.tile([0, 1, 0] if d[1] == 0 else [1, 0, 0], (n, 1)), np.full(n, 400.0))
        gp = gpu.GPUPhotons(ph)
        gp.propagate(gg, rng, max_steps=1)
        out = gp.get()
        assert not np.isnan(out.pos).any() and not np.isnan(out.dir).any() and not np.isnan(out.pol).any()
        gp.propagate(gg, rng, max_steps=10)
        assert (gp.flags.get() & event.NAN_ABORT).sum() == 0


def test_c3_detector_bomb_vs_oracle(ctx, oracle_mod):
    "config C3 shape on a small demo detector: photon bomb, full surface physics"
    n = 200000
    ph = F.isotropic_photons(n, seed=11, wavelength_range=(400, 800), t0=100.0)
    det = F.small_detector()
    out_g, out_o, gg, gp, rng, orc = run_both(oracle_mod, det, ph, seed=5, max_steps=100, detector=True)
    same = out_g.flags == out_o.flags
    assert same.mean() > 0.98, same.mean()
    assert (out_g.last_hit_triangles[same] == out_o.last_hit_triangles[same]).mean() > 0.999
    eff_g = ((out_g.flags & event.SURFACE_DETECT) != 0).mean()
    eff_o = ((out_o.flags & event.SURFACE_DETECT) != 0).mean()
    assert abs(eff_g - eff_o) <= 0.005 * eff_o
    for b, fa, fb, ok in S.fractions_agree(out_g.flags, out_o.flags):
        assert ok, (b, fa, fb)
    dg = (out_g.flags & event.SURFACE_DETECT) != 0
    do = (out_o.flags & event.SURFACE_DETECT) != 0
    assert S.ks_p(out_g.t[dg], out_o.t[do]) > 0.01
    # per-channel hit counts from get_flat_hits vs the oracle's photon list
    hits = gp.get_flat_hits(gg)
    assert len(hits) == dg.sum()  # every detection in this geometry is on a PMT
    nch = det.num_channels()
    solid = det.solid_id[out_o.last_hit_triangles[do]]
    ch_o = np.bincount(det.solid_id_to_channel_index[solid], minlength=nch)
    ch_g = np.bincount(hits.channel.astype(np.int64), minlength=nch)
    assert S.chi2_counts_p(ch_g, ch_o) > 0.01
    # a statistically independent engine run (other seed) agrees with the oracle as well
    gp2 = gpu.GPUPhotons(ph)
    gp2.propagate(gg, gpu.get_rng_states(1, seed=99), max_steps=100)
    out2 = gp2.get()
    for b, fa, fb, ok in S.fractions_agree(out2.flags, out_o.flags):
        assert ok, (b, fa, fb)
    eff2 = ((out2.flags & event.SURFACE_DETECT) != 0).mean()
    assert abs(eff2 - eff_o) < 4 * np.sqrt(2) * S.binomial_sigma(eff_o, n)


def test_c4_scintillator_wls_dichroic(ctx, oracle_mod):
    "config C4 (synthetic): bulk reemission, WLS, dichroic and thin-film surfaces"
    n = 200000
    ph = F.isotropic_photons(n, seed=3, wavelength=380.0, pos=(100, 50, -30))
    det = F.scintillator_detector()
    out_g, out_o, gg, gp, rng, orc = run_both(oracle_mod, det, ph, seed=5, max_steps=100, detector=True)
    assert (out_g.flags == out_o.flags).mean() > 0.95
    for b, fa, fb, ok in S.fractions_agree(out_g.flags, out_o.flags):
        assert ok, (b, fa, fb)
    for bit in (event.BULK_REEMIT, event.SURFACE_REEMIT, event.SURFACE_TRANSMIT, event.SURFACE_DETECT):
        assert ((out_g.flags & bit) != 0).mean() > 0.05  # every model is exercised
    dg = (out_g.flags & event.SURFACE_DETECT) != 0
    do = (out_o.flags & event.SURFACE_DETECT) != 0
    assert S.ks_p(out_g.t[dg], out_o.t[do]) > 0.01
    hg, _ = np.histogram(out_g.wavelengths[dg], bins=40, range=(300, 700))
    ho, _ = np.histogram(out_o.wavelengths[do], bins=40, range=(300, 700))
    assert S.chi2_counts_p(hg, ho) > 0.01


def test_weights_and_scatter_first(ctx, oracle_mod):
    "use_weights / scatter_first modes (photon.h:201-232, :700-718)"
    n = 50000
    ph = F.isotropic_photons(n, seed=21, wavelength=450.0)
    for kw in (dict(use_weights=True), dict(scatter_first=1), dict(scatter_first=-1)):
        out_g, out_o, *_ = run_both(oracle_mod, F.small_detector(), ph, seed=2, max_steps=20,
                                    detector=True, **kw)
        assert (out_g.flags == out_o.flags).mean() > 0.97, kw
        assert abs(out_g.weights.mean() - out_o.weights.mean()) < 5e-3 * max(out_o.weights.mean(), 1e-3), kw
        if kw.get('scatter_first') == 1:  # forced unless absorption comes first
            assert ((out_g.flags & event.RAYLEIGH_SCATTER) != 0).mean() > 0.97
        if kw.get('use_weights'):
            assert ((out_g.flags & event.BULK_ABSORB) != 0).sum() == 0


def test_repeated_calls_and_terminal_skipping(ctx, oracle_mod):
    "propagate may be called repeatedly; terminal photons are left untouched (propagate.cu:258)"
    n = 50000
    ph = F.isotropic_photons(n, seed=4, wavelength_range=(400, 800))
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    gg = gpu.GPUDetector(det)
    rng = gpu.get_rng_states(1, seed=8)
    gp = gpu.GPUPhotons(ph)
    gp.propagate(gg, rng, max_steps=1)
    first = gp.get()
    done = (first.flags & TERMINAL) != 0
    assert 0 < done.sum() < n
    for _ in range(60):
        gp.propagate(gg, rng, max_steps=1)
    final = gp.get()
    assert np.array_equal(final.pos[done], first.pos[done])
    assert np.array_equal(final.flags[done], first.flags[done])
    assert ((final.flags & TERMINAL) != 0).mean() > 0.999
    # single-stepping is statistically the same as one call
    gp2 = gpu.GPUPhotons(ph)
    gp2.propagate(gg, rng, max_steps=61)
    for b, fa, fb, ok in S.fractions_agree(final.flags, gp2.flags.get()):
        assert ok, (b, fa, fb)
    # empty photon list
    empty = gpu.GPUPhotons(event.Photons())
    empty.propagate(gg, rng)
    assert len(empty.get()) == 0
