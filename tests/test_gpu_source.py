"""GPU parity of the device photon source (chroma_photons_generate) against
the CPU oracle on the same Philox streams, and its use through GPUPhotons / Simulation."""
import numpy as np
import pytest

import fixtures as F
import stats_util as S
from chroma_b200 import event, gpu
from chroma_b200.sim import Simulation

pytestmark = pytest.mark.gpu

# Tolerance of the floating-point tier: every operation of the sampler is IEEE-rounded and ordered
# identically on both sides except sin/cos(2 pi u) (device sincospif, oracle double sin/cos rounded
# to float): <= 2 ulp of a value <= 1 on the cone vector, amplified by at most 1/|cross| in the
# normalised polarisation.
DIR_ATOL = 3e-7


def oracle_bomb(oracle_mod, src, n, id0):
    return oracle_mod.generate_photons(n, id0, src.seed, src.sequence, src.rotation(), src.cos_half_angle,
                                       src.wavelength_lo, src.wavelength_hi)


@pytest.mark.parametrize('kw', [dict(), dict(phi=0.3, direction=(0, 1, 1), wavelengths=500.0),
                                dict(phi=2.5, direction=(-1, 0.2, -0.4), wavelengths=(300.0, 650.0))])
def test_generate_vs_oracle(ctx, oracle_mod, kw):
    n = 300001
    src = event.PhotonSource(n, pos=(10.0, -20.0, 30.0), t0=7.5, weight=0.25, seed=17, sequence=4, **kw)
    got = gpu.GPUPhotons(src).get()
    d, p, w = oracle_bomb(oracle_mod, src, n, 0)
    assert np.abs(got.dir - d).max() <= DIR_ATOL
    # where the cross product is not tiny the polarisation agrees to the same few ulp
    assert np.quantile(np.abs(got.pol - p).max(axis=1), 0.999) <= 1e-5
    assert np.abs(got.pol - p).max() <= 1e-3
    assert (np.abs(got.dir - d).max(axis=1) == 0).mean() > 0.5     # most photons are bit-identical
    assert np.array_equal(got.wavelengths, w)
    assert (got.pos == np.float32([10.0, -20.0, 30.0])).all()
    assert (got.t == np.float32(7.5)).all() and (got.weights == np.float32(0.25)).all()
    assert (got.flags == 0).all() and (got.last_hit_triangles == -1).all() and (got.evidx == 0).all()
    assert np.abs(np.linalg.norm(got.dir, axis=1) - 1).max() < 5e-7
    assert np.abs(np.linalg.norm(got.pol, axis=1) - 1).max() < 5e-7


def test_generate_sharded_and_copies(ctx):
    "id_base = first photon of the shard reproduces the single-container photons; ncopies replicates"
    src = event.PhotonSource(50000, seed=3)
    whole = gpu.GPUPhotons(src).get()
    part = event.PhotonSource(20000, seed=3)
    shard = gpu.GPUPhotons(part, id_base=30000).get()
    assert np.array_equal(whole.dir[30000:], shard.dir) and np.array_equal(whole.pol[30000:], shard.pol)
    assert np.array_equal(whole.wavelengths[30000:], shard.wavelengths)
    rep = gpu.GPUPhotons(src, ncopies=3)
    assert len(rep) == 150000 and rep.true_nphotons == 50000
    out = rep.get()
    for k in range(3):
        assert np.array_equal(out.dir[k * 50000:(k + 1) * 50000], whole.dir)
    empty = gpu.GPUPhotons(event.PhotonSource(0))
    assert len(empty) == 0 and len(empty.get()) == 0


def test_window_load_mixes_host_and_device_photons(ctx):
    host = F.isotropic_photons(1000, seed=2, wavelength=450.0)
    host.evidx[:] = 1
    src = event.PhotonSource(3000, wavelengths=410.0, seed=1)
    src.evidx = 0
    src2 = event.PhotonSource(500, wavelengths=420.0, seed=1, sequence=9)
    src2.evidx = 2
    g = gpu.GPUPhotons.empty(4500)
    g.load(0, src)
    g.load(3000, host)
    g.load(4000, src2)
    out = g.get()
    assert (out.wavelengths[:3000] == 410).all() and (out.wavelengths[4000:] == 420).all()
    assert np.array_equal(out.dir[3000:4000], host.dir) and np.array_equal(out.pol[3000:4000], host.pol)
    assert np.array_equal(out.evidx, np.repeat([0, 1, 2], [3000, 1000, 500]).astype(np.uint32))
    assert (out.flags == 0).all() and (out.last_hit_triangles == -1).all() and (out.weights == 1).all()
    alone = gpu.GPUPhotons(src).get()
    assert np.array_equal(alone.dir, out.dir[:3000])


def test_simulation_from_sources_matches_host_bomb(ctx, oracle_mod):
    """Statistical tier: events started from device sources detect like events started from host
    photon bombs (detection efficiency within 0.5 % + 3 sigma, history-bit fractions within 4 sigma,
    hit-time KS p > 0.01)."""
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    sim = Simulation(det, seed=5)
    n, nev = 100000, 3
    srcs = [event.PhotonSource(n, wavelengths=(400.0, 800.0), t0=100.0, seed=40, sequence=k) for k in range(nev)]
    evs = list(sim.simulate(srcs, keep_photons_beg=True, keep_photons_end=True, run_daq=True, max_steps=100,
                            photons_per_batch=2 * n))
    assert [ev.id for ev in evs] == [0, 1, 2] and all(ev.nphotons == n for ev in evs)
    assert all(len(ev.photons_beg) == n and len(ev.photons_end) == n for ev in evs)
    assert not np.array_equal(evs[0].photons_beg.dir, evs[1].photons_beg.dir)
    assert (evs[1].photons_end.evidx == 1).all() and (evs[2].photons_end.evidx == 0).all()
    hosts = [F.isotropic_photons(n, seed=70 + k, wavelength_range=(400, 800), t0=100.0) for k in range(nev)]
    ref = list(sim.simulate(hosts, keep_photons_end=True, run_daq=True, max_steps=100, photons_per_batch=2 * n))
    fa = np.concatenate([ev.photons_end.flags for ev in evs])
    fb = np.concatenate([ev.photons_end.flags for ev in ref])
    assert all(ok for *_, ok in S.fractions_agree(fa, fb, nsigma=4.0))
    ea = ((fa & event.SURFACE_DETECT) != 0).mean()
    eb = ((fb & event.SURFACE_DETECT) != 0).mean()
    assert abs(ea - eb) < 0.005 * eb + 3 * S.binomial_sigma(eb, len(fa)) * np.sqrt(2)
    ta = np.concatenate([ev.channels.t[ev.channels.hit] for ev in evs])
    tb = np.concatenate([ev.channels.t[ev.channels.hit] for ev in ref])
    assert S.ks_p(ta, tb) > 0.01
    # a single source goes straight in, like a single Photons
    one = next(sim.simulate(event.PhotonSource(1000, seed=1), keep_photons_end=True, max_steps=10))
    assert len(one.photons_end) == 1000
