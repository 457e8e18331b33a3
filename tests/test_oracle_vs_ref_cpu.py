"""Pins the oracle port (oracle/chroma_oracle.cpp) against the reference's OWN source run here:
chroma/cuda/propagate.cu, daq.cu and the headers they include, compiled for the host by
`make -C oracle ref_cpu` (oracle/ref_cpu_wrapper.cpp -> oracle/_ref/libchroma_ref_cpu.so) with the
engine's Philox streams behind curand_uniform.  The only arithmetic difference between the two is
that the port restates nvcc's FMA contraction in intersect_triangle / get_node (it is bit-exact
against the reference's sm_100a cubin, tests/test_reference_gpu.py) while the host build of the
reference is contraction-free IEEE: first-hit triangles and photon histories must be identical,
distances and positions equal to a few float ulp."""
import numpy as np
import pytest

import fixtures as F
from chroma_b200 import event
from oracle import ref_cpu

pytestmark = pytest.mark.skipif(not ref_cpu.available(),
                                reason='oracle/_ref/libchroma_ref_cpu.so not built (needs /root/reference at build time)')

DIST_RTOL = 2e-5      # FMA contraction vs separate multiply-add in the Moller-Trumbore test
TERMINAL = event.NO_HIT | event.BULK_ABSORB | event.SURFACE_DETECT | event.SURFACE_ABSORB | event.NAN_ABORT


@pytest.fixture(scope='module', params=['small_detector', 'scintillator_detector', 'water_sphere'])
def pair(request, oracle_mod):
    det = F.attach_bvh(getattr(F, request.param)(), oracle_mod.build_bvh)
    orc = oracle_mod.Oracle(det)
    return request.param, det, orc, ref_cpu.RefCpu(det, orc)


def test_first_hit(pair):
    name, det, orc, ref = pair
    for pos, d, last in [F.isotropic_rays(40000, seed=2) + (None,),
                         F.isotropic_rays(20000, seed=3, radius=900.0) + (None,)]:
        to, do = orc.intersect(pos, d, last)
        tr, dr = ref.intersect(pos, d, last)
        assert np.array_equal(to, tr)
        hit = to >= 0
        assert hit.mean() > 0.5
        assert np.abs(do[hit] - dr[hit]).max() <= DIST_RTOL * do[hit].max()
        assert (dr[~hit] == -1.0).all() and (do[~hit] == -1.0).all()
    # second bounce: the last hit triangle is excluded by both
    pos, d = F.isotropic_rays(20000, seed=4)
    t1, d1 = ref.intersect(pos, d)
    ok = t1 >= 0
    p2 = (pos + d * d1[:, None])[ok].astype(np.float32)
    d2 = (-d[ok]).astype(np.float32)
    to, _ = orc.intersect(p2, d2, t1[ok])
    tr, _ = ref.intersect(p2, d2, t1[ok])
    assert np.array_equal(to, tr) and (tr != t1[ok]).all()


@pytest.mark.parametrize('kw', [dict(max_steps=100), dict(max_steps=3), dict(max_steps=100, use_weights=True),
                                dict(max_steps=100, scatter_first=1)])
def test_propagate_histories(pair, kw):
    name, det, orc, ref = pair
    wl = (380, 700) if name == 'scintillator_detector' else (400, 800)
    ph = F.isotropic_photons(30000, seed=11, wavelength_range=wl, t0=5.0)
    po = orc.propagate(ph, seed=21, epoch=3, id_base=1000, **kw)
    pr = ref.propagate(ph, seed=21, epoch=3, id_base=1000, **kw)
    same = po.flags == pr.flags
    assert same.mean() >= 0.9995          # a trajectory may fork where a comparison sits on a rounding edge
    assert (po.last_hit_triangles[same] == pr.last_hit_triangles[same]).all()
    scale = np.abs(pr.pos).max()
    assert np.abs(po.pos[same] - pr.pos[same]).max() <= 2e-5 * scale
    assert np.abs(po.t[same] - pr.t[same]).max() <= 2e-5 * max(1.0, np.abs(pr.t).max())
    assert np.allclose(po.dir[same], pr.dir[same], atol=2e-5) and np.allclose(po.pol[same], pr.pol[same], atol=2e-4)
    assert np.array_equal(po.wavelengths[same], pr.wavelengths[same])
    assert np.allclose(po.weights[same], pr.weights[same], rtol=1e-4)   # products of exp(-d/L) over the path
    if kw['max_steps'] == 100:
        assert ((pr.flags & TERMINAL) != 0).mean() > 0.99
    # terminal photons are left alone by both
    again = ref.propagate(pr, seed=1, epoch=9, **kw)
    done = (pr.flags & TERMINAL) != 0
    assert np.array_equal(again.pos[done], pr.pos[done]) and np.array_equal(again.flags[done], pr.flags[done])


def test_run_daq(pair):
    name, det, orc, ref = pair
    if not hasattr(det, 'num_channels'):
        pytest.skip('no channels')
    ph = F.isotropic_photons(60000, seed=13, wavelength_range=(400, 700), t0=50.0)
    end = ref.propagate(ph, seed=2, max_steps=100)
    co, cr = orc.run_daq(end, seed=2, epoch=1), ref.run_daq(end, seed=2, epoch=1)
    assert cr.hit.sum() > 0
    assert np.array_equal(co.hit, cr.hit) and np.array_equal(co.flags, cr.flags)
    assert np.array_equal(co.t, cr.t) and np.array_equal(co.q, cr.q)
    half = orc.run_daq(end, seed=2, epoch=1, weight=0.5), ref.run_daq(end, seed=2, epoch=1, weight=0.5)
    assert np.array_equal(half[0].t, half[1].t) and half[1].q.sum() < cr.q.sum()
