"""GPU tests of the two-level ("instanced") geometry, SURVEY.md 8(f) rank 3 (chroma_b200/csrc/inst.cuh,
engine_inst.cu): placed solids stored once, detector assembled on the device from
Geometry.placements() instead of the flattened mesh of chroma/geometry.py:360-367.

Tier: statistical (the triangle test runs in the template frame, so distances differ from the flat
mesh in the last bits).  What is asserted:
  * first hits: same triangle id as the flat engine for all but grazing rays, distances equal to 1e-5;
  * photon histories: the Philox streams are keyed per photon, so the flat and the two-level engine give
    the SAME history for every photon whose trajectory does not fork on a rounding edge (> 99 %);
  * against the reference's own propagate.cubin on the flattened mesh: history-bit fractions within
    4 sigma, detection efficiency within 0.5 %, per-channel chi-square and hit-time KS p > 0.01;
  * run_daq / get_flat_hits see the same solids and channels as on the flat geometry;
  * the lazily assembled flat arrays (vertices, triangles, material codes, solid ids) equal the
    host-flattened mesh."""
import numpy as np
import pytest

import fixtures as F
import stats_util as S
from chroma_b200 import event, gpu
from chroma_b200.detector import Detector
from chroma_b200.geometry import Solid
from chroma_b200 import make
from chroma_b200.demo.optics import water, glass, black_surface, r7081hqe_photocathode
from chroma_b200.transform import make_rotation_matrix

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def pair(ctx, oracle_mod):
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)      # add_pmt path: one Solid, many placements
    flat = gpu.GPUDetector(det)
    inst = gpu.GPUDetector(det, instanced=True)
    return det, flat, inst


def test_tree_info(pair):
    det, flat, inst = pair
    tf, ti = flat.tree(), inst.tree()
    assert tf['kind'] == 'wide tree' and ti['kind'] == 'two-level tree'
    assert ti['placements'] == len(det.solids) and ti['templates'] == 2
    nt_templates = len(det.solids[0].mesh.triangles) + len(det.solids[1].mesh.triangles)
    assert ti['triangle_slots'] == nt_templates < tf['triangle_slots'] == len(det.mesh.triangles)
    assert ti['bytes'] * 3 < tf['bytes']            # 32 placements only; 8.9 MB vs 11.9 GB at 29,023


def test_first_hits_agree_with_flat(pair):
    det, flat, inst = pair
    for pos, d in (F.isotropic_rays(400000, seed=2), F.isotropic_rays(300000, seed=3, radius=1500.0)):
        tf, df = gpu.intersect(flat, pos, d)
        ti, di = gpu.intersect(inst, pos, d)
        same = tf == ti
        assert same.mean() > 0.9995, same.mean()
        hit = same & (tf >= 0)
        assert hit.mean() > 0.99
        # float rounding of the template-frame ray, amplified 1/cos(incidence) on grazing hits
        err = np.abs(df[hit] - di[hit]) / df[hit]
        assert np.quantile(err, 0.999) <= 1e-5 and err.max() <= 1e-3, (np.quantile(err, 0.999), err.max())
        assert ((tf < 0) == (ti < 0)).mean() > 0.99999
        # where the triangle differs, the two hits are (nearly) at the same distance: a shared edge
        bad = ~same & (tf >= 0) & (ti >= 0)
        if bad.any():
            assert np.abs(df[bad] - di[bad]).max() <= 1e-3
    # second bounce: the triangle just hit is excluded (mesh.h:82), given by its flat-mesh id
    pos, d = F.isotropic_rays(200000, seed=4)
    t1, d1 = gpu.intersect(inst, pos, d)
    ok = t1 >= 0
    p2 = (pos + d * d1[:, None])[ok].astype(np.float32)
    _, dir2 = F.isotropic_rays(int(ok.sum()), seed=5)
    tf, df = gpu.intersect(flat, p2, dir2, t1[ok])
    ti, di = gpu.intersect(inst, p2, dir2, t1[ok])
    assert (ti != t1[ok]).all()
    assert (tf == ti).mean() > 0.999


def test_histories_equal_flat_engine(pair):
    det, flat, inst = pair
    n = 400000
    ph = F.isotropic_photons(n, seed=11, wavelength_range=(400, 800), t0=100.0)
    outs = []
    for gg in (flat, inst):
        gp = gpu.GPUPhotons(ph)
        gp.propagate(gg, gpu.get_rng_states(1, seed=5), max_steps=100)
        outs.append(gp.get())
    of, oi = outs
    same = of.flags == oi.flags
    assert same.mean() > 0.99, same.mean()
    assert (of.last_hit_triangles[same] == oi.last_hit_triangles[same]).mean() > 0.999
    assert np.abs(of.pos[same] - oi.pos[same]).max(axis=1).mean() < 0.1     # mm, after up to 100 steps
    assert ((oi.flags & event.NAN_ABORT) != 0).sum() == 0
    # single-stepping through repeated calls (last_hit_triangles are mapped back to placements)
    gp = gpu.GPUPhotons(ph)
    rng = gpu.get_rng_states(1, seed=5)
    for _ in range(4):
        gp.propagate(inst, rng, max_steps=1)
    o4 = gp.get()
    gp = gpu.GPUPhotons(ph)
    rng = gpu.get_rng_states(1, seed=5)
    for _ in range(4):
        gp.propagate(flat, rng, max_steps=1)
    f4 = gp.get()
    assert (o4.flags == f4.flags).mean() > 0.99


@S.two_stage
def test_statistics_vs_reference_cubin(pair, ctx):
    from oracle import ref_gpu
    if not ref_gpu.available():
        pytest.fail('oracle/_ref/*.cubin missing')
    det, flat, inst = pair
    n = 400000
    ph = F.isotropic_photons(n, seed=12, wavelength_range=(400, 800), t0=100.0)
    rc = ref_gpu.RefContext(0)
    try:
        rg = ref_gpu.RefGeometry(rc, det)
        rp = ref_gpu.RefPhotons(rc, ph)
        rp.propagate(rg, ref_gpu.RefRng(rc, 64 * 1024, seed=7), max_steps=10)
        out_r = rp.get()
    finally:
        rc.close()
    gp = gpu.GPUPhotons(ph)
    rng = gpu.get_rng_states(1, seed=7)
    gp.propagate(inst, rng, max_steps=10)
    out_g = gp.get()
    for bit, fr, fg, ok in S.fractions_agree(out_r.flags, out_g.flags, nsigma=4.0):
        assert ok, 'history bit %d: reference %.5f two-level engine %.5f' % (bit, fr, fg)
    det_r = ((out_r.flags & event.SURFACE_DETECT) != 0).mean()
    det_g = ((out_g.flags & event.SURFACE_DETECT) != 0).mean()
    assert abs(det_r - det_g) <= 0.005 * det_r + 3 * np.sqrt(2) * S.binomial_sigma(det_r, n)
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
    # DAQ and hit compaction resolve triangle -> solid -> channel through the placement table
    daq = gpu.GPUDaq(inst)
    daq.begin_acquire(); daq.acquire(gp, rng)
    ch = daq.end_acquire().get()
    want = np.zeros(nch, bool)
    want[np.unique(ch_g)] = True
    assert np.array_equal(ch.hit, want)
    hits_g = gp.get_flat_hits(inst)
    assert len(hits_g) == len(ch_g) and np.array_equal(np.sort(hits_g.channel), np.sort(ch_g))


def test_flat_arrays_assembled_on_device(pair):
    det, flat, inst = pair
    tri = inst.triangles.get().view(np.uint32).reshape(-1, 3)
    ver = inst.vertices.get().view(np.float32).reshape(-1, 3)
    assert len(tri) == len(det.mesh.triangles)
    got = ver[tri]                                            # (ntri, 3, 3) corner coordinates
    want = np.asarray(det.mesh.vertices, np.float32)[np.asarray(det.mesh.triangles)]
    assert np.abs(got - want).max() < 2e-3                    # float32 placement vs float64 rounded once
    codes = (((det.inner_material_index & 0xff) << 24) | ((det.outer_material_index & 0xff) << 16) |
             ((det.surface_index & 0xff) << 8)).astype(np.uint32)
    assert np.array_equal(inst.material_codes.get(), codes)
    assert np.array_equal(inst.solid_id_map.get(), np.asarray(det.solid_id, np.uint32))
    assert np.array_equal(inst.colors.get(), np.asarray(det.colors, np.uint32))


def test_mixed_templates_rotations_and_errors(ctx, oracle_mod):
    "several templates, rotated and displaced placements, a solid used once; a non-rotation is refused"
    def build():
        det = Detector(water)
        det.add_solid(Solid(make.sphere(3000.0, nsteps=32), water, water, surface=black_surface))
        cube = Solid(make.cube(200.0), glass, water, surface=r7081hqe_photocathode)
        ball = Solid(make.sphere(120.0, nsteps=12), glass, water, surface=r7081hqe_photocathode)
        rs = np.random.RandomState(2)
        for k in range(60):
            axis = rs.normal(size=3)
            pos = rs.normal(size=3)
            pos *= rs.uniform(600, 2200) / np.linalg.norm(pos)
            det.add_pmt(cube if k % 2 else ball, make_rotation_matrix(rs.uniform(0, 3), axis / np.linalg.norm(axis)), pos)
        det.set_time_dist_gaussian(1.5, -7.5, 7.5)
        det.set_charge_dist_gaussian(1.0, 0.1, 0.0, 1.5)
        return det
    det = build()
    inst = gpu.GPUDetector(det, instanced=True)               # before any flatten(): placements only
    assert not hasattr(det, 'mesh') and inst.tree()['templates'] == 3
    F.attach_bvh(det, oracle_mod.build_bvh)
    flat = gpu.GPUDetector(det)
    pos, d = F.isotropic_rays(300000, seed=8, radius=500.0)
    tf, df = gpu.intersect(flat, pos, d)
    ti, di = gpu.intersect(inst, pos, d)
    assert (tf == ti).mean() > 0.9995 and (tf >= 0).all()
    hit = tf == ti
    assert np.quantile(np.abs(df[hit] - di[hit]) / df[hit], 0.999) <= 1e-5
    ph = F.isotropic_photons(200000, seed=3, wavelength_range=(400, 800))
    outs = []
    for gg in (flat, inst):
        gp = gpu.GPUPhotons(ph)
        gp.propagate(gg, gpu.get_rng_states(1, seed=2), max_steps=20)
        outs.append(gp.get().flags)
    assert (outs[0] == outs[1]).mean() > 0.99
    bad = build()
    bad.solid_rotations[5] = (bad.solid_rotations[5] * 1.5).astype(np.float32)
    from chroma_b200._lib import ChromaError
    with pytest.raises(ChromaError, match='not a pure rotation'):
        gpu.GPUDetector(bad, instanced=True)
