"""Photon source, CPU side: the host description (event.PhotonSource) against
the reference-pinned samplers, and the oracle restatement of the device sampler against the
distributions of chroma/sample.py + chroma/benchmark.py:75-83."""
import numpy as np
import pytest
from scipy import stats

from chroma_b200 import event, sample
from chroma_b200.transform import normalize


def cone(theta, u):
    c = np.sqrt(1 - u ** 2)
    return np.stack([c * np.cos(theta), c * np.sin(theta), u], axis=1)


@pytest.mark.parametrize('direction', [(0, 0, 1), (0, 1, 1), (1, 0, 0), (-0.3, 0.2, -0.9), (2, -1, 0.5)])
def test_rotation_is_the_flashlight_rotation(direction):
    """sample.flashlight (pinned to the reference by tests/golden, test_host_golden.py) draws
    (theta, u) and rotates the cone vectors; PhotonSource.rotation() must be that rotation."""
    phi = 0.4
    np.random.seed(21)
    want = sample.flashlight(phi=phi, direction=direction, size=200)
    np.random.seed(21)
    theta = np.random.uniform(0.0, 2 * np.pi, 200)
    u = np.random.uniform(np.cos(phi), 1, 200)
    src = event.PhotonSource(200, direction=direction, phi=phi)
    got = cone(theta, u) @ src.rotation().T
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-12)
    np.testing.assert_allclose(src.rotation() @ np.array([0, 0, 1.0]), normalize(np.asarray(direction, float)),
                               atol=1e-12)


def test_source_description():
    s = event.PhotonSource(10, wavelengths=430.0)
    assert len(s) == 10 and s.wavelength_lo == s.wavelength_hi == 430.0
    assert s.cos_half_angle == -1.0 and np.array_equal(s.rotation(), np.eye(3))
    s = event.PhotonSource(10, direction=(0, 0, -1), phi=0.1)
    np.testing.assert_allclose(s.rotation() @ np.array([0, 0, 1.0]), [0, 0, -1], atol=1e-15)
    assert abs(s.cos_half_angle - np.cos(0.1)) < 1e-15


def test_oracle_bomb_distribution(oracle_mod):
    "isotropic bomb: unit vectors, transverse polarisation, the reference sampler's distributions"
    n = 200000
    d, p, w = oracle_mod.generate_photons(n, 0, seed=5, sequence=0, rot=np.eye(3), cos_half=-1.0,
                                          wl_lo=400.0, wl_hi=800.0)
    assert np.abs(np.linalg.norm(d, axis=1) - 1).max() < 5e-7
    assert np.abs(np.linalg.norm(p, axis=1) - 1).max() < 5e-7
    assert np.abs((d * p).sum(axis=1)).max() < 1e-5   # cross of nearly parallel vectors, renormalised
    np.random.seed(3)
    ref = sample.uniform_sphere(n)
    ref_pol = normalize(np.cross(sample.uniform_sphere(n), ref))
    for k in range(3):
        assert stats.ks_2samp(d[:, k], ref[:, k]).pvalue > 0.01
        assert stats.ks_2samp(p[:, k], ref_pol[:, k]).pvalue > 0.01
    assert stats.kstest(np.arctan2(d[:, 1], d[:, 0]), 'uniform', args=(-np.pi, 2 * np.pi)).pvalue > 0.01
    # polarisation angle about the direction is uniform: project on a fixed transverse frame
    e1 = normalize(np.cross(d, np.array([0.3, -0.5, 0.8])))
    e2 = np.cross(d, e1)
    ang = np.arctan2((p * e2).sum(axis=1), (p * e1).sum(axis=1))
    assert stats.kstest(ang, 'uniform', args=(-np.pi, 2 * np.pi)).pvalue > 0.01
    assert w.min() > 400.0 and w.max() <= 800.0
    assert stats.kstest(w, 'uniform', args=(400, 400)).pvalue > 0.01


def test_oracle_flashlight_cone(oracle_mod):
    phi, direction = 0.3, (0, 1, 1)
    src = event.PhotonSource(100000, direction=direction, phi=phi, wavelengths=500.0)
    d, p, w = oracle_mod.generate_photons(len(src), 0, 9, 1, src.rotation(), src.cos_half_angle, 500.0, 500.0)
    axis = normalize(np.asarray(direction, float))
    cosang = d @ axis
    assert cosang.min() >= np.cos(phi) - 1e-6
    assert stats.kstest(cosang, 'uniform', args=(np.cos(phi), 1 - np.cos(phi))).pvalue > 0.01
    np.random.seed(8)
    ref = sample.flashlight(phi=phi, direction=direction, size=len(src))
    for k in range(3):
        assert stats.ks_2samp(d[:, k], ref[:, k]).pvalue > 0.01
    assert (w == np.float32(500.0)).all()
    assert np.abs((d * p).sum(axis=1)).max() < 1e-5   # cross of nearly parallel vectors, renormalised


def test_oracle_source_is_keyed_by_photon_index(oracle_mod):
    "splitting a source over shards (id0 = first photon of the shard) gives the same photons"
    args = dict(seed=11, sequence=2, rot=np.eye(3), cos_half=-1.0, wl_lo=400.0, wl_hi=800.0)
    d, p, w = oracle_mod.generate_photons(1000, 0, **args)
    d2, p2, w2 = oracle_mod.generate_photons(400, 600, **args)
    assert np.array_equal(d[600:], d2) and np.array_equal(p[600:], p2) and np.array_equal(w[600:], w2)
    d3, _, _ = oracle_mod.generate_photons(1000, 0, **dict(args, sequence=3))
    assert not np.array_equal(d, d3)
