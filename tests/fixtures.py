"""Geometry and photon fixtures shared by the tests."""
import os

import numpy as np

from chroma_b200 import event, make, sample
from chroma_b200.detector import Detector
from chroma_b200.geometry import Geometry, Solid, Material, Surface, DichroicProps, vacuum, \
    standard_wavelengths
from chroma_b200.demo.optics import water, glass, black_surface, shiny_surface, \
    r7081hqe_photocathode


def attach_bvh(geometry, builder):
    geometry.flatten()
    geometry.bvh = builder(geometry.mesh)
    return geometry


def cube_geometry(size=1000.0):
    "test/test_ray_intersection.py setup: a 1000 mm cube"
    g = Geometry(vacuum)
    g.add_solid(Solid(make.cube(size), vacuum, vacuum))
    return g


def vacuum_box(size=100.0):
    "test/test_propagation.py:21-24"
    g = Geometry(vacuum)
    g.add_solid(Solid(make.box(size, size, size), vacuum, vacuum))
    return g


def water_box(size=100.0):
    "test/test_rayleigh.py:15-18"
    g = Geometry(water)
    g.add_solid(Solid(make.box(size, size, size), water, water))
    return g


def water_sphere(radius=20000.0, nsteps=64):
    "config C1: water sphere with a black wall"
    g = Geometry(water)
    g.add_solid(Solid(make.sphere(radius, nsteps=nsteps), water, water, surface=black_surface))
    return g


def small_detector(spiral_step=900.0):
    "a tiny()-style demo detector (tiny() itself is 53 PMTs, 387k triangles)"
    from chroma_b200 import demo
    return demo.detector(2000.0, 2500.0, spiral_step, fast=False)


def gaussian_cdf(grid, mean, sigma):
    pdf = np.exp(-0.5 * ((grid - mean) / sigma) ** 2)
    cdf = np.cumsum(pdf)
    cdf -= cdf[0]
    return cdf / cdf[-1]


def scintillator_detector():
    """config C4 (synthetic; no such geometry ships with the reference): a scintillator ball
    with two reemission components inside water, a spherical dichroic shell, a WLS shell, a
    thin-film (complex model) photocathode shell with channels, all inside a black sphere."""
    w = standard_wavelengths.astype(np.float64)
    times = np.arange(0, 1000, 0.05)
    scint = Material('scintillator')
    scint.set('refractive_index', 1.5)
    scint.set('absorption_length', 800.0)
    scint.set('scattering_length', 5000.0)
    for frac, mean, tau in ((0.7, 430.0, 5.0), (0.3, 480.0, 20.0)):
        scint.comp_absorption_length.append(np.column_stack([w, np.full(len(w), 800.0 / frac)]).astype(np.float32))
        scint.comp_reemission_prob.append(np.column_stack([w, np.where(w < 450, 0.8, 0.1)]).astype(np.float32))
        scint.comp_reemission_wvl_cdf.append(np.column_stack([w, gaussian_cdf(w, mean, 30.0)]).astype(np.float32))
        tc = 1 - np.exp(-times / tau)
        scint.comp_reemission_time_cdf.append(np.column_stack([times, tc / tc[-1]]))

    wls = Surface('wls', model=2)
    wls.set('absorb', np.where(w < 420, 0.6, 0.05))
    wls.set('reemit', 0.9)
    wls.set('reflect_specular', 0.05)
    wls.set('reflect_diffuse', 0.05)
    wls.set('reemission_cdf', gaussian_cdf(w, 500.0, 20.0))

    angles = np.linspace(0, np.pi / 2, 5)
    refl = [np.column_stack([w, np.where(w > 450 + 20 * i, 0.85, 0.05)]) for i in range(5)]
    tran = [np.column_stack([w, np.where(w > 450 + 20 * i, 0.10, 0.90)]) for i in range(5)]
    dich = Surface('dichroic', model=3)
    dich.dichroic_props = DichroicProps(angles, refl, tran)

    cathode = Surface('thin_film_cathode', model=1)
    cathode.set('detect', 0.3)
    cathode.set('eta', 2.0)
    cathode.set('k', 1.5)
    cathode.set('reflect_diffuse', 0.1)
    cathode.thickness = 25e-6  # mm
    cathode.transmissive = 1

    det = Detector(water)
    det.add_solid(Solid(make.sphere(1000.0, nsteps=24), scint, water))
    det.add_solid(Solid(make.sphere(1300.0, nsteps=24), water, water, surface=dich))
    det.add_solid(Solid(make.sphere(1600.0, nsteps=24), water, water, surface=wls))
    # eight cathode patches (octants of a sphere) as separate channels
    ball = make.sphere(1900.0, nsteps=24)
    centers = ball.get_triangle_centers()
    octant = (centers[:, 0] > 0).astype(int) + 2 * (centers[:, 1] > 0) + 4 * (centers[:, 2] > 0)
    from chroma_b200.geometry import Mesh
    for k in range(8):
        tri = ball.triangles[octant == k]
        used, inv = np.unique(tri, return_inverse=True)
        patch = Mesh(ball.vertices[used], inv.reshape(-1, 3), remove_null_triangles=False)
        det.add_pmt(Solid(patch, glass, water, surface=cathode))
    det.add_solid(Solid(make.sphere(2300.0, nsteps=24), water, water, surface=black_surface))
    det.set_time_dist_gaussian(1.5, -7.5, 7.5)
    det.set_charge_dist_gaussian(1.0, 0.1, 0.0, 1.5)
    return det


def isotropic_photons(n, seed=0, wavelength=400.0, pos=(0, 0, 0), t0=0.0, wavelength_range=None):
    "photon bomb: benchmark.py:77-83 shape (isotropic, random perpendicular polarisation)"
    rs = np.random.get_state()
    # the repetition of a statistical test (stats_util.two_stage) draws a fresh sample
    np.random.seed(seed + 100003 * int(os.environ.get('CHROMA_TEST_ATTEMPT', '0')))
    dirs = sample.uniform_sphere(n)
    pol = np.cross(sample.uniform_sphere(n), dirs)
    pol /= np.linalg.norm(pol, axis=1)[:, None]
    if wavelength_range is not None:
        wl = np.random.uniform(wavelength_range[0], wavelength_range[1], n)
    else:
        wl = np.full(n, wavelength)
    np.random.set_state(rs)
    position = np.tile(np.asarray(pos, dtype=np.float32), (n, 1))
    return event.Photons(position, dirs, pol, wl, t=np.full(n, t0, dtype=np.float32))


def isotropic_rays(n, seed=0, radius=0.0):
    rs = np.random.get_state()
    np.random.seed(seed)
    dirs = sample.uniform_sphere(n).astype(np.float32)
    if radius > 0:
        pos = (sample.uniform_sphere(n) * radius * np.random.uniform(0, 1, n)[:, None] ** (1 / 3.)).astype(np.float32)
    else:
        pos = np.zeros((n, 3), dtype=np.float32)
    np.random.set_state(rs)
    return pos, dirs


def c4_detector_small():
    "config C4's detector (chroma_b200/demo/scintillator.py) with a sparser PMT shell, for CPU-sized tests"
    from chroma_b200.demo import scintillator
    return scintillator.detector(spiral_step=900.0)
