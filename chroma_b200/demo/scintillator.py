"""Synthetic detector of config C4 (no such geometry ships with the reference, only the
models do: chroma/cuda/photon.h:193-308 bulk reemission, :592-637 WLS, :640-670 dichroic;
chroma/geometry.py:257-281 Surface / DichroicProps): a two-component scintillator ball in water,
surrounded by a shell of 8-inch PMTs with light concentrators (demo.pmt), each behind a wavelength-
shifting plate and a short-pass dichroic disc, inside a black sphere.

One placed solid ("module") = PMT + concentrator + dichroic disc + WLS plate, so the detector is
`npmt` placements of one template plus two single solids (what the two-level geometry wants)."""
import numpy as np

from chroma_b200.detector import Detector
from chroma_b200.geometry import Solid, Material, Surface, DichroicProps, standard_wavelengths
from chroma_b200.make import sphere, rotate_extrude
from chroma_b200.demo import pmt_placements, _set_daq_dists
from chroma_b200.demo.optics import water, black_surface
from chroma_b200.demo.pmt import build_8inch_pmt_with_lc


def gaussian_cdf(grid, mean, sigma):
    pdf = np.exp(-0.5 * ((grid - mean) / sigma) ** 2)
    cdf = np.cumsum(pdf)
    cdf -= cdf[0]
    return cdf / cdf[-1]


def scintillator_material():
    "absorption 2 m shared by two reemitting components: 70 % -> 430 +- 30 nm / 5 ns, 30 % -> 480 +- 30 nm / 20 ns"
    w = standard_wavelengths.astype(np.float64)
    times = np.arange(0, 1000, 0.05)
    m = Material('c4_scintillator')
    m.set('refractive_index', 1.5)
    m.set('absorption_length', np.where(w < 440, 2000.0, 20000.0))
    m.set('scattering_length', 8000.0)
    for frac, mean, tau in ((0.7, 430.0, 5.0), (0.3, 480.0, 20.0)):
        m.comp_absorption_length.append(
            np.column_stack([w, np.where(w < 440, 2000.0, 20000.0) / frac]).astype(np.float32))
        m.comp_reemission_prob.append(np.column_stack([w, np.where(w < 440, 0.8, 0.05)]).astype(np.float32))
        m.comp_reemission_wvl_cdf.append(np.column_stack([w, gaussian_cdf(w, mean, 30.0)]).astype(np.float32))
        tc = 1 - np.exp(-times / tau)
        m.comp_reemission_time_cdf.append(np.column_stack([times, tc / tc[-1]]))
    return m


def wls_surface():
    "absorbs below 400 nm, re-emits 90 % of that at 420 +- 20 nm (photon.h:592-637)"
    w = standard_wavelengths.astype(np.float64)
    s = Surface('c4_wls', model=2)
    s.set('absorb', np.where(w < 400, 0.8, 0.02))
    s.set('reemit', 0.9)
    s.set('reflect_specular', 0.02)
    s.set('reflect_diffuse', 0.02)
    s.set('reemission_cdf', gaussian_cdf(w, 420.0, 20.0))
    return s


def dichroic_surface(nangles=5):
    "short-pass at 450 nm, the edge moving down with the angle of incidence (photon.h:640-670)"
    w = standard_wavelengths.astype(np.float64)
    angles = np.linspace(0, np.pi / 2, nangles)
    edge = [450.0 - 40.0 * np.sin(a) ** 2 for a in angles]
    refl = [np.column_stack([w, np.where(w > e, 0.90, 0.04)]) for e in edge]
    tran = [np.column_stack([w, np.where(w > e, 0.06, 0.94)]) for e in edge]
    s = Surface('c4_dichroic', model=3)
    s.dichroic_props = DichroicProps(angles, refl, tran)
    return s


def disc(radius, y, nsteps=24):
    "flat disc of triangles normal to y (the axis the demo PMT faces)"
    return rotate_extrude([0.0, 0.5 * radius, radius], [y, y, y], nsteps)


def build_module():
    "PMT + concentrator (mouth radius 134 mm at y = 130 mm) + dichroic disc at y = 140 + WLS plate at y = 160"
    module = build_8inch_pmt_with_lc()
    module = module + Solid(disc(134.0, 140.0), water, water, surface=dichroic_surface(), color=0x80ff00ff)
    module = module + Solid(disc(134.0, 160.0), water, water, surface=wls_surface(), color=0x8000ffff)
    return module


def detector(pmt_radius=4200.0, sphere_radius=4800.0, spiral_step=350.0, scintillator_radius=1000.0):
    """Default: 909 modules on a 4.2 m sphere around a 1 m scintillator ball ("a
    tiny()-style shell of 53-1,000 PMTs each behind a dichroic disc and a WLS plate")."""
    det = Detector(water)
    det.add_solid(Solid(sphere(scintillator_radius, nsteps=64), scintillator_material(), water, color=0x4000ff00))
    det.add_solid(Solid(sphere(sphere_radius, nsteps=96), water, water, surface=black_surface, color=0xBBFFFFFF))
    module = build_module()
    rotations, positions = pmt_placements(pmt_radius, spiral_step)
    for rotation, position in zip(rotations, positions):
        det.add_pmt(module, rotation, position)
    _set_daq_dists(det)
    return det


def event_vertex(gid, radius=1000.0):
    "vertex of event `gid`: uniform in the scintillator ball"
    rs = np.random.RandomState(200003 + int(gid))
    v = rs.normal(size=3)
    return (v / np.linalg.norm(v) * 0.98 * radius * rs.uniform() ** (1.0 / 3.0)).astype(np.float32)
