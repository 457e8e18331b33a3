"""An 8-inch PMT with light concentrator for the demo detector.

The reference revolves digitised SNO PMT / concentrator outlines
(``chroma/demo/sno_pmt.txt``, ``sno_cone.txt`` through ``chroma/pmt.py:40-85``).  Here the
outlines are analytic (ellipsoidal bulb on a cylindrical neck; conical concentrator) with
the same dimensions to a few mm and about the same number of profile points, so a PMT has
5,808 triangles (reference: 5,856).  The PMT faces +y; y = 0 is roughly the equator.
"""
import numpy as np

from chroma_b200.geometry import Solid
from chroma_b200.make import rotate_extrude
from chroma_b200.demo.optics import water, glass, vacuum, shiny_surface, r7081hqe_photocathode

NECK_RADIUS = 38.0     # mm
BASE_Y = -184.0        # mm
BULB_RADIUS = 101.0    # mm, equatorial radius of the bulb
BULB_HALF_HEIGHT = 75.0


def pmt_profile(npoints=55):
    "(x, y) outline from the base centre to the face centre, counter-clockwise."
    n_neck = 10
    n_bulb = npoints - n_neck - 2
    # ellipse centred at (0, 0); it meets the neck where x = NECK_RADIUS on the lower half
    t_neck = -np.arccos(NECK_RADIUS / BULB_RADIUS)
    y_join = BULB_HALF_HEIGHT * np.sin(t_neck)
    xs = [0.0] + [NECK_RADIUS] * n_neck
    ys = [BASE_Y] + list(np.linspace(BASE_Y, y_join, n_neck, endpoint=False))
    t = np.linspace(t_neck, np.pi / 2, n_bulb + 1)
    xs += list(BULB_RADIUS * np.cos(t))
    ys += list(BULB_HALF_HEIGHT * np.sin(t))
    xs[-1] = 0.0
    return np.array(xs), np.array(ys)


def offset_profile(x, y, d):
    "Move every profile point by `d` along the inward normal of the outline (closed at the axis)."
    p = np.column_stack([x, y])
    tang = np.gradient(p, axis=0)
    tang /= np.linalg.norm(tang, axis=1)[:, None]
    inward = np.column_stack([-tang[:, 1], tang[:, 0]])  # left of a counter-clockwise path
    q = p + d * inward
    q[0, 0] = 0.0
    q[-1, 0] = 0.0
    q[:, 0] = np.maximum(q[:, 0], 0.0)
    return q[:, 0], q[:, 1]


def build_8inch_pmt(outer_material=water, nsteps=24, glass_thickness=3.0):
    x, y = pmt_profile()
    xi, yi = offset_profile(x, y, glass_thickness)
    outer_mesh = rotate_extrude(x, y, nsteps)
    inner_mesh = rotate_extrude(xi, yi, nsteps)
    outer = Solid(outer_mesh, glass, outer_material)
    photocathode = np.mean(inner_mesh.assemble(), axis=1)[:, 1] > 0
    inner = Solid(inner_mesh, vacuum, glass,
                  surface=np.where(photocathode, r7081hqe_photocathode, shiny_surface),
                  color=np.where(photocathode, 0xff00, 0xff0000))
    pmt = outer + inner
    pmt.profile = np.column_stack([x, y])
    pmt.nsteps = nsteps
    return pmt


def build_light_collector(outer_material=water, nsteps=24, npoints=16):
    "Conical reflector in front of the bulb: radius 98 -> 134 mm over y = 22 -> 130 mm."
    y = np.linspace(22.0, 130.0, npoints)
    s = (y - y[0]) / (y[-1] - y[0])
    r = 98.0 + 36.0 * np.sqrt(np.clip(1 - (1 - s) ** 2, 0, 1))
    return Solid(rotate_extrude(r, y, nsteps), outer_material, outer_material,
                 surface=shiny_surface)


def build_8inch_pmt_with_lc(outer_material=water, nsteps=24):
    return build_8inch_pmt(outer_material, nsteps) + build_light_collector(outer_material, nsteps)
