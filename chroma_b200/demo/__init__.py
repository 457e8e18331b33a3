"""Demo detector: PMTs on a spherical spiral inside a black water sphere
(API of ``chroma/demo/__init__.py``: ``detector()``, ``tiny()``)."""
from math import sin, cos, sqrt

import numpy as np

from chroma_b200.detector import Detector
from chroma_b200.geometry import Solid, Mesh
from chroma_b200.make import sphere
from chroma_b200.transform import make_rotation_matrix, normalize
from chroma_b200.demo.optics import water, black_surface
from chroma_b200.demo.pmt import build_8inch_pmt_with_lc


def spherical_spiral(radius, spacing):
    "Points roughly `spacing` apart on a spiral wrapping a sphere (demo/__init__.py:19-30)."
    dl = spacing / radius
    t = 0.0
    a = np.pi / dl
    while t < np.pi:
        yield np.array([sin(t) * sin(a * t), sin(t) * cos(a * t), cos(t)]) * radius
        t += dl / sqrt(1 + a ** 2 * sin(t) ** 2)


def pmt_placements(pmt_radius, spiral_step):
    "(rotations (n,3,3), positions (n,3)) of PMTs facing the centre."
    rotations, positions = [], []
    y_axis = np.array((0.0, 1.0, 0.0))
    for position in spherical_spiral(pmt_radius, spiral_step):
        direction = -normalize(position)
        axis = np.cross(direction, y_axis)
        angle = np.arccos(np.clip(np.dot(y_axis, direction), -1, 1))
        if np.linalg.norm(axis) < 1e-12:
            axis = np.array((1.0, 0.0, 0.0))
        rotations.append(make_rotation_matrix(angle, axis))
        positions.append(position)
    return np.array(rotations), np.array(positions)


def _set_daq_dists(geo):
    time_rms, charge_mean, charge_rms = 1.5, 1.0, 0.1
    geo.set_time_dist_gaussian(time_rms, -5 * time_rms, 5 * time_rms)
    geo.set_charge_dist_gaussian(charge_mean, charge_rms, 0.0, charge_mean + 5 * charge_rms)


def detector(pmt_radius=14000.0, sphere_radius=14500.0, spiral_step=350.0, fast=None):
    """The demo detector.  Defaults give 10,055 PMTs; spiral_step=206 gives 29,023.

    fast=True assembles the flattened arrays directly (vectorised instancing, no vertex
    de-duplication) instead of going through 10^4 add_pmt calls and np.unique; the default
    picks it above 2,000 PMTs."""
    pmt = build_8inch_pmt_with_lc()
    rotations, positions = pmt_placements(pmt_radius, spiral_step)
    outer = Solid(sphere(sphere_radius, nsteps=200), water, water, surface=black_surface,
                  color=0xBBFFFFFF)
    if fast is None:
        fast = len(positions) > 2000
    geo = Detector(water)
    if not fast:
        geo.add_solid(outer)
        for rotation, position in zip(rotations, positions):
            geo.add_pmt(pmt, rotation, position)
    else:
        # the flattened arrays (168.6 M triangles at 29,023 PMTs: ~9 s and 14 GB of host memory) are
        # assembled on the first flatten(); a two-level GPUGeometry(instanced=True) never needs them
        n = len(positions)
        rot32 = np.concatenate([np.identity(3, dtype=np.float32)[None], rotations.astype(np.float32)])
        pos32 = np.concatenate([np.zeros((1, 3), np.float32), positions.astype(np.float32)])
        tmpl = np.ones(n + 1, dtype=np.uint32)
        tmpl[0] = 0
        geo._placements = ([outer, pmt], tmpl, rot32, pos32)
        geo.solids = [outer, pmt]  # placements are not expanded into 10^4 python objects
        geo.solid_id_to_channel_index = np.concatenate([[-1], np.arange(n)]).astype(np.int32)
        geo.channel_index_to_solid_id = np.arange(1, n + 1, dtype=np.int32)
        geo.channel_index_to_channel_type = np.arange(n, dtype=np.int32)
        geo.channel_index_to_position = positions.astype(np.float32)
        geo._deferred_flatten = lambda: _fast_assemble(geo, outer, pmt, rotations, positions)
    _set_daq_dists(geo)
    return geo


def _fast_assemble(geo, outer, pmt, rotations, positions):
    "Fill the attributes Geometry.flatten()/Detector.flatten() would produce."
    n = len(positions)
    rot32 = rotations.astype(np.float32)
    pos32 = positions.astype(np.float32)
    pv = pmt.mesh.vertices
    nv0, nt0 = len(outer.mesh.vertices), len(outer.mesh.triangles)
    nvp, ntp = len(pv), len(pmt.mesh.triangles)
    vertices = np.empty((nv0 + n * nvp, 3), dtype=np.float32)
    vertices[:nv0] = outer.mesh.vertices
    # np.inner(v, R) + d for every placement (chroma/geometry.py:363-364)
    inst = np.einsum('vk,njk->nvj', pv.astype(np.float64), rot32.astype(np.float64)) + \
        pos32[:, None, :].astype(np.float64)
    vertices[nv0:] = inst.reshape(-1, 3).astype(np.float32)
    del inst
    triangles = np.empty((nt0 + n * ntp, 3), dtype=np.uint32)
    triangles[:nt0] = outer.mesh.triangles
    offs = (nv0 + np.arange(n, dtype=np.uint32) * np.uint32(nvp))[:, None, None]
    triangles[nt0:] = (pmt.mesh.triangles.astype(np.uint32)[None, :, :] + offs).reshape(-1, 3)
    geo.mesh = Mesh(vertices, triangles, remove_duplicate_vertices=False, round=False,
                    remove_null_triangles=False)
    geo.solids = [outer, pmt]  # placements are not expanded into 10^4 python objects
    geo.colors = np.concatenate([outer.color, np.tile(pmt.color, n)])
    geo.solid_id = np.concatenate([np.zeros(nt0, np.uint32),
                                   np.repeat(np.arange(1, n + 1, dtype=np.uint32), ntp)])
    mats, seen = [], set()
    for m in list(outer.unique_materials) + list(pmt.unique_materials):
        if id(m) not in seen:
            seen.add(id(m))
            mats.append(m)
    surfs, seen = [], set()
    for s in list(outer.unique_surfaces) + list(pmt.unique_surfaces):
        if id(s) not in seen:
            seen.add(id(s))
            surfs.append(s)
    geo.unique_materials, geo.unique_surfaces = mats, surfs
    ml = {id(m): i for i, m in enumerate(mats)}
    sl = {id(s): i for i, s in enumerate(surfs)}

    def idx(a, lookup):
        return np.fromiter((lookup[id(x)] for x in a), dtype=np.int32, count=len(a))

    def both(attr, lookup):
        return np.concatenate([idx(getattr(outer, attr), lookup),
                               np.tile(idx(getattr(pmt, attr), lookup), n)])

    geo.inner_material_index = both('inner_material', ml)
    geo.outer_material_index = both('outer_material', ml)
    geo.surface_index = both('surface', sl)
    if id(None) in sl:
        geo.surface_index[geo.surface_index == sl[id(None)]] = -1
    geo.solid_id_to_channel_index = np.concatenate([[-1], np.arange(n)]).astype(np.int32)
    geo.channel_index_to_solid_id = np.arange(1, n + 1, dtype=np.int32)
    geo.channel_index_to_channel_type = np.arange(n, dtype=np.int32)
    geo.channel_index_to_position = pos32


def tiny():
    return detector(2000.0, 2500.0, 700.0)
