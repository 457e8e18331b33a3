"""ctypes front-end of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / reference arm.  Never imported from chroma_b200/.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, 'liboracle.so')


class OracleGeometry(C.Structure):
    _fields_ = [
        ('vertices', C.c_void_p), ('triangles', C.c_void_p), ('material_codes', C.c_void_p),
        ('solid_id', C.c_void_p), ('nvertices', C.c_uint64), ('ntriangles', C.c_uint64),
        ('nodes', C.c_void_p), ('nnodes', C.c_uint64),
        ('world_origin', C.c_float * 3), ('world_scale', C.c_float),
        ('nwavelengths', C.c_uint32), ('wavelength_start', C.c_float), ('wavelength_step', C.c_float),
        ('ntimes', C.c_uint32), ('time_start', C.c_float), ('time_step', C.c_float),
        ('nmaterials', C.c_uint32), ('material_props', C.c_void_p), ('material_num_comp', C.c_void_p),
        ('ncomp_total', C.c_uint32),
        ('comp_reemission_prob', C.c_void_p), ('comp_reemission_wvl_cdf', C.c_void_p),
        ('comp_absorption_length', C.c_void_p), ('comp_reemission_time_cdf', C.c_void_p),
        ('nsurfaces', C.c_uint32), ('surface_present', C.c_void_p), ('surface_props', C.c_void_p),
        ('surface_model', C.c_void_p), ('surface_transmissive', C.c_void_p),
        ('surface_thickness', C.c_void_p),
        ('dichroic_nangles', C.c_void_p), ('dichroic_angles', C.c_void_p),
        ('dichroic_reflect', C.c_void_p), ('dichroic_transmit', C.c_void_p),
        ('solid_id_to_channel_index', C.c_void_p), ('nsolids', C.c_uint64), ('nchannels', C.c_uint32),
        ('time_cdf_x', C.c_void_p), ('time_cdf_y', C.c_void_p), ('time_cdf_len', C.c_uint32),
        ('charge_cdf_x', C.c_void_p), ('charge_cdf_y', C.c_void_p), ('charge_cdf_len', C.c_uint32),
    ]


_lib = None


def build(force=False):
    "Compile liboracle.so (g++, OpenMP) if it is missing or stale."
    src = [os.path.join(HERE, f) for f in ('chroma_oracle.cpp', 'chroma_oracle.h')]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in src):
        subprocess.run(['make', '-C', HERE, 'oracle'], check=True, capture_output=True)
    return LIB


def load():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
        _lib.oracle_num_threads.restype = C.c_int
        _lib.oracle_build_bvh.restype = C.c_uint64
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


def num_threads():
    return load().oracle_num_threads()


def set_num_threads(n):
    load().oracle_set_num_threads(int(n))


class Oracle(object):
    """CPU oracle bound to one (flattened, BVH-carrying) Geometry or Detector."""

    def __init__(self, geometry, wavelengths=None, times=None):
        from chroma_b200.geometry import standard_wavelengths
        from chroma_b200.gpu.geometry import pack_optics  # host-side table packing only
        self.lib = load()
        if wavelengths is None:
            wavelengths = standard_wavelengths
        if times is None:
            times = np.arange(0, 1000, 0.05)
        if not hasattr(geometry, 'mesh'):
            geometry.flatten()
        opt = pack_optics(geometry, np.asarray(wavelengths, dtype=np.float32), times)
        k = self.keep = dict(opt)
        k['vertices'] = _c(geometry.mesh.vertices, np.float32)
        k['triangles'] = _c(geometry.mesh.triangles, np.uint32)
        k['material_codes'] = (((geometry.inner_material_index & 0xff) << 24) |
                               ((geometry.outer_material_index & 0xff) << 16) |
                               ((geometry.surface_index & 0xff) << 8)).astype(np.uint32)
        k['solid_id'] = _c(geometry.solid_id, np.uint32)
        k['colors'] = _c(geometry.colors, np.uint32)
        k['nodes'] = np.ascontiguousarray(geometry.bvh.nodes).view(np.uint32).reshape(-1, 4)
        g = self.g = OracleGeometry()
        g.nvertices, g.ntriangles, g.nnodes = len(k['vertices']), len(k['triangles']), len(k['nodes'])
        g.world_origin = (C.c_float * 3)(*[float(x) for x in geometry.bvh.world_coords.world_origin])
        g.world_scale = float(geometry.bvh.world_coords.world_scale)
        g.nwavelengths, g.wavelength_start = len(wavelengths), float(wavelengths[0])
        g.wavelength_step = float(wavelengths[1] - wavelengths[0])
        g.ntimes, g.time_start, g.time_step = len(times), float(times[0]), float(times[1] - times[0])
        g.nmaterials, g.nsurfaces = len(geometry.unique_materials), len(geometry.unique_surfaces)
        g.ncomp_total = opt['ncomp_total']
        for name in ('vertices', 'triangles', 'material_codes', 'solid_id', 'nodes', 'material_props',
                     'material_num_comp', 'comp_reemission_prob', 'comp_reemission_wvl_cdf',
                     'comp_absorption_length', 'comp_reemission_time_cdf', 'surface_present',
                     'surface_props', 'surface_model', 'surface_transmissive', 'surface_thickness',
                     'dichroic_nangles', 'dichroic_angles', 'dichroic_reflect', 'dichroic_transmit'):
            setattr(g, name, _p(k[name]) if k[name].size else None)
        self.nchannels = 0
        if hasattr(geometry, 'num_channels'):
            k['s2c'] = _c(geometry.solid_id_to_channel_index, np.int32)
            k['tx'], k['ty'] = _c(geometry.time_cdf[0], np.float32), _c(geometry.time_cdf[1], np.float32)
            k['qx'], k['qy'] = _c(geometry.charge_cdf[0], np.float32), _c(geometry.charge_cdf[1], np.float32)
            g.solid_id_to_channel_index, g.nsolids = _p(k['s2c']), len(k['s2c'])
            g.nchannels = self.nchannels = geometry.num_channels()
            g.time_cdf_x, g.time_cdf_y, g.time_cdf_len = _p(k['tx']), _p(k['ty']), len(k['tx'])
            g.charge_cdf_x, g.charge_cdf_y, g.charge_cdf_len = _p(k['qx']), _p(k['qy']), len(k['qx'])

    def intersect(self, pos, dir, last_hit=None, counts=False):
        pos, dir = _c(pos, np.float32), _c(dir, np.float32)
        n = len(pos)
        last = _c(last_hit, np.int32) if last_hit is not None else None
        tri, dist = np.empty(n, np.int32), np.empty(n, np.float32)
        cnt = np.empty((n, 2), np.uint32) if counts else None
        self.lib.oracle_intersect(C.byref(self.g), C.c_uint64(n), _p(pos), _p(dir), _p(last), _p(tri),
                                  _p(dist), _p(cnt))
        return (tri, dist, cnt) if counts else (tri, dist)

    def intersect_ordered(self, pos, dir, last_hit=None):
        "near-first search of the same tree: (tri, dist, counts[n][2]); see chroma_oracle.cpp"
        pos, dir = _c(pos, np.float32), _c(dir, np.float32)
        n = len(pos)
        last = _c(last_hit, np.int32) if last_hit is not None else None
        tri, dist = np.empty(n, np.int32), np.empty(n, np.float32)
        cnt = np.empty((n, 2), np.uint32)
        self.lib.oracle_intersect_ordered(C.byref(self.g), C.c_uint64(n), _p(pos), _p(dir), _p(last),
                                          _p(tri), _p(dist), _p(cnt))
        return tri, dist, cnt

    def propagate(self, photons, seed, id_base=0, epoch=0, max_steps=10, use_weights=False,
                  scatter_first=0, stats=False):
        "Returns a new event.Photons (input untouched); stats=True also (steps,nodes,tris) per photon."
        from chroma_b200 import event
        n = len(photons)
        pos, dir, pol = [np.array(a, dtype=np.float32, order='C') for a in (photons.pos, photons.dir, photons.pol)]
        wl, t = np.array(photons.wavelengths, np.float32), np.array(photons.t, np.float32)
        last, flags = np.array(photons.last_hit_triangles, np.int32), np.array(photons.flags, np.uint32)
        w = np.array(photons.weights, np.float32)
        st = np.zeros((n, 3), np.uint64) if stats else None
        self.lib.oracle_propagate(C.byref(self.g), C.c_uint64(n), _p(pos), _p(dir), _p(pol), _p(wl), _p(t),
                                  _p(last), _p(flags), _p(w), C.c_uint64(seed), C.c_uint32(id_base),
                                  C.c_uint32(epoch), C.c_int(max_steps), C.c_int(int(use_weights)),
                                  C.c_int(scatter_first), _p(st))
        out = event.Photons(pos, dir, pol, wl, t, last, flags, w, np.array(photons.evidx, np.uint32))
        return (out, st) if stats else out

    def render(self, pos, dir, alpha_depth=10, bg_color=0, state=None):
        """render.cu:35-181 on the host.  Returns (pixels, state) with state = (dx, dxlen, color) to
        pass back in for keep_last_render."""
        pos, dir = _c(pos, np.float32), _c(dir, np.float32)
        n = len(pos)
        if state is None:
            state = (np.zeros((n, alpha_depth), np.float32), np.zeros(n, np.uint32),
                     np.zeros((n, alpha_depth, 4), np.float32))
        dx, dxlen, color = state
        pixels = np.empty(n, np.uint32)
        colors = self.keep['colors']
        self.lib.oracle_render(C.byref(self.g), _p(colors), C.c_uint64(n), _p(pos), _p(dir),
                               C.c_uint32(alpha_depth), _p(pixels), _p(dx), _p(dxlen), _p(color),
                               C.c_uint32(bg_color))
        return pixels, state

    def run_daq(self, photons, seed, id_base=0, epoch=0, weight=1.0):
        from chroma_b200 import event
        n = len(photons)
        t, q = np.empty(self.nchannels, np.float32), np.empty(self.nchannels, np.float32)
        h = np.empty(self.nchannels, np.uint32)
        self.lib.oracle_run_daq(C.byref(self.g), C.c_uint64(n), _p(_c(photons.t, np.float32)),
                                _p(_c(photons.flags, np.uint32)), _p(_c(photons.last_hit_triangles, np.int32)),
                                _p(_c(photons.weights, np.float32)), C.c_uint64(seed), C.c_uint32(id_base),
                                C.c_uint32(epoch), C.c_float(weight), _p(t), _p(q), _p(h))
        return event.Channels(t < 1e8, t, q, h)

    def run_daq_many(self, photons, seed, ndaq, id_base=0, epoch=0, weight=1.0, convert_all=True):
        "daq.cu:88-150 with `ndaq` replicas -> (t, q, flags) arrays of shape [ndaq, nchannels]"
        return _run_daq_many(self.lib, 'oracle_run_daq_many', self.g, self.nchannels, photons, seed, id_base,
                             epoch, weight, ndaq, convert_all)


def _run_daq_many(lib, fn, g, nchannels, photons, seed, id_base, epoch, weight, ndaq, convert_all):
    "shared driver of oracle_run_daq_many / ref_cpu_run_daq_many -> (t, q, flags), each [ndaq, nchannels]"
    n = len(photons)
    tot = nchannels * ndaq
    t, q, h = np.empty(tot, np.float32), np.empty(tot, np.float32), np.empty(tot, np.uint32)
    getattr(lib, fn)(C.byref(g), C.c_uint64(n), _p(_c(photons.t, np.float32)), _p(_c(photons.flags, np.uint32)),
                     _p(_c(photons.last_hit_triangles, np.int32)), _p(_c(photons.weights, np.float32)),
                     C.c_uint64(seed), C.c_uint32(id_base), C.c_uint32(epoch), C.c_float(weight), C.c_int(ndaq),
                     C.c_int(int(convert_all)), _p(t), _p(q), _p(h))
    return t.reshape(ndaq, nchannels), q.reshape(ndaq, nchannels), h.reshape(ndaq, nchannels)


def bin_hits(channels, hitcount, pdf, trange, qrange):
    "In-place pdf.cu:9-32 on numpy arrays (pdf shape [nch, tbins, qbins])."
    lib = load()
    nch, tbins, qbins = pdf.shape
    lib.oracle_bin_hits(C.c_int(nch), _p(_c(channels.q, np.float32)), _p(_c(channels.t, np.float32)),
                        _p(hitcount), C.c_int(tbins), C.c_float(trange[0]), C.c_float(trange[1]),
                        C.c_int(qbins), C.c_float(qrange[0]), C.c_float(qrange[1]), _p(pdf))


def accumulate_pdf_eval(event_hit, event_time, mc_time, hitcount, bincount, nearest_mc, min_twidth,
                        trange, min_bin_content):
    """In-place pdf.cu:34-219 (accumulate_bincount + accumulate_nearest_neighbor) on numpy arrays:
    mc_time [ndaq, nch] float32, hitcount/bincount uint32[nch], nearest_mc float32[nhit, k]."""
    lib = load()
    mc_time = _c(mc_time, np.float32).reshape(-1, len(event_hit))
    assert nearest_mc.shape == (int(np.count_nonzero(event_hit)), min_bin_content)
    lib.oracle_accumulate_pdf_eval(C.c_int(len(event_hit)), C.c_int(mc_time.shape[0]),
                                   _p(_c(event_hit, np.uint32)), _p(_c(event_time, np.float32)), _p(mc_time),
                                   _p(hitcount), _p(bincount), C.c_float(min_twidth), C.c_float(trange[0]),
                                   C.c_float(trange[1]), C.c_int(min_bin_content), _p(nearest_mc))


def accumulate_moments(time_only, mc_time, mc_charge, trange, qrange, mom0, tmom1, tmom2, qmom1, qmom2):
    "In-place pdf.cu:223-266."
    lib = load()
    lib.oracle_accumulate_moments(C.c_int(int(time_only)), C.c_int(len(mom0)), _p(_c(mc_time, np.float32)),
                                  _p(_c(mc_charge, np.float32)), C.c_float(trange[0]), C.c_float(trange[1]),
                                  C.c_float(qrange[0]), C.c_float(qrange[1]), _p(mom0), _p(tmom1), _p(tmom2),
                                  _p(qmom1), _p(qmom2))


def accumulate_kernel_eval(time_only, event_hit, event_time, event_charge, mc_time, mc_charge, trange, qrange,
                           inv_tbw, inv_qbw, hitcount, time_pdf, charge_pdf):
    "In-place pdf.cu:268-368 (terms evaluated in double, see chroma_oracle.cpp)."
    lib = load()
    lib.oracle_accumulate_kernel_eval(C.c_int(int(time_only)), C.c_int(len(hitcount)),
                                      _p(_c(event_hit, np.uint32)), _p(_c(event_time, np.float32)),
                                      _p(_c(event_charge, np.float32)), _p(_c(mc_time, np.float32)),
                                      _p(_c(mc_charge, np.float32)), C.c_float(trange[0]), C.c_float(trange[1]),
                                      C.c_float(qrange[0]), C.c_float(qrange[1]), _p(_c(inv_tbw, np.float32)),
                                      _p(_c(inv_qbw, np.float32)), _p(hitcount), _p(time_pdf), _p(charge_pdf))


def make_leaves(mesh, world_origin, world_scale):
    lib = load()
    v, t = _c(mesh.vertices, np.float32), _c(mesh.triangles, np.uint32)
    nodes, codes = np.empty((len(t), 4), np.uint32), np.empty(len(t), np.uint64)
    wo = (C.c_float * 3)(*[float(x) for x in world_origin])
    lib.oracle_make_leaves(_p(v), _p(t), C.c_uint64(len(t)), wo, C.c_float(world_scale), _p(nodes), _p(codes))
    return nodes, codes


def build_bvh(mesh, target_degree=3, leaves=None, codes=None, world_coords=None):
    """BVH(world_coords, nodes, layer_offsets) built by the CPU restatement of bvh/grid.py.
    With `leaves` (uint32[n][4]), `codes` (uint64[n], 48-bit) and `world_coords` given, only the
    host part of the builder runs on them (the reference makes its leaves on the GPU)."""
    from chroma_b200.bvh.bvh import BVH, WorldCoords
    from chroma_b200.gpu.tools import uint4
    lib = load()
    v, t = _c(mesh.vertices, np.float32), _c(mesh.triangles, np.uint32)
    nodes = np.empty((2 * len(t) + 1024, 4), np.uint32)
    offsets = np.empty(64, np.uint64)
    nlayers, wo, ws = C.c_uint32(), (C.c_float * 3)(), C.c_float()
    if leaves is not None:
        lv = np.ascontiguousarray(leaves).view(np.uint32).reshape(-1, 4)
        cd = _c(codes, np.uint64)
        assert len(lv) == len(t) == len(cd)
        lib.oracle_build_bvh_from_leaves.restype = C.c_uint64
        n = lib.oracle_build_bvh_from_leaves(_p(lv), _p(cd), C.c_uint64(len(t)), C.c_int(target_degree),
                                             _p(nodes), _p(offsets), C.byref(nlayers))
        wc = world_coords
    else:
        n = lib.oracle_build_bvh(_p(v), C.c_uint64(len(v)), _p(t), C.c_uint64(len(t)), C.c_int(target_degree),
                                 _p(nodes), _p(offsets), C.byref(nlayers), wo, C.byref(ws))
        wc = WorldCoords(np.array(list(wo), np.float32), np.float32(ws.value))
    if n == 0:
        raise RuntimeError('oracle_build_bvh failed')
    packed = np.ascontiguousarray(nodes[:n]).view(uint4)[:, 0]
    return BVH(wc, packed, [int(x) for x in offsets[:nlayers.value]])


def uniform_stream(seed, id, epoch, stream, n):
    out = np.empty(n, np.float32)
    load().oracle_uniform_stream(C.c_uint64(seed), C.c_uint32(id), C.c_uint32(epoch), C.c_uint32(stream),
                                 C.c_uint32(n), _p(out))
    return out


def generate_photons(n, id0, seed, sequence, rot, cos_half, wl_lo, wl_hi):
    "(dir, pol, wavelengths) of the device photon source, restated on the CPU (chroma_oracle.cpp)."
    d = np.empty((n, 3), np.float32)
    p = np.empty((n, 3), np.float32)
    w = np.empty(n, np.float32)
    r = np.ascontiguousarray(rot, dtype=np.float32).ravel()
    load().oracle_generate_photons(C.c_uint64(n), C.c_uint32(id0), C.c_uint64(seed), C.c_uint32(sequence),
                                   _p(r), C.c_float(cos_half), C.c_float(wl_lo), C.c_float(wl_hi),
                                   _p(d), _p(p), _p(w))
    return d, p, w
