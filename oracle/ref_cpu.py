"""ctypes binding of oracle/_ref/libchroma_ref_cpu.so: the reference's OWN propagate.cu / daq.cu /
mesh.h compiled for the host (oracle/ref_cpu_wrapper.cpp, `make -C oracle ref_cpu`).

TEST INFRASTRUCTURE ONLY: used by tests/ to pin the oracle port against the reference's source,
and by bench.py's CPU legs (`cpu_baseline.kind = "reference"`).  Same call shapes as
oracle.Oracle so the two can be compared line by line."""
import ctypes as C
import os

import numpy as np

from oracle import oracle as O

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, '_ref', 'libchroma_ref_cpu.so')
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def load():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
        _lib.ref_cpu_num_threads.restype = C.c_int
    return _lib


def num_threads():
    return int(load().ref_cpu_num_threads())


class RefCpu(object):
    "The reference's kernels on the host, over the flat geometry description oracle.Oracle builds."

    def __init__(self, geometry, oracle=None):
        self.lib = load()
        self.oracle = oracle if oracle is not None else O.Oracle(geometry)   # owns the packed arrays
        self.g = self.oracle.g
        self.nchannels = self.oracle.nchannels

    def intersect(self, pos, dir, last_hit=None):
        pos, dir = O._c(pos, np.float32), O._c(dir, np.float32)
        n = len(pos)
        last = O._c(last_hit, np.int32) if last_hit is not None else None
        tri, dist = np.empty(n, np.int32), np.empty(n, np.float32)
        self.lib.ref_cpu_intersect(C.byref(self.g), C.c_uint64(n), O._p(pos), O._p(dir), O._p(last), O._p(tri),
                                   O._p(dist))
        return tri, dist

    def propagate(self, photons, seed, id_base=0, epoch=0, max_steps=10, use_weights=False, scatter_first=0):
        "Returns a new event.Photons (input untouched)."
        from chroma_b200 import event
        n = len(photons)
        pos, dir, pol = [np.array(a, dtype=np.float32, order='C') for a in (photons.pos, photons.dir, photons.pol)]
        wl, t = np.array(photons.wavelengths, np.float32), np.array(photons.t, np.float32)
        last, flags = np.array(photons.last_hit_triangles, np.int32), np.array(photons.flags, np.uint32)
        w, ev = np.array(photons.weights, np.float32), np.array(photons.evidx, np.uint32)
        self.lib.ref_cpu_propagate(C.byref(self.g), C.c_uint64(n), O._p(pos), O._p(dir), O._p(pol), O._p(wl),
                                   O._p(t), O._p(last), O._p(flags), O._p(w), O._p(ev), C.c_uint64(seed),
                                   C.c_uint32(id_base), C.c_uint32(epoch), C.c_int(max_steps),
                                   C.c_int(int(use_weights)), C.c_int(scatter_first))
        return event.Photons(pos, dir, pol, wl, t, last, flags, w, ev)

    def run_daq(self, photons, seed, id_base=0, epoch=0, weight=1.0):
        from chroma_b200 import event
        n = len(photons)
        t, q = np.empty(self.nchannels, np.float32), np.empty(self.nchannels, np.float32)
        h = np.empty(self.nchannels, np.uint32)
        self.lib.ref_cpu_run_daq(C.byref(self.g), C.c_uint64(n), O._p(np.array(photons.t, np.float32)),
                                 O._p(np.array(photons.flags, np.uint32)),
                                 O._p(np.array(photons.last_hit_triangles, np.int32)),
                                 O._p(np.array(photons.weights, np.float32)), C.c_uint64(seed), C.c_uint32(id_base),
                                 C.c_uint32(epoch), C.c_float(weight), O._p(t), O._p(q), O._p(h))
        return event.Channels(t < 1e8, t, q, h)


    def run_daq_many(self, photons, seed, ndaq, id_base=0, epoch=0, weight=1.0, convert_all=True):
        "the reference's run_daq_many (daq.cu:88-150) -> (t, q, flags) of shape [ndaq, nchannels]"
        return O._run_daq_many(self.lib, 'ref_cpu_run_daq_many', self.g, self.nchannels, photons, seed, id_base,
                               epoch, weight, ndaq, convert_all)


# ---- chroma/cuda/pdf.cu and render.cu: their own translation units, their own libraries ----------
_pdf_lib = _render_lib = None


def _sub(name):
    return C.CDLL(os.path.join(_HERE, '_ref', name))


def pdf_lib():
    global _pdf_lib
    if _pdf_lib is None:
        _pdf_lib = _sub('libchroma_ref_cpu_pdf.so')
    return _pdf_lib


def render_lib():
    global _render_lib
    if _render_lib is None:
        _render_lib = _sub('libchroma_ref_cpu_render.so')
    return _render_lib


def bin_hits(channels, hitcount, pdf, trange, qrange):
    "In-place reference bin_hits (pdf.cu:9-32) on numpy arrays (pdf shape [nch, tbins, qbins])."
    nch, tbins, qbins = pdf.shape
    pdf_lib().ref_cpu_bin_hits(C.c_int(nch), O._p(O._c(channels.q, np.float32)), O._p(O._c(channels.t, np.float32)),
                               O._p(hitcount), C.c_int(tbins), C.c_float(trange[0]), C.c_float(trange[1]),
                               C.c_int(qbins), C.c_float(qrange[0]), C.c_float(qrange[1]), O._p(pdf))


def accumulate_pdf_eval(event_hit, event_time, mc_time, hitcount, bincount, nearest_mc, min_twidth, trange,
                        min_bin_content):
    "In-place reference accumulate_bincount + accumulate_nearest_neighbor_block (same shapes as oracle.py)."
    mc_time = O._c(mc_time, np.float32).reshape(-1, len(event_hit))
    assert nearest_mc.shape == (int(np.count_nonzero(event_hit)), min_bin_content)
    pdf_lib().ref_cpu_accumulate_pdf_eval(C.c_int(len(event_hit)), C.c_int(mc_time.shape[0]),
                                          O._p(O._c(event_hit, np.uint32)), O._p(O._c(event_time, np.float32)),
                                          O._p(mc_time), O._p(hitcount), O._p(bincount), C.c_float(min_twidth),
                                          C.c_float(trange[0]), C.c_float(trange[1]), C.c_int(min_bin_content),
                                          O._p(nearest_mc))


def accumulate_moments(time_only, mc_time, mc_charge, trange, qrange, mom0, tmom1, tmom2, qmom1, qmom2):
    pdf_lib().ref_cpu_accumulate_moments(C.c_int(int(time_only)), C.c_int(len(mom0)), O._p(O._c(mc_time, np.float32)),
                                         O._p(O._c(mc_charge, np.float32)), C.c_float(trange[0]), C.c_float(trange[1]),
                                         C.c_float(qrange[0]), C.c_float(qrange[1]), O._p(mom0), O._p(tmom1),
                                         O._p(tmom2), O._p(qmom1), O._p(qmom2))


def accumulate_kernel_eval(time_only, event_hit, event_time, event_charge, mc_time, mc_charge, trange, qrange,
                           inv_tbw, inv_qbw, hitcount, time_pdf, charge_pdf):
    pdf_lib().ref_cpu_accumulate_kernel_eval(
        C.c_int(int(time_only)), C.c_int(len(hitcount)), O._p(O._c(event_hit, np.uint32)),
        O._p(O._c(event_time, np.float32)), O._p(O._c(event_charge, np.float32)), O._p(O._c(mc_time, np.float32)),
        O._p(O._c(mc_charge, np.float32)), C.c_float(trange[0]), C.c_float(trange[1]), C.c_float(qrange[0]),
        C.c_float(qrange[1]), O._p(O._c(inv_tbw, np.float32)), O._p(O._c(inv_qbw, np.float32)), O._p(hitcount),
        O._p(time_pdf), O._p(charge_pdf))


def render(oracle, pos, dir, alpha_depth=10, bg_color=0, state=None):
    "Reference render (render.cu:35-181) on the host: (pixels, state) like oracle.Oracle.render."
    pos, dir = O._c(pos, np.float32), O._c(dir, np.float32)
    n = len(pos)
    if state is None:
        state = (np.zeros((n, alpha_depth), np.float32), np.zeros(n, np.uint32),
                 np.zeros((n, alpha_depth, 4), np.float32))
    dx, dxlen, color = state
    pixels = np.empty(n, np.uint32)
    render_lib().ref_cpu_render(C.byref(oracle.g), O._p(oracle.keep['colors']), C.c_uint64(n), O._p(pos), O._p(dir),
                                C.c_uint32(alpha_depth), O._p(pixels), O._p(dx), O._p(dxlen), O._p(color),
                                C.c_uint32(bg_color))
    return pixels, state
