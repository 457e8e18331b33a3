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
