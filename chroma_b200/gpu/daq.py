"""GPUDaq / GPUChannels (API of ``chroma/gpu/daq.py``)."""
import ctypes as C

import numpy as np

from chroma_b200 import _lib, event
from chroma_b200.gpu.tools import DeviceArray


class GPUChannels(object):
    def __init__(self, t, q, flags, ndaq=1, stride=None):
        self.t, self.q, self.flags = t, q, flags
        self.ndaq = ndaq
        self.stride = len(t) if stride is None else stride

    def iterate_copies(self):
        for i in range(self.ndaq):
            w = slice(i * self.stride, (i + 1) * self.stride)
            yield GPUChannels(self.t[w], self.q[w], self.flags[w])

    def get(self):
        t, q = self.t.get(), self.q.get()
        # every channel with a small enough time counts as hit (chroma/gpu/daq.py:32)
        return event.Channels(t < 1e8, t, q, self.flags.get())

    def __len__(self):
        return self.t.size


class GPUDaq(object):
    def __init__(self, gpu_detector, ndaq=1):
        self.ctx = gpu_detector.ctx
        self.gpu_detector = gpu_detector
        h = C.c_void_p()
        _lib.call('chroma_daq_create', gpu_detector.handle, int(ndaq), C.byref(h))
        self.handle = h
        self.ndaq = ndaq
        self.stride = gpu_detector.nchannels
        t, q, f, n = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_uint64()
        _lib.call('chroma_daq_arrays', self.handle, C.byref(t), C.byref(q), C.byref(f), C.byref(n))
        self.earliest_time_gpu = DeviceArray(n.value, np.float32, ctx=self.ctx, ptr=t.value, base=self)
        self.channel_q_gpu = DeviceArray(n.value, np.float32, ctx=self.ctx, ptr=q.value, base=self)
        self.channel_history_gpu = DeviceArray(n.value, np.uint32, ctx=self.ctx, ptr=f.value, base=self)
        ti, qi = C.c_void_p(), C.c_void_p()
        _lib.call('chroma_daq_int_arrays', self.handle, C.byref(ti), C.byref(qi), C.byref(n))
        self.earliest_time_int_gpu = DeviceArray(n.value, np.uint32, ctx=self.ctx, ptr=ti.value, base=self)
        self.channel_q_int_gpu = DeviceArray(n.value, np.uint32, ctx=self.ctx, ptr=qi.value, base=self)

    def begin_acquire(self, nthreads_per_block=64):
        _lib.call('chroma_daq_begin_acquire', self.handle)

    def acquire(self, gpuphotons, rng_states, nthreads_per_block=64, max_blocks=1024,
                start_photon=None, nphotons=None, weight=1.0):
        start_photon = 0 if start_photon is None else start_photon
        if nphotons is None:
            nphotons = len(gpuphotons) - start_photon
        _lib.call('chroma_daq_acquire', self.handle, gpuphotons.handle, rng_states.handle,
                  gpuphotons._offset + start_photon, nphotons, float(weight))

    def end_acquire(self, nthreads_per_block=64):
        # stream-ordered, like the reference's convert_* launches (chroma/gpu/daq.py:93-99): consumers on
        # the context's stream (bin_hits, PDF evaluation) follow in order, .get() synchronises
        _lib.call('chroma_daq_end_acquire', self.handle)
        return GPUChannels(self.earliest_time_gpu, self.channel_q_gpu, self.channel_history_gpu,
                           self.ndaq, self.stride)

    def __del__(self):
        try:
            if self.handle is not None and self.ctx.handle is not None:
                _lib.call('chroma_daq_destroy', self.handle)
                self.handle = None
        except Exception:
            pass
