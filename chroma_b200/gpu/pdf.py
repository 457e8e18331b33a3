"""GPUPDF: per-channel (time, charge) histograms (API of ``chroma/gpu/pdf.py:177-227``).
The variable-bin / kernel-estimation evaluators of the reference (GPUKernelPDF,
setup_pdf_eval ...) are "next" rows of the scope table and not built yet."""
import ctypes as C

import numpy as np

from chroma_b200 import _lib
from chroma_b200.gpu.tools import DeviceArray, current_context


class GPUPDF(object):
    def __init__(self):
        self.ctx = current_context()
        self.handle = None

    def setup_pdf(self, nchannels, tbins, trange, qbins, qrange):
        """nchannels x tbins x qbins uint32 histogram + per-channel hit counts."""
        self._release()
        self.events_in_histogram = 0
        h = C.c_void_p()
        _lib.call('chroma_pdf_create', self.ctx.handle, int(nchannels), int(tbins),
                  float(trange[0]), float(trange[1]), int(qbins), float(qrange[0]),
                  float(qrange[1]), C.byref(h))
        self.handle = h
        self.nchannels, self.tbins, self.trange, self.qbins, self.qrange = \
            nchannels, tbins, trange, qbins, qrange
        hc, pdf, n = C.c_void_p(), C.c_void_p(), C.c_uint64()
        _lib.call('chroma_pdf_arrays', self.handle, C.byref(hc), C.byref(pdf), C.byref(n))
        self.hitcount_gpu = DeviceArray(nchannels, np.uint32, ctx=self.ctx, ptr=hc.value, base=self)
        self.pdf_gpu = DeviceArray((nchannels, tbins, qbins), np.uint32, ctx=self.ctx,
                                   ptr=pdf.value, base=self)

    def clear_pdf(self):
        _lib.call('chroma_pdf_clear', self.handle)
        self.events_in_histogram = 0

    def add_hits_to_pdf(self, gpuchannels, nthreads_per_block=64):
        _lib.call('chroma_pdf_add_device', self.handle, C.c_void_p(gpuchannels.t.ptr),
                  C.c_void_p(gpuchannels.q.ptr))
        self.events_in_histogram += 1

    def get_pdfs(self):
        "(hitcount uint32[nch], pdf uint32[nch, tbins, qbins])"
        hitcount = np.empty(self.nchannels, dtype=np.uint32)
        pdf = np.empty((self.nchannels, self.tbins, self.qbins), dtype=np.uint32)
        _lib.call('chroma_pdf_get', self.handle, _lib.ptr(hitcount), _lib.ptr(pdf))
        return hitcount, pdf

    def _release(self):
        if self.handle is not None and self.ctx.handle is not None:
            _lib.call('chroma_pdf_destroy', self.handle)
        self.handle = None

    def __del__(self):
        try:
            self._release()
        except Exception:
            pass
