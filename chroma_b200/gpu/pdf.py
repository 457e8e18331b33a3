"""GPUPDF and GPUKernelPDF (API of ``chroma/gpu/pdf.py``): per-channel (time, charge)
histograms, the variable-bin PDF evaluation at the observed hit times, and the
kernel-estimated PDF.  All arithmetic on device data is done by libchroma_b200.so
(csrc/engine_daq.cu, csrc/engine_pdfeval.cu); the host parts kept here are the ones the
reference also does in numpy (bandwidth rule, final division)."""
import ctypes as C

import numpy as np

from chroma_b200 import _lib
from chroma_b200.gpu.tools import DeviceArray, current_context


def _release(owner, attr, destroy):
    h = getattr(owner, attr, None)
    if h is not None and owner.ctx.handle is not None:
        _lib.call(destroy, h)
    setattr(owner, attr, None)


class GPUKernelPDF(object):
    """Kernel-estimated PDF values at the observed hits (chroma/gpu/pdf.py:7-175)."""

    def __init__(self):
        self.ctx = current_context()
        self.handle = None

    def setup_moments(self, nchannels, trange, qrange, time_only=True):
        "Allocate the per-channel moment accumulators (pdf.py:13-32)."
        _release(self, 'handle', 'chroma_kernel_pdf_destroy')
        h = C.c_void_p()
        _lib.call('chroma_kernel_pdf_create', self.ctx.handle, int(nchannels), float(trange[0]),
                  float(trange[1]), float(qrange[0]), float(qrange[1]), int(bool(time_only)),
                  C.byref(h))
        self.handle = h
        self.nchannels, self.trange, self.qrange, self.time_only = nchannels, trange, qrange, time_only

    def clear_moments(self):
        _lib.call('chroma_kernel_pdf_clear_moments', self.handle)

    def accumulate_moments(self, gpuchannels, nthreads_per_block=64):
        "Add the latest run_daq() result to the moments (pdf.py:42-59)."
        if len(gpuchannels.t) != self.nchannels:
            raise ValueError('channel arrays do not match the number of channels given to setup_moments()')
        _lib.call('chroma_kernel_pdf_accumulate_moments', self.handle, C.c_void_p(gpuchannels.t.ptr),
                  C.c_void_p(gpuchannels.q.ptr))

    def get_moments(self):
        "(hitcount, tmom1, tmom2, qmom1, qmom2) as host arrays."
        hitcount = np.empty(self.nchannels, np.uint32)
        m = [np.empty(self.nchannels, np.float32) for _ in range(4)]
        _lib.call('chroma_kernel_pdf_get_moments', self.handle, _lib.ptr(hitcount), *[_lib.ptr(a) for a in m])
        return (hitcount,) + tuple(m)

    def compute_bandwidth(self, event_hit, event_time, event_charge, scale_factor=1.0):
        """Bandwidths from the accumulated moments; the rule and its numpy evaluation order are
        the reference's (pdf.py:61-113), including the missing square in the exponent."""
        rho = 1.0
        hitcount, tmom1, tmom2, qmom1, qmom2 = self.get_moments()
        mom0 = np.maximum(hitcount, 1)
        tmean = tmom1 / mom0
        tvar = np.maximum(tmom2 / mom0 - tmean ** 2, 0.0)
        trms = tvar ** 0.5
        d = 1 if self.time_only else 2
        with np.errstate(divide='ignore', invalid='ignore', over='ignore'):
            dimensionality_factor = ((4.0 / (d + 2)) / (mom0 / scale_factor)) ** (-1.0 / (d + 4))
            gaussian_density = np.minimum(1.0 / trms, (1.0 / np.sqrt(2.0 * np.pi)) *
                                          np.exp(-0.5 * ((event_time - tmean) / trms)) / trms)
            time_bandwidths = dimensionality_factor / gaussian_density * rho
            inv_time_bandwidths = np.zeros_like(time_bandwidths)
            pos = time_bandwidths > 0
            inv_time_bandwidths[pos] = time_bandwidths[pos] ** -1
            self.inv_time_bandwidths = inv_time_bandwidths.astype(np.float32)
            if self.time_only:
                self.inv_charge_bandwidths = np.zeros_like(self.inv_time_bandwidths)
            else:
                qmean = qmom1 / mom0
                qrms = (qmom2 / mom0 - qmean ** 2) ** 0.5
                gaussian_density = np.minimum(1.0 / qrms, (1.0 / np.sqrt(2.0 * np.pi)) *
                                              np.exp(-0.5 * ((event_charge - qmean) / qrms)) / qrms)
                charge_bandwidths = dimensionality_factor / gaussian_density * rho
                self.inv_charge_bandwidths = (charge_bandwidths ** -1).astype(np.float32)

    def setup_kernel(self, event_hit, event_time, event_charge):
        "Upload the event of interest and zero the evaluation accumulators (pdf.py:115-134)."
        self.event_hit = np.ascontiguousarray(np.asarray(event_hit).astype(np.uint32))
        et = np.ascontiguousarray(np.asarray(event_time).astype(np.float32))
        eq = np.ascontiguousarray(np.asarray(event_charge).astype(np.float32))
        if not hasattr(self, 'inv_time_bandwidths'):
            raise RuntimeError('compute_bandwidth() must be called before setup_kernel()')
        _lib.call('chroma_kernel_pdf_setup_kernel', self.handle, _lib.ptr(self.event_hit), _lib.ptr(et),
                  _lib.ptr(eq), _lib.ptr(self.inv_time_bandwidths), _lib.ptr(self.inv_charge_bandwidths))

    def clear_kernel(self):
        _lib.call('chroma_kernel_pdf_clear_kernel', self.handle)

    def accumulate_kernel(self, gpuchannels, nthreads_per_block=64):
        "Add the latest run_daq() result to the kernel evaluation (pdf.py:141-160)."
        if len(gpuchannels.t) != self.nchannels:
            raise ValueError('channel arrays do not match the number of channels given to setup_moments()')
        _lib.call('chroma_kernel_pdf_accumulate_kernel', self.handle, C.c_void_p(gpuchannels.t.ptr),
                  C.c_void_p(gpuchannels.q.ptr))

    def get_kernel_sums(self):
        "(hitcount, time sum, charge sum): the raw device accumulators."
        hitcount = np.empty(self.nchannels, np.uint32)
        tp, qp = np.empty(self.nchannels, np.float32), np.empty(self.nchannels, np.float32)
        _lib.call('chroma_kernel_pdf_get_kernel', self.handle, _lib.ptr(hitcount), _lib.ptr(tp), _lib.ptr(qp))
        return hitcount, tp, qp

    def get_kernel_eval(self):
        "(hitcount, pdf value, zeros) per channel (pdf.py:163-175)."
        hitcount, time_pdf_values, charge_pdf_values = self.get_kernel_sums()
        time_pdf_values /= np.maximum(1, hitcount)
        charge_pdf_values /= np.maximum(1, hitcount)
        pdf_values = time_pdf_values if self.time_only else time_pdf_values * charge_pdf_values
        return hitcount, pdf_values, np.zeros_like(pdf_values)

    def __del__(self):
        try:
            _release(self, 'handle', 'chroma_kernel_pdf_destroy')
        except Exception:
            pass


class GPUPDF(object):
    def __init__(self):
        self.ctx = current_context()
        self.handle = None
        self.eval_handle = None

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

    # ---- PDF value at the observed hit of each channel (pdf.py:229-372) -----------------------
    def setup_pdf_eval(self, event_hit, event_time, event_charge, min_twidth, trange, min_qwidth,
                       qrange, min_bin_content=10, time_only=True):
        """The bin around each observed hit time is `min_twidth` wide, or wide enough to hold
        `min_bin_content` Monte Carlo hits (k-nearest-neighbour distance)."""
        assert time_only  # as the reference: only the time observable is supported
        _release(self, 'eval_handle', 'chroma_pdf_eval_destroy')
        event_hit = np.asarray(event_hit)
        self.event_hit = event_hit.astype(bool)
        self.event_nhit = int(np.count_nonzero(event_hit))
        self.map_hit_offset_to_channel_id = np.where(event_hit)[0].astype(np.uint32)
        hit = np.ascontiguousarray(event_hit.astype(np.uint32))
        et = np.ascontiguousarray(np.asarray(event_time).astype(np.float32))
        h = C.c_void_p()
        _lib.call('chroma_pdf_eval_create', self.ctx.handle, len(hit), _lib.ptr(hit), _lib.ptr(et),
                  float(min_twidth), float(trange[0]), float(trange[1]), int(min_bin_content), C.byref(h))
        self.eval_handle = h
        self.eval_nchannels = len(hit)
        self.min_twidth, self.trange, self.min_qwidth, self.qrange = min_twidth, trange, min_qwidth, qrange
        self.min_bin_content, self.time_only = min_bin_content, time_only

    def clear_pdf_eval(self):
        _lib.call('chroma_pdf_eval_clear', self.eval_handle)

    def accumulate_pdf_eval(self, gpuchannels, nthreads_per_block=64, max_blocks=10000):
        "Add the latest run_daq() result (all `ndaq` replicas) to the evaluation; one launch."
        if gpuchannels.stride != self.eval_nchannels:
            raise ValueError('channel arrays do not match the event given to setup_pdf_eval()')
        _lib.call('chroma_pdf_eval_accumulate', self.eval_handle, C.c_void_p(gpuchannels.t.ptr),
                  int(gpuchannels.ndaq))

    def get_pdf_eval_arrays(self):
        "(hitcount, bincount, nearest_mc[nhit, min_bin_content]): the raw device accumulators."
        hitcount = np.empty(self.eval_nchannels, np.uint32)
        bincount = np.empty(self.eval_nchannels, np.uint32)
        nearest = np.empty((self.event_nhit, self.min_bin_content), np.float32)
        nhit = C.c_uint32()
        _lib.call('chroma_pdf_eval_get', self.eval_handle, _lib.ptr(hitcount), _lib.ptr(bincount),
                  _lib.ptr(nearest) if self.event_nhit else None, C.byref(nhit))
        assert nhit.value == self.event_nhit
        return hitcount, bincount, nearest

    def get_pdf_eval(self):
        "(hitcount, pdf value, pdf uncertainty) per channel (pdf.py:331-372)."
        hitcount, bincount, nearest_mc_by_hit = self.get_pdf_eval_arrays()
        return pdf_eval_values(self.event_hit, hitcount, bincount, nearest_mc_by_hit,
                               self.map_hit_offset_to_channel_id, self.min_twidth,
                               self.min_bin_content)

    def _release(self):
        _release(self, 'handle', 'chroma_pdf_destroy')

    def __del__(self):
        try:
            self._release()
            _release(self, 'eval_handle', 'chroma_pdf_eval_destroy')
        except Exception:
            pass


def pdf_eval_values(evhit, hitcount, bincount, nearest_mc_by_hit, hit_channels, min_twidth,
                    min_bin_content):
    """Host tail of get_pdf_eval (pdf.py:336-372): fixed-bin estimate where the minimum bin
    filled up, k-nearest-neighbour estimate elsewhere."""
    pdf_value = np.zeros(len(hitcount), dtype=float)
    pdf_frac_uncert = np.zeros_like(pdf_value)
    high_stats = bincount >= min_bin_content
    if high_stats.any():
        pdf_value[high_stats] = bincount[high_stats].astype(float) / hitcount[high_stats] / min_twidth
        pdf_frac_uncert[high_stats] = 1.0 / np.sqrt(bincount[high_stats])
    low_stats = ~high_stats & (hitcount > 0) & evhit
    nearest_mc = np.full((len(hitcount), min_bin_content), 1e9, dtype=np.float32)
    nearest_mc[hit_channels, :] = nearest_mc_by_hit
    last_valid_entry = np.maximum(0, (nearest_mc < 1e9).astype(int).sum(axis=1) - 1)
    distance = nearest_mc[np.arange(len(last_valid_entry)), last_valid_entry]
    if low_stats.any():
        pdf_value[low_stats] = ((last_valid_entry[low_stats] + 1).astype(float) / hitcount[low_stats] /
                                distance[low_stats] / 2.0)
        pdf_frac_uncert[low_stats] = 1.0 / np.sqrt(last_valid_entry[low_stats] + 1)
    return hitcount, pdf_value, pdf_value * pdf_frac_uncert
