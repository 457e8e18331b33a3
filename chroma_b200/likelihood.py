"""Likelihood of a detector event under a vertex hypothesis (API of ``chroma/likelihood.py``).

The reference returns ``uncertainties.ufloat`` values; that package is an optional import here:
without it the results are the small ``Value`` pair below (``nominal_value``, ``std_dev``)."""
from itertools import islice
from math import sqrt

import numpy as np

from chroma_b200.tools import profile_if_possible

try:  # pragma: no cover - not installed in the build image
    from uncertainties import ufloat as _ufloat

    def Value(nominal_value, std_dev=0.0):
        return _ufloat(nominal_value, std_dev)
except ImportError:
    class Value(object):
        "A number with a standard deviation; only what Likelihood needs of ufloat."
        def __init__(self, nominal_value, std_dev=0.0):
            self.nominal_value, self._std_dev = float(nominal_value), float(std_dev)

        def std_dev(self):
            return self._std_dev

        def __neg__(self):
            return Value(-self.nominal_value, self._std_dev)

        def __add__(self, other):
            if isinstance(other, Value):
                return Value(self.nominal_value + other.nominal_value,
                             sqrt(self._std_dev ** 2 + other._std_dev ** 2))
            return Value(self.nominal_value + other, self._std_dev)
        __radd__ = __add__

        def __float__(self):
            return self.nominal_value

        def __repr__(self):
            return '%r+/-%r' % (self.nominal_value, self._std_dev)


class Likelihood(object):
    "Evaluates likelihoods for detector events (likelihood.py:7-42)."

    def __init__(self, sim, event=None, tbins=100, trange=(-0.5, 999.5), qbins=10,
                 qrange=(-0.5, 49.5), time_only=True):
        self.sim = sim
        self.tbins, self.trange, self.qbins, self.qrange = tbins, trange, qbins, qrange
        self.time_only = time_only
        if event is not None:
            self.set_event(event)

    def set_event(self, event):
        "Set the detector event being reconstructed."
        self.event = event

    def _pdf_floor(self):
        floor = 1.0 / (self.trange[1] - self.trange[0])
        if not self.time_only:
            floor /= self.qrange[1] - self.qrange[0]
        return floor

    def eval_channel_vbin(self, vertex_generator, nevals, nreps=16, ndaq=50):
        """(hit probability, PDF value, PDF uncertainty) per channel with the variable-bin
        window method (likelihood.py:44-82)."""
        ntotal = nevals * nreps * ndaq
        hitcount, pdf_prob, pdf_prob_uncert = self.sim.eval_pdf(
            self.event.channels, islice(vertex_generator, nevals), 0.2, self.trange, 1, self.qrange,
            nreps=nreps, ndaq=ndaq, time_only=self.time_only, min_bin_content=320)
        hit_prob = hitcount.astype(np.float32) / ntotal
        bad_value = (pdf_prob <= 0.0) | np.isnan(pdf_prob)
        pdf_prob[bad_value] = self._pdf_floor()
        pdf_prob_uncert[bad_value] = self._pdf_floor()
        self.channels_without_data = int((bad_value & self.event.channels.hit).sum())
        return hit_prob, pdf_prob, pdf_prob_uncert

    def eval(self, vertex_generator, nevals, nreps=16, ndaq=50):
        "Negative log likelihood of the event under `vertex_generator` (likelihood.py:84-113)."
        ntotal = nevals * nreps * ndaq
        hit_prob, pdf_prob, _ = self.eval_channel_vbin(vertex_generator, nevals, nreps, ndaq)
        hit = self.event.channels.hit
        hit_prob[~hit] = 1.0 - hit_prob[~hit]
        hit_prob = np.maximum(hit_prob, 0.5 / ntotal)
        log_likelihood = Value(np.log(hit_prob).sum(), 0.0)
        log_likelihood += Value(np.log(pdf_prob[hit]).sum(), 0.0)
        return -log_likelihood

    def setup_kernel(self, vertex_generator, nevals, nreps, ndaq, oversample_factor):
        self.sim.setup_kernel(self.event.channels, islice(vertex_generator, nevals * oversample_factor),
                              self.trange, self.qrange, nreps=nreps, ndaq=ndaq,
                              time_only=self.time_only, scale_factor=oversample_factor)

    @profile_if_possible
    def eval_kernel(self, vertex_generator, nevals, nreps=16, ndaq=50, navg=10):
        """Negative log likelihood from the kernel-estimated PDF, averaged over `navg` passes
        (likelihood.py:127-184; as there, the hit/no-hit term is left out)."""
        hit = self.event.channels.hit
        mom0, mom1, mom2 = 0, 0.0, 0.0
        for _ in range(navg):
            hitcount, pdf_prob, pdf_prob_uncert = self.sim.eval_kernel(
                self.event.channels, islice(vertex_generator, nevals), self.trange, self.qrange,
                nreps=nreps, ndaq=ndaq, time_only=self.time_only)
            bad_value = (pdf_prob <= 0.0) | np.isnan(pdf_prob)
            pdf_prob[bad_value] = self._pdf_floor()
            self.channels_without_data = int((bad_value & hit).sum())
            log_likelihood = float(np.log(pdf_prob[hit]).sum())
            if np.isfinite(log_likelihood):
                mom0 += 1
                mom1 += log_likelihood
                mom2 += log_likelihood ** 2
        avg_like = mom1 / mom0
        rms_like = max(mom2 / mom0 - avg_like ** 2, 0.0) ** 0.5
        return Value(-avg_like, rms_like / sqrt(mom0))
