"""Simulation driver (API of ``chroma/sim.py``): batching, propagate, hits, per-event DAQ
and PDF accumulation on top of chroma_b200.gpu."""
import itertools
import os
import time

import numpy as np

from chroma_b200 import event, gpu
from chroma_b200.tools import profile_if_possible


def pick_seed():
    "A seed mixed from the current time and the process id."
    return int(time.time()) ^ (os.getpid() << 16) & 2 ** 32 - 1


def _peek(iterable):
    it = iter(iterable)
    first = next(it)
    return first, itertools.chain([first], it)


class Simulation(object):
    def __init__(self, detector, seed=None, cuda_device=None, particle_tracking=False,
                 photon_tracking=False, geant4_processes=0, nthreads_per_block=64,
                 max_blocks=1024, instanced=False):
        """`geant4_processes` is accepted for compatibility; the Geant4 photon generator is
        outside this engine, so events must already carry ``photons_beg``.  `instanced=True` (not in
        the reference) uses the two-level geometry built from the detector's placements: no
        flattening, no BVH build, a few MB instead of tens of GB; statistical tier."""
        self.detector = detector
        self.nthreads_per_block = nthreads_per_block
        self.max_blocks = max_blocks
        self.photon_tracking = photon_tracking
        self.seed = pick_seed() if seed is None else seed
        np.random.seed(self.seed % (2 ** 32))
        self.photon_generator = None

        self.context = gpu.create_cuda_context(cuda_device)
        if hasattr(detector, 'num_channels'):
            self.gpu_geometry = gpu.GPUDetector(detector, instanced=instanced)
            self.gpu_daq = gpu.GPUDaq(self.gpu_geometry)
            self.gpu_pdf = gpu.GPUPDF()
            self.gpu_pdf_kernel = gpu.GPUKernelPDF()
        else:
            self.gpu_geometry = gpu.GPUGeometry(detector, instanced=instanced)
        self.rng_states = gpu.get_rng_states(self.nthreads_per_block * self.max_blocks,
                                             seed=self.seed)
        self.pdf_config = None

    def _simulate_batch(self, batch_events, keep_photons_beg=False, keep_photons_end=False,
                        keep_hits=True, keep_flat_hits=True, run_daq=False, max_steps=100,
                        verbose=False):
        "Propagate one batch of events (evidx = index in the batch) and yield them completed."
        bounds = np.cumsum([0] + [len(ev.photons_beg) for ev in batch_events])
        # One container for the batch; every event is loaded straight into its window: host photons
        # are copied from the event's own arrays (no concatenated copy of the batch on the host, the
        # reference joins first: chroma/sim.py:62-72), an event.PhotonSource is sampled in place.
        # Triangles and weights of incoming photons are not used (chroma/sim.py:70-72).
        gpu_photons = gpu.GPUPhotons.empty(int(bounds[-1]))
        # the event index of a window goes by value: the events' own arrays are not written (an event
        # object that appears twice in a batch, or page-locked arrays still being read, stay intact)
        for i, (ev, start) in enumerate(zip(batch_events, bounds[:-1])):
            # borrow: the events' arrays stay as they are until propagate() below has returned
            gpu_photons.load(int(start), ev.photons_beg, evidx=i, borrow=True)
        if keep_photons_beg:
            for ev, start, end in zip(batch_events, bounds[:-1], bounds[1:]):
                if isinstance(ev.photons_beg, event.PhotonSource):   # read back what was sampled
                    ev.photons_beg = gpu_photons._download(int(start), int(end - start))
        tracking = gpu_photons.propagate(self.gpu_geometry, self.rng_states,
                                         nthreads_per_block=self.nthreads_per_block,
                                         max_blocks=self.max_blocks, max_steps=max_steps,
                                         track=self.photon_tracking)
        has_channels = hasattr(self.detector, 'num_channels')
        if keep_photons_end:
            batch_photons_end = gpu_photons.get()
        if has_channels and (keep_hits or keep_flat_hits):
            batch_hits = gpu_photons.get_flat_hits(self.gpu_geometry)

        for i, (ev, start, end) in enumerate(zip(batch_events, bounds[:-1], bounds[1:])):
            if not keep_photons_beg:
                ev.photons_beg = None
            if self.photon_tracking:
                step_ids, step_photons = tracking
                tracks = [[] for _ in range(end - start)]
                for ids, photons in zip(step_ids, step_photons):
                    mask = np.logical_and(ids >= start, ids < end)
                    if np.count_nonzero(mask) == 0:
                        break
                    selected = photons[mask]
                    for k, pid in enumerate(ids[mask] - start):
                        tracks[pid].append(selected[k])
                ev.photon_tracks = [event.Photons.join(t, concatenate=False) if t else event.Photons()
                                    for t in tracks]
            if keep_photons_end:
                ev.photons_end = batch_photons_end[start:end]
            if has_channels and (keep_hits or keep_flat_hits):
                ev_hits = batch_hits[batch_hits.evidx == i]
                if keep_hits:
                    ev.hits = {int(ch): ev_hits[ev_hits.channel == ch]
                               for ch in np.unique(ev_hits.channel)}
                if keep_flat_hits:
                    ev.flat_hits = ev_hits
            if hasattr(self, 'gpu_daq') and run_daq:
                # the DAQ runs per event (chroma/sim.py:127-137)
                self.gpu_daq.begin_acquire()
                self.gpu_daq.acquire(gpu_photons, self.rng_states, start_photon=int(start),
                                     nphotons=int(end - start),
                                     nthreads_per_block=self.nthreads_per_block,
                                     max_blocks=self.max_blocks)
                ev.channels = self.gpu_daq.end_acquire().get()
            yield ev

    def simulate(self, iterable, keep_photons_beg=False, keep_photons_end=False, keep_hits=True,
                 keep_flat_hits=True, run_daq=False, max_steps=1000, photons_per_batch=1000000,
                 evid_start=0):
        """Generator over completed events.  `iterable`: a Photons object, or an iterable of
        Photons / of Events with ``photons_beg`` set."""
        if isinstance(iterable, (event.Photons, event.PhotonSource)):
            first, iterable = iterable, [iterable]
        else:
            first, iterable = _peek(iterable)
        if isinstance(first, (event.Photons, event.PhotonSource)):
            iterable = (event.Event(photons_beg=x) for x in iterable)
        elif isinstance(first, event.Vertex) or (isinstance(first, event.Event) and
                                                 first.photons_beg is None):
            raise NotImplementedError('generating photons from vertices needs the Geant4 '
                                      'generator, which is not part of chroma_b200')
        kw = dict(keep_photons_beg=keep_photons_beg, keep_photons_end=keep_photons_end,
                  keep_hits=keep_hits, keep_flat_hits=keep_flat_hits, run_daq=run_daq,
                  max_steps=max_steps)
        nphotons, batch, evid = 0, [], evid_start
        for ev in iterable:
            ev.id = evid
            evid += 1
            ev.nphotons = len(ev.photons_beg)
            nphotons += ev.nphotons
            batch.append(ev)
            if nphotons >= photons_per_batch:
                yield from self._simulate_batch(batch, **kw)
                nphotons, batch = 0, []
        if batch:
            yield from self._simulate_batch(batch, **kw)

    def create_pdf(self, iterable, tbins, trange, qbins, qrange, nreps=1, event_ids=None, max_steps=10):
        """(hitcount[nch], pdf[nch, tbins, qbins]) accumulated over the events (sim.py:188-217).

        `event_ids` (not in the reference): the GLOBAL number of every event of `iterable`, for jobs that
        deal the events of one PDF to several GPUs (dist.deal_events).  Each event's random streams are
        then keyed by its number (dist.event_epoch) instead of by how many launches this process has made,
        so the PDF summed over the ranks (Communicator.reduce_pdf) is the same, bin for bin, for any
        number of ranks, and no two ranks draw the same uniforms."""
        first, iterable = _peek(iterable)
        if isinstance(first, (event.Photons, event.PhotonSource)):
            iterable = (event.Event(photons_beg=x) for x in iterable)
        pdf_config = (tbins, trange, qbins, qrange)
        if pdf_config != self.pdf_config:
            self.pdf_config = pdf_config
            self.gpu_pdf.setup_pdf(self.detector.num_channels(), tbins, trange, qbins, qrange)
        else:
            self.gpu_pdf.clear_pdf()
        if nreps > 1:
            iterable = (ev for ev in iterable for _ in range(nreps))
            if event_ids is not None:   # every repetition is an event of its own
                event_ids = (int(g) * nreps + r for g in event_ids for r in range(nreps))
        ids = iter(event_ids) if event_ids is not None else None
        for ev in iterable:
            if ids is not None:
                from chroma_b200.dist import event_epoch
                self.rng_states.epoch = event_epoch(next(ids))
            gpu_photons = gpu.GPUPhotons(ev.photons_beg, borrow=True)
            gpu_photons.propagate(self.gpu_geometry, self.rng_states,
                                  nthreads_per_block=self.nthreads_per_block,
                                  max_blocks=self.max_blocks, max_steps=max_steps)
            self.gpu_daq.begin_acquire()
            self.gpu_daq.acquire(gpu_photons, self.rng_states,
                                 nthreads_per_block=self.nthreads_per_block,
                                 max_blocks=self.max_blocks)
            self.gpu_pdf.add_hits_to_pdf(self.gpu_daq.end_acquire())
        return self.gpu_pdf.get_pdfs()

    def _photon_events(self, iterable):
        first, iterable = _peek(iterable)
        if isinstance(first, (event.Photons, event.PhotonSource)):
            return (event.Event(photons_beg=x) for x in iterable)
        if isinstance(first, event.Event) and first.photons_beg is None:
            raise NotImplementedError('generating photons from vertices needs the Geant4 '
                                      'generator, which is not part of chroma_b200')
        return iterable

    @profile_if_possible
    def eval_pdf(self, event_channels, iterable, min_twidth, trange, min_qwidth, qrange,
                 min_bin_content=100, nreps=1, ndaq=1, nscatter=1, time_only=True):
        """(hitcount, pdf value, pdf uncertainty) per channel at the observed hit times
        (chroma/sim.py:219-283): unscattered photons forced with weights plus `nscatter`
        forced-scatter copies, each DAQ'd ndaq//64 times with 64 replicas."""
        ndaq_per_rep = 64
        ndaq_reps = ndaq // ndaq_per_rep
        gpu_daq = gpu.GPUDaq(self.gpu_geometry, ndaq=ndaq_per_rep)
        self.gpu_pdf.setup_pdf_eval(event_channels.hit, event_channels.t, event_channels.q,
                                    min_twidth, trange, min_qwidth, qrange,
                                    min_bin_content=min_bin_content, time_only=True)
        launch = dict(nthreads_per_block=self.nthreads_per_block, max_blocks=self.max_blocks)
        for ev in self._photon_events(iterable):
            no_scatter = gpu.GPUPhotons(ev.photons_beg, ncopies=nreps)
            scatter = gpu.GPUPhotons(ev.photons_beg, ncopies=nreps * nscatter)
            no_scatter.propagate(self.gpu_geometry, self.rng_states, use_weights=True,
                                 scatter_first=-1, max_steps=10, **launch)
            scatter.propagate(self.gpu_geometry, self.rng_states, use_weights=True,
                              scatter_first=1, max_steps=5, **launch)
            nphotons = no_scatter.true_nphotons
            for i in range(no_scatter.ncopies):
                direct = no_scatter.select(event.SURFACE_DETECT, start_photon=i * nphotons,
                                           nphotons=nphotons)
                scattered = [scatter.select(event.SURFACE_DETECT,
                                            start_photon=(nscatter * i + j) * nphotons,
                                            nphotons=nphotons) for j in range(nscatter)]
                if len(direct) == 0:
                    continue
                for _ in range(ndaq_reps):
                    gpu_daq.begin_acquire()
                    gpu_daq.acquire(direct, self.rng_states, **launch)
                    for s in scattered:
                        gpu_daq.acquire(s, self.rng_states, weight=1.0 / nscatter, **launch)
                    self.gpu_pdf.accumulate_pdf_eval(gpu_daq.end_acquire(),
                                                     nthreads_per_block=ndaq_per_rep)
        return self.gpu_pdf.get_pdf_eval()

    def _daq_each_copy(self, iterable, nreps, ndaq, sink):
        "propagate every event `nreps` times and hand `ndaq` DAQ passes of each copy to `sink`."
        launch = dict(nthreads_per_block=self.nthreads_per_block, max_blocks=self.max_blocks)
        for ev in self._photon_events(iterable):
            gpu_photons = gpu.GPUPhotons(ev.photons_beg, ncopies=nreps)
            gpu_photons.propagate(self.gpu_geometry, self.rng_states, **launch)
            for photon_slice in gpu_photons.iterate_copies():
                for _ in range(ndaq):
                    self.gpu_daq.begin_acquire()
                    self.gpu_daq.acquire(photon_slice, self.rng_states, **launch)
                    sink(self.gpu_daq.end_acquire())

    def setup_kernel(self, event_channels, bandwidth_iterable, trange, qrange, nreps=1, ndaq=1,
                     time_only=True, scale_factor=1.0):
        "Accumulate moments and pick the kernel bandwidths (chroma/sim.py:285-315)."
        self.gpu_pdf_kernel.setup_moments(len(event_channels.hit), trange, qrange,
                                          time_only=time_only)
        self._daq_each_copy(bandwidth_iterable, nreps, ndaq, self.gpu_pdf_kernel.accumulate_moments)
        self.gpu_pdf_kernel.compute_bandwidth(event_channels.hit, event_channels.t,
                                              event_channels.q, scale_factor=scale_factor)

    def eval_kernel(self, event_channels, kernel_iterable, trange, qrange, nreps=1, ndaq=1,
                    naverage=1, time_only=True):
        "(hitcount, pdf value, zeros) from the kernel estimate (chroma/sim.py:317-343)."
        self.gpu_pdf_kernel.setup_kernel(event_channels.hit, event_channels.t, event_channels.q)
        self._daq_each_copy(kernel_iterable, nreps, ndaq, self.gpu_pdf_kernel.accumulate_kernel)
        return self.gpu_pdf_kernel.get_kernel_eval()

    def __del__(self):
        try:
            for name in ('gpu_daq', 'gpu_pdf', 'gpu_pdf_kernel', 'gpu_geometry', 'rng_states'):
                if hasattr(self, name):
                    delattr(self, name)
            self.context.pop()
        except Exception:
            pass
