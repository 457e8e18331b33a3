#!/usr/bin/env python
"""Micro-benchmarks with the names, arguments and loop structure of ``chroma/benchmark.py``:
``intersect`` (:22-46), ``load_photons`` (:48-66), ``propagate`` (:68-95), ``pdf`` (:97-153) and
``pdf_eval`` (:155-242), plus ``generate_photons`` for the device photon source.

Differences from the reference, all forced by what exists on the box:
  * rates are returned as ``(mean, std)`` tuples (the `uncertainties` package is not installed);
  * the 100 MeV electron events of ``pdf`` / ``pdf_eval`` come from Geant4 in the reference
    (``G4ParallelGenerator``); here ``events`` defaults to photon bombs of `photons_per_event`
    photons sampled on the GPU (event.PhotonSource), and any iterable of Events / Photons with
    the same meaning may be passed instead;
  * timing is wall clock around a context synchronize, as in the reference; the first pass is
    discarded as it is there.

Run as ``python -m chroma_b200.benchmark [ray load generate propagate pdf pdf_eval]``.
"""
import itertools
import sys
import time

import numpy as np

from chroma_b200 import event, gpu, sample, tools
from chroma_b200.transform import normalize

PHOTONS_PER_EVENT = 25000   # stand-in for the Cherenkov yield of a 100 MeV electron in water


def _rate(count, run_times):
    "count / time as (mean, std), first-order propagation of the spread of the run times"
    t = np.asarray(run_times, dtype=np.float64)
    if len(t) == 0:
        return (float('nan'), float('nan'))
    mean = count / t.mean()
    return (float(mean), float(mean * t.std() / t.mean()))


def rate_to_str(rate):
    return '%.4g +/- %.2g' % rate


def _sync():
    gpu.current_context().synchronize()


def bomb_events(nevents, photons_per_event=PHOTONS_PER_EVENT, seed=0, first=0):
    "events started from isotropic photon bombs at the centre, sampled on the device"
    for i in range(first, first + nevents):
        yield event.Event(photons_beg=event.PhotonSource(photons_per_event, wavelengths=(400.0, 800.0),
                                                         seed=seed, sequence=i))


def intersect(gpu_geometry, number=100, nphotons=500000, nthreads_per_block=64, max_blocks=1024):
    "Returns the average number of ray intersections per second (chroma/benchmark.py:22-46)."
    run_times = []
    for i in range(number):
        pos = gpu.zeros(nphotons * 3, np.float32)
        dir = sample.uniform_sphere(nphotons)
        reorder = tools.argsort_direction(dir)
        dir = gpu.to_gpu(np.ascontiguousarray(dir[reorder], dtype=np.float32).ravel())
        _sync()
        t0 = time.time()
        gpu.intersect(gpu_geometry, pos, dir)
        _sync()
        elapsed = time.time() - t0
        if i > 0:
            run_times.append(elapsed)
    return _rate(nphotons, run_times)


def load_photons(number=100, nphotons=500000):
    "Returns the average number of photons moved to the GPU per second (chroma/benchmark.py:48-66)."
    pos = np.zeros((nphotons, 3))
    dir = sample.uniform_sphere(nphotons)
    pol = normalize(np.cross(sample.uniform_sphere(nphotons), dir))
    wavelengths = np.random.uniform(400, 800, size=nphotons)
    photons = event.Photons(pos, dir, pol, wavelengths)
    run_times = []
    for i in range(number):
        t0 = time.time()
        gpu_photons = gpu.GPUPhotons(photons)
        _sync()
        elapsed = time.time() - t0
        del gpu_photons
        if i > 0:
            run_times.append(elapsed)
    return _rate(nphotons, run_times)


def generate_photons(number=100, nphotons=500000):
    """Photons per second of the same bomb sampled on the device (no reference counterpart: it
    replaces the host sampling + ``load_photons`` pair)."""
    run_times = []
    for i in range(number):
        src = event.PhotonSource(nphotons, wavelengths=(400.0, 800.0), seed=1, sequence=i)
        _sync()
        t0 = time.time()
        gpu_photons = gpu.GPUPhotons(src)
        _sync()
        elapsed = time.time() - t0
        del gpu_photons
        if i > 0:
            run_times.append(elapsed)
    return _rate(nphotons, run_times)


def propagate(gpu_detector, number=10, nphotons=500000, nthreads_per_block=64, max_blocks=1024):
    "Returns the average number of photons propagated per second (chroma/benchmark.py:68-95)."
    rng_states = gpu.get_rng_states(nthreads_per_block * max_blocks)
    run_times = []
    for i in range(number):
        pos = np.zeros((nphotons, 3))
        dir = sample.uniform_sphere(nphotons)
        reorder = tools.argsort_direction(dir)
        dir = dir[reorder]
        pol = normalize(np.cross(sample.uniform_sphere(nphotons), dir))
        wavelengths = np.random.uniform(400, 800, size=nphotons)
        photons = event.Photons(pos, dir, pol, wavelengths)
        gpu_photons = gpu.GPUPhotons(photons)
        _sync()
        t0 = time.time()
        gpu_photons.propagate(gpu_detector, rng_states, nthreads_per_block, max_blocks)
        _sync()
        elapsed = time.time() - t0
        if i > 0:
            run_times.append(elapsed)
    return _rate(nphotons, run_times)


@tools.profile_if_possible
def pdf(gpu_detector, npdfs=10, nevents=100, nreps=16, ndaq=1, nthreads_per_block=64, max_blocks=1024,
        events=None, communicator=None):
    """Returns the average number of events histogrammed per second (chroma/benchmark.py:97-153):
    each event is propagated `nreps` times, each copy read out `ndaq` times into the
    (channel, time, charge) histogram.  `events(nevents)` yields the events of one PDF.  With a
    `communicator` (chroma_b200.dist.Communicator) the events of a PDF are dealt over the ranks
    and the histograms summed over NCCL before they are read back."""
    rng_states = gpu.get_rng_states(nthreads_per_block * max_blocks)
    gpu_daq = gpu.GPUDaq(gpu_detector)
    gpu_pdf = gpu.GPUPDF()
    gpu_pdf.setup_pdf(gpu_detector.nchannels, 100, (-0.5, 999.5), 10, (-0.5, 9.5))
    events = bomb_events if events is None else events
    run_times = []
    for i in range(npdfs):
        t0 = time.time()
        gpu_pdf.clear_pdf()
        evs = events(nevents)
        if communicator is not None:
            evs = itertools.islice(evs, communicator.rank, None, communicator.world)
        for ev in evs:
            gpu_photons = gpu.GPUPhotons(ev.photons_beg, ncopies=nreps)
            gpu_photons.propagate(gpu_detector, rng_states, nthreads_per_block, max_blocks)
            for gpu_photon_slice in gpu_photons.iterate_copies():
                for idaq in range(ndaq):
                    gpu_daq.begin_acquire()
                    gpu_daq.acquire(gpu_photon_slice, rng_states, nthreads_per_block, max_blocks)
                    gpu_channels = gpu_daq.end_acquire()
                    gpu_pdf.add_hits_to_pdf(gpu_channels, nthreads_per_block)
        if communicator is not None:
            communicator.reduce_pdf(gpu_pdf)
        hitcount, pdf = gpu_pdf.get_pdfs()
        elapsed = time.time() - t0
        if i > 0:
            run_times.append(elapsed)
    return _rate(nevents * nreps * ndaq, run_times)


@tools.profile_if_possible
def pdf_eval(gpu_detector, npdfs=10, nevents=25, nreps=16, ndaq=128, nthreads_per_block=64,
             max_blocks=1024, events=None):
    """Returns the average number of events per second accumulated in the PDF-evaluation data
    structure (chroma/benchmark.py:155-242)."""
    rng_states = gpu.get_rng_states(nthreads_per_block * max_blocks)
    events = bomb_events if events is None else events

    # Make data event
    data_ev = next(iter(events(1)))
    gpu_photons = gpu.GPUPhotons(data_ev.photons_beg)
    gpu_photons.propagate(gpu_detector, rng_states, nthreads_per_block, max_blocks)
    gpu_daq = gpu.GPUDaq(gpu_detector)
    gpu_daq.begin_acquire()
    gpu_daq.acquire(gpu_photons, rng_states, nthreads_per_block, max_blocks)
    data_ev_channels = gpu_daq.end_acquire().get()

    # Setup PDF evaluation
    gpu_daq = gpu.GPUDaq(gpu_detector, ndaq=ndaq)
    gpu_pdf = gpu.GPUPDF()
    gpu_pdf.setup_pdf_eval(data_ev_channels.hit, data_ev_channels.t, data_ev_channels.q, 0.05,
                           (-0.5, 999.5), 1.0, (-0.5, 20), min_bin_content=20, time_only=True)
    run_times = []
    for i in range(npdfs):
        t0 = time.time()
        gpu_pdf.clear_pdf_eval()
        for ev in events(nevents):
            gpu_photons = gpu.GPUPhotons(ev.photons_beg, ncopies=nreps)
            gpu_photons.propagate(gpu_detector, rng_states, nthreads_per_block, max_blocks)
            for gpu_photon_slice in gpu_photons.iterate_copies():
                gpu_daq.begin_acquire()
                gpu_photon_slice = gpu_photon_slice.select(event.SURFACE_DETECT)
                gpu_daq.acquire(gpu_photon_slice, rng_states, nthreads_per_block, max_blocks)
                gpu_channels = gpu_daq.end_acquire()
                gpu_pdf.accumulate_pdf_eval(gpu_channels, nthreads_per_block)
        _sync()
        elapsed = time.time() - t0
        if i > 0:
            run_times.append(elapsed)
    return _rate(nevents * nreps * ndaq, run_times)


def main(argv=None):
    from chroma_b200 import demo
    from chroma_b200.loader import create_geometry_from_obj

    argv = sys.argv[1:] if argv is None else argv
    tests = argv if argv else ['ray', 'load', 'generate', 'propagate', 'pdf', 'pdf_eval']
    context = gpu.create_cuda_context()
    detector = create_geometry_from_obj(demo.detector())
    gpu_detector = gpu.GPUDetector(detector)
    if 'ray' in tests:
        print('%s ray intersections/sec.' % rate_to_str(intersect(gpu_detector)))
    if 'load' in tests:
        print('%s photons loaded/sec.' % rate_to_str(load_photons()))
    if 'generate' in tests:
        print('%s photons sampled on the device/sec.' % rate_to_str(generate_photons()))
    if 'propagate' in tests:
        print('%s photons propagated/sec.' % rate_to_str(propagate(gpu_detector)))
    if 'pdf' in tests:
        print('%s events histogrammed/s' % rate_to_str(pdf(gpu_detector)))
    if 'pdf_eval' in tests:
        print('%s events/s accumulated in PDF evaluation data structure (25 events x 16 Chroma x 128 DAQ)'
              % rate_to_str(pdf_eval(gpu_detector)))
    context.pop()


if __name__ == '__main__':
    main()
