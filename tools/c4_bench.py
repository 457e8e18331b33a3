"""Config C4 (BASELINE.json: scintillator + wavelength-shifter reemission, 100 events x 2.5e6 photons
with dichroic surfaces) through Simulation.simulate on one GPU.

The reference ships no such geometry, only the models; the detector is the synthetic one BASELINE.md
describes (chroma_b200/demo/scintillator.py): a two-component scintillator ball in water inside a shell
of 907 PMT modules — 8-inch PMT + light concentrator behind a short-pass dichroic disc and a
wavelength-shifting plate — and a black outer sphere (5.4 M triangles flattened).  Every event is a
bomb at a random vertex of the scintillator, wavelengths U(350,500) nm.  Reports photons/s end to
end through the public API with device sources and with host bombs (H2D inside the timed region),
the propagate launch alone, the history-bit fractions, and the same fractions from the CPU oracle on
a sample (same Philox streams).  --instanced uses the two-level geometry; --fixture the small nested-
sphere fixture of the parity tests (tests/fixtures.py scintillator_detector)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

BITS = ['NO_HIT', 'BULK_ABSORB', 'SURFACE_DETECT', 'SURFACE_ABSORB', 'RAYLEIGH_SCATTER', 'REFLECT_DIFFUSE',
        'REFLECT_SPECULAR', 'SURFACE_REEMIT', 'SURFACE_TRANSMIT', 'BULK_REEMIT']


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--events', type=int, default=100)
    ap.add_argument('--photons', type=int, default=2_500_000)
    ap.add_argument('--max-steps', type=int, default=100)
    ap.add_argument('--oracle-sample', type=int, default=200000)
    ap.add_argument('--out', default=None)
    ap.add_argument('--instanced', action='store_true')
    ap.add_argument('--fixture', action='store_true', help='the small nested-sphere parity fixture instead')
    args = ap.parse_args()
    import fixtures as F
    from chroma_b200 import event, gpu
    from chroma_b200.demo import scintillator
    from chroma_b200.sim import Simulation
    from oracle import oracle as O

    O.set_num_threads(os.cpu_count() or 1)
    det = F.attach_bvh(F.scintillator_detector() if args.fixture else scintillator.detector(), O.build_bvh)
    sim = Simulation(det, seed=3, instanced=args.instanced)
    wl = 380.0 if args.fixture else (350.0, 500.0)
    vertex = (lambda g: (0, 0, 0)) if args.fixture else scintillator.event_vertex
    ctx = sim.context
    n, nev = args.photons, args.events

    def run(sources, keep_first=False):
        flags = np.zeros(len(BITS), np.float64)
        nhit_ch = 0
        first_end = None
        ctx.synchronize()
        l0 = ctx.launch_count()
        t0 = time.perf_counter()
        for ev in sim.simulate(sources, keep_photons_end=keep_first, keep_hits=False, keep_flat_hits=False,
                               run_daq=True, max_steps=args.max_steps, photons_per_batch=4 * n):
            nhit_ch += int(ev.channels.hit.sum())
            if keep_first and ev.photons_end is not None:
                f = ev.photons_end.flags
                flags += [((f >> b) & 1).sum() for b in range(len(BITS))]
                if first_end is None:
                    first_end = ev.photons_end
        ctx.synchronize()
        dt = time.perf_counter() - t0
        return dt, flags, nhit_ch, ctx.launch_count() - l0, first_end

    def device_sources(k, seq0):
        return [event.PhotonSource(n, pos=vertex(seq0 + i), wavelengths=wl, t0=0.0, seed=11, sequence=seq0 + i)
                for i in range(k)]

    # flags of a few events (needs the D2H of photons_end: outside the rate measurement)
    _, fl, _, _, _ = run(device_sources(4, 1000), keep_first=True)
    frac = {name: fl[i] / (4 * n) for i, name in enumerate(BITS)}
    # oracle fractions on a sample with the engine's photons as input
    m = min(args.oracle_sample, n)
    src = event.PhotonSource(m, pos=vertex(5000), wavelengths=wl, seed=11, sequence=5000)
    beg = gpu.GPUPhotons(src).get()
    orc = O.Oracle(det)
    t0 = time.perf_counter()
    end_o = orc.propagate(beg, seed=99, epoch=0, max_steps=args.max_steps)
    t_cpu = time.perf_counter() - t0
    rng = gpu.get_rng_states(1, seed=99)
    gp = gpu.GPUPhotons(beg)
    gp.propagate(sim.gpu_geometry, rng, max_steps=args.max_steps)
    end_g = gp.get()
    agree = float((end_g.flags == end_o.flags).mean())
    frac_o = {name: float(((end_o.flags >> i) & 1).mean()) for i, name in enumerate(BITS)}

    # the propagate launch alone on one event-sized bomb (CUDA events)
    g1 = gpu.GPUPhotons(event.PhotonSource(n, pos=vertex(7), wavelengths=wl, seed=11, sequence=7))
    pristine = gpu.GPUPhotons(event.PhotonSource(n, pos=vertex(7), wavelengths=wl, seed=11, sequence=7))
    from chroma_b200 import _lib
    kms = []
    for _ in range(4):
        _lib.call('chroma_photons_copy', g1.handle, pristine.handle)
        ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin()
        g1.propagate(sim.gpu_geometry, rng, max_steps=args.max_steps)
        kms.append(ctx.timer_end())
    del g1, pristine
    run(device_sources(4, 2000))                                     # warm-up
    dt_dev, _, hits_dev, launches, _ = run(device_sources(nev, 0))
    host = F.isotropic_photons(n, seed=1, wavelength_range=(wl, wl) if args.fixture else wl, pos=vertex(3))
    dt_host, _, hits_host, _, _ = run([host for _ in range(nev)])    # same host arrays each event: H2D every batch
    line = {
        'config': 'C4: %d events x %d photons, scintillator + WLS + dichroic (%s, %d channels), '
                  'Simulation.simulate(run_daq=True, max_steps=%d)'
                  % (nev, n, 'tests/fixtures.scintillator_detector' if args.fixture else 'chroma_b200.demo.scintillator',
                     det.num_channels(), args.max_steps),
        'geometry': sim.gpu_geometry.tree(),
        'propagate_launch_ms': float(np.median(kms[1:])), 'propagate_launch_photons_per_s': n / (float(np.median(kms[1:])) * 1e-3),
        'photons_per_s_device_sources': nev * n / dt_dev, 'seconds_device_sources': dt_dev,
        'photons_per_s_host_bombs': nev * n / dt_host, 'seconds_host_bombs': dt_host,
        'gpu_launches': launches, 'channels_hit_per_event': hits_dev / nev,
        'channels_hit_per_event_host': hits_host / nev,
        'history_fractions_engine': frac, 'history_fractions_oracle_sample': frac_o,
        'oracle_sample': m, 'oracle_flags_identical': agree,
        'cpu_oracle_photons_per_s': m / t_cpu, 'cpu_threads': O.num_threads() if hasattr(O, 'num_threads') else None,
        'triangles': int(len(det.mesh.triangles)),
    }
    s = json.dumps(line)
    print(s)
    if args.out:
        with open(args.out, 'w') as f:
            f.write(s + '\n')


if __name__ == '__main__':
    main()
