"""Measure ONCE, with the CPU oracle's counters (chroma/cuda/mesh.h:65-66,75,84 restated in
oracle/chroma_oracle.cpp), the visit statistics of the REFERENCE traversal that define the
algorithmic bytes of SURVEY.md 8(d) for every BASELINE.json config, and freeze them in
profiles/frozen_visit_stats.json (copied into BASELINE.md section 5).  bench.py and the tools read
the frozen file; nothing is re-measured at bench time.

  python tools/freeze_visit_stats.py            (CPU only: ~15 GB of host memory, a few minutes)

B_ray    = 24 + 8 + 16 N_node + 48 N_tri
B_photon = 120 + sum_steps (16 N_node + 48 N_tri + 4 + 48) + 24 (DAQ)"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
OUT = os.path.join(ROOT, 'profiles', 'frozen_visit_stats.json')


def b_ray(nodes, tris):
    return 24.0 + 8.0 + 16.0 * nodes + 48.0 * tris


def b_photon(steps, nodes, tris):
    "nodes / tris are per PHOTON (summed over its steps)"
    return 120.0 + 16.0 * nodes + 48.0 * tris + 52.0 * steps + 24.0


def ray_stats(orc, pos, d):
    _, _, cnt = orc.intersect(pos, d, counts=True)
    nodes, tris = float(cnt[:, 0].mean()), float(cnt[:, 1].mean())
    return dict(rays=len(pos), nodes=nodes, tris=tris, bytes_per_ray=b_ray(nodes, tris))


def photon_stats(orc, ph, max_steps, seed=1):
    _, st = orc.propagate(ph, seed=seed, max_steps=max_steps, stats=True)
    steps, nodes, tris = [float(st[:, k].mean()) for k in range(3)]
    return dict(photons=len(ph), max_steps=max_steps, steps=steps, nodes=nodes, tris=tris,
                bytes_per_photon=b_photon(steps, nodes, tris))


def main():
    import fixtures as F
    import bench
    from chroma_b200 import demo, event
    from chroma_b200.demo import scintillator
    from chroma_b200.tools import argsort_direction
    from oracle import oracle as O
    out = {'how': 'oracle/chroma_oracle.cpp counters of the reference traversal (mesh.h:65-66,75,84), '
                  'tools/freeze_visit_stats.py', 'configs': {}}
    C = out['configs']
    t0 = time.time()
    # C1: water sphere, 1e5 photons, 400 nm, max_steps=100
    g = F.attach_bvh(F.water_sphere(), O.build_bvh)
    C['C1'] = photon_stats(O.Oracle(g), F.isotropic_photons(100000, seed=0, wavelength=400.0), 100)
    print('C1', C['C1'], time.time() - t0, flush=True)
    # C4: scintillator + WLS + dichroic, bombs at random vertices in the scintillator, U(350,500) nm
    det = F.attach_bvh(scintillator.detector(), O.build_bvh)
    orc = O.Oracle(det)
    parts = [F.isotropic_photons(10000, seed=40 + k, wavelength_range=(350, 500), pos=scintillator.event_vertex(k))
             for k in range(4)]
    C['C4'] = photon_stats(orc, event.Photons.join(parts), 100)
    C['C4'].update(pmts=det.num_channels(), triangles=int(len(det.mesh.triangles)))
    print('C4', C['C4'], time.time() - t0, flush=True)
    del det, orc
    for name, step in (('10055', 350.0), ('29023', 206.0)):
        det = demo.detector(spiral_step=step)
        det = F.attach_bvh(det, O.build_bvh)
        orc = O.Oracle(det)
        print(name, 'built', time.time() - t0, flush=True)
        n = 100000
        pos, d = F.isotropic_rays(n, seed=0)
        d = d[argsort_direction(d)]
        C['C2_%s_pmts' % name] = dict(ray_stats(orc, pos, d), pmts=det.num_channels(),
                                      triangles=int(len(det.mesh.triangles)),
                                      rays_from='origin, isotropic, direction-sorted (chroma/benchmark.py:22-46)')
        print('C2', name, C['C2_%s_pmts' % name], time.time() - t0, flush=True)
        if name == '29023':
            C['C3'] = dict(photon_stats(orc, bench.make_bomb(40000, seed=1234), 10), pmts=det.num_channels(),
                           triangles=int(len(det.mesh.triangles)))
            print('C3', C['C3'], time.time() - t0, flush=True)
            sys.path.insert(0, os.path.join(ROOT, 'tools'))
            import pdf_bench
            parts = [F.isotropic_photons(5000, seed=50 + k, wavelength_range=(400, 800), pos=pdf_bench.event_vertex(k))
                     for k in range(8)]
            C['C5'] = dict(photon_stats(orc, event.Photons.join(parts), 10), pmts=det.num_channels(),
                           vertices='uniform in r < 10 m (tools/pdf_bench.py event_vertex)')
            print('C5', C['C5'], time.time() - t0, flush=True)
        del det, orc
    with open(OUT, 'w') as f:
        json.dump(out, f, indent=1)
    print('wrote', OUT)


if __name__ == '__main__':
    main()
