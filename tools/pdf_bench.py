"""Config C5 as BASELINE.md states it (chroma/benchmark.py:99-154 `pdf` shape): 1,000 events x 1e6
photons — isotropic bombs at uniformly random vertices inside r < 10 m, wavelengths U(400,800) nm —
through Simulation.create_pdf into a (channel, 100 time bins, 10 charge bins) PDF with
trange=(-0.5, 999.5), qrange=(-0.5, 9.5); events dealt round-robin to the ranks
(dist.deal_events), one NCCL all-reduce of the 116 MB PDF + hit counts at the end.

  python tools/pdf_bench.py [--events 1000 --photons 1000000 --instanced]                   # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
         --master-port P tools/pdf_bench.py ...

Every event is keyed by its GLOBAL number (source sequence, vertex, and the call epoch of its
propagate / run_daq: Simulation.create_pdf(event_ids=...)), so the merged PDF is the same array for
any number of ranks: `pdf_sha256` of the N = 1, 2, 4, 8 lines must be equal.  Checked in the line:
hitcount[i] == pdf[i].sum() for every channel (test/test_pdf.py:51-53).
Timed between barrier + synchronize pairs, max over ranks; the all-reduce also on its own."""
import argparse
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def event_vertex(gid, radius=10000.0):
    "vertex of global event `gid`: uniform in the ball r < 10 m (mm)"
    rs = np.random.RandomState(100003 + gid)
    v = rs.normal(size=3)
    return (v / np.linalg.norm(v) * radius * rs.uniform() ** (1.0 / 3.0)).astype(np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--events', type=int, default=1000, help='events per PDF, all ranks together')
    ap.add_argument('--photons', type=int, default=1_000_000)
    ap.add_argument('--spiral-step', type=float, default=206.0)
    ap.add_argument('--max-steps', type=int, default=10)
    ap.add_argument('--instanced', action='store_true', help='two-level geometry (no host flattening)')
    ap.add_argument('--out', default=None)
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    from chroma_b200 import demo, dist, event, gpu
    from chroma_b200.loader import build
    from chroma_b200.sim import Simulation

    comm = dist.Communicator(world, rank, local_rank) if world > 1 else None
    t0 = time.time()
    det = demo.detector(spiral_step=args.spiral_step)
    ctx0 = gpu.create_cuda_context(local_rank)          # the BVH builder needs a context
    if not args.instanced:
        det.flatten()
        build(det, fetch_nodes=False)
    sim = Simulation(det, seed=1, cuda_device=local_rank, instanced=args.instanced)
    setup_s = time.time() - t0
    nch = det.num_channels()
    mine = dist.deal_events(args.events, world, rank)

    def sources(ids):
        for g in ids:
            yield event.PhotonSource(args.photons, pos=event_vertex(g), wavelengths=(400.0, 800.0), t0=0.0,
                                     seed=7, sequence=g)

    def one_pdf(ids):
        sim.create_pdf(sources(ids), 100, (-0.5, 999.5), 10, (-0.5, 9.5), event_ids=ids, max_steps=args.max_steps)
        sim.context.synchronize()
        t_red = 0.0
        if comm is not None:
            t1 = time.perf_counter()
            comm.reduce_pdf(sim.gpu_pdf)
            t_red = time.perf_counter() - t1
        return t_red

    one_pdf(mine[:2])                                     # warm-up: pools, NCCL rings
    if comm is not None:
        comm.barrier()
    sim.context.synchronize()
    l0 = sim.context.launch_count()
    t1 = time.perf_counter()
    t_red = one_pdf(mine)
    sim.context.synchronize()
    if comm is not None:
        comm.barrier()
    dt = time.perf_counter() - t1
    launches = sim.context.launch_count() - l0
    if comm is not None:
        dt, t_red = comm.max_float(dt), comm.max_float(t_red)
    if rank == 0:
        hitcount, hist = sim.gpu_pdf.get_pdfs()
        nphot = args.events * args.photons
        line = {
            'metric': 'photons propagated + histogrammed/sec (chroma.benchmark pdf shape, config C5)',
            'value': nphot / dt, 'unit': 'photons/s', 'n_gpus': world, 'scaling': 'strong (one PDF of %d events dealt to the ranks)' % args.events,
            'events': args.events, 'photons_per_event': args.photons, 'events_per_s': args.events / dt,
            's_per_pdf': dt, 'nccl_reduce_ms': t_red * 1e3, 'nccl_reduce_bytes': int(nch * 1000 * 4 + nch * 4),
            'pmts': nch, 'max_steps': args.max_steps, 'geometry': gpu_tree(sim), 'setup_s': setup_s,
            'gpu_launches_rank0': int(launches), 'source': 'device (event.PhotonSource keyed by the global event number)',
            'hitcount_equals_pdf_sum': bool(np.array_equal(hitcount, hist.reshape(nch, -1).sum(axis=1))),
            'pdf_entries': int(hist.sum()), 'pdf_sha256': hashlib.sha256(np.ascontiguousarray(hist).tobytes()).hexdigest(),
            'hitcount_sha256': hashlib.sha256(np.ascontiguousarray(hitcount).tobytes()).hexdigest(),
            'timing': 'wall clock between barrier+synchronize pairs, max over ranks; D2H of the PDF not included',
        }
        s = json.dumps(line)
        print(s)
        if args.out:
            with open(args.out, 'w') as f:
                f.write(s + '\n')
    if comm is not None:
        comm.barrier()
        comm.close()


def gpu_tree(sim):
    t = sim.gpu_geometry.tree()
    return {'kind': t['kind'], 'search_structure_bytes': t['bytes']}


if __name__ == '__main__':
    main()
