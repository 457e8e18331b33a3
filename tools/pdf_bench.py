"""Config C5 (BASELINE.json: `chroma.benchmark pdf`, 1e9 photons sharded over 8 B200s, per-channel
time/charge PDF accumulation with an NCCL reduce) on N GPUs of one box.

  python tools/pdf_bench.py                                    # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
         --master-port P tools/pdf_bench.py

Every rank owns `--events-per-gpu` whole events (weak scaling, events dealt by rank); an event is a
`--photons` bomb sampled on the device (event.PhotonSource), propagated, read out by run_daq and
binned by bin_hits into the rank's (channel, 100 time bins, 10 charge bins) histogram; once per PDF
the histograms are summed with one NCCL all-reduce (116 MB at 29,023 channels) and read back on
rank 0.  Timed with a barrier + synchronize on both sides, max over ranks; the all-reduce is also
timed on its own.  Prints one JSON line on rank 0."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--events-per-gpu', type=int, default=50)
    ap.add_argument('--photons', type=int, default=2_500_000)
    ap.add_argument('--npdfs', type=int, default=3)
    ap.add_argument('--events-per-batch', type=int, default=4)
    ap.add_argument('--spiral-step', type=float, default=206.0)
    ap.add_argument('--max-steps', type=int, default=10)
    ap.add_argument('--out', default=None)
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    from chroma_b200 import demo, dist, event, gpu
    from chroma_b200.loader import build

    if world > 1:
        dist.bind_to_gpu_numa(local_rank)
    comm = dist.Communicator(world, rank, local_rank) if world > 1 else None
    ctx = gpu.create_cuda_context(local_rank)
    det = demo.detector(spiral_step=args.spiral_step)
    det.flatten()
    build(det, fetch_nodes=False)
    gg = gpu.GPUDetector(det)
    nch, ntri = det.num_channels(), len(det.mesh.triangles)
    gg.geometry = None
    det = None
    rng = gpu.get_rng_states(1, seed=1)
    daq = gpu.GPUDaq(gg)
    pdf = gpu.GPUPDF()
    pdf.setup_pdf(nch, 100, (-0.5, 999.5), 10, (-0.5, 9.5))
    nev = args.events_per_gpu
    launches = []

    def one_pdf(ipdf):
        pdf.clear_pdf()
        nb = max(1, args.events_per_batch)
        for k0 in range(0, nev, nb):
            # a batch of whole events in one container, as Simulation.simulate batches them
            # (chroma/sim.py:139-186): one propagate launch, then DAQ + binning per event window
            m = min(nb, nev - k0)
            gp = gpu.GPUPhotons.empty(m * args.photons)
            for j in range(m):
                gid = (ipdf * world + rank) * nev + k0 + j     # global event number: keys the source
                src = event.PhotonSource(args.photons, wavelengths=(400.0, 800.0), t0=100.0, seed=7, sequence=gid)
                src.evidx = j
                gp.load(j * args.photons, src)
            gp.propagate(gg, rng, max_steps=args.max_steps)
            for j in range(m):
                daq.begin_acquire()
                daq.acquire(gp, rng, start_photon=j * args.photons, nphotons=args.photons)
                pdf.add_hits_to_pdf(daq.end_acquire())
        ctx.synchronize()
        t_red = 0.0
        if comm is not None:
            t0 = time.perf_counter()
            comm.reduce_pdf(pdf)
            t_red = time.perf_counter() - t0
        hitcount, hist = pdf.get_pdfs() if rank == 0 else (None, None)
        return t_red, hitcount, hist

    one_pdf(0)                                             # warm-up (allocation pools, NCCL rings)
    total_ms, red_ms = [], []
    hitcount = hist = None
    for i in range(args.npdfs):
        if comm is not None:
            comm.barrier()
        ctx.synchronize()
        l0 = ctx.launch_count()
        t0 = time.perf_counter()
        t_red, hitcount, hist = one_pdf(1 + i)
        ctx.synchronize()
        if comm is not None:
            comm.barrier()
        dt = (time.perf_counter() - t0) * 1e3
        launches.append(ctx.launch_count() - l0)
        total_ms.append(comm.max_float(dt) if comm is not None else dt)
        red_ms.append(comm.max_float(t_red * 1e3) if comm is not None else 0.0)
    if rank == 0:
        nphot = world * nev * args.photons
        ms = float(np.mean(total_ms))
        line = {
            'metric': 'photons propagated + histogrammed/sec (chroma.benchmark pdf shape, config C5)',
            'value': nphot / (ms * 1e-3), 'unit': 'photons/s', 'n_gpus': world, 'scaling': 'weak',
            'events_per_s': world * nev / (ms * 1e-3), 'ms_per_pdf': ms, 'ms_per_pdf_all': [round(x, 2) for x in total_ms],
            'nccl_reduce_ms': float(np.mean(red_ms)), 'nccl_reduce_bytes': int(nch * 1000 * 4 + nch * 4),
            'photons_per_pdf': nphot, 'events_per_gpu': nev, 'photons_per_event': args.photons, 'events_per_batch': args.events_per_batch,
            'pmts': nch, 'triangles': int(ntri), 'max_steps': args.max_steps, 'gpu_launches_per_pdf': launches[-1],
            'source': 'device (event.PhotonSource): no H2D of photons',
            'pdf_entries': int(hist.sum()), 'hit_events': int(hitcount.sum()),
            'timing': 'wall clock between barrier+synchronize pairs, max over ranks, includes the D2H of the PDF on rank 0',
        }
        s = json.dumps(line)
        print(s)
        if args.out:
            with open(args.out, 'w') as f:
                f.write(s + '\n')
    if comm is not None:
        comm.barrier()
        comm.close()
    ctx.pop()


if __name__ == '__main__':
    main()
