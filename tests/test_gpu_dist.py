"""Config C5's check on one GPU box: the PDF of a set of events dealt to several ranks
(dist.deal_events + Simulation.create_pdf(event_ids=...), merged by an all-reduce) is the SAME array,
bin for bin, as the PDF one process accumulates over all the events — every event's random streams are
keyed by its global number (dist.event_epoch), not by which rank ran it or what the rank ran before.
Two processes share cuda:0 here and merge over gloo; on N GPUs tools/pdf_bench.py does the same with
NCCL (its `pdf_sha256` must agree for N = 1, 2, 4, 8)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NEVENTS, NPHOTONS = 6, 20000
PDF_SHAPE = (100, (-0.5, 999.5), 10, (-0.5, 999.5))     # charge = photons seen by the channel


def _sources(ids):
    from chroma_b200 import event
    rs = np.random.RandomState(5)
    vertices = rs.uniform(-800, 800, size=(NEVENTS, 3)).astype(np.float32)
    return [event.PhotonSource(NPHOTONS, pos=vertices[g], wavelengths=(400.0, 800.0), t0=0.0, seed=7, sequence=g)
            for g in ids]


def _pdf_of(ids, instanced=True):
    import fixtures as F
    from chroma_b200.sim import Simulation
    det = F.small_detector()
    sim = Simulation(det, seed=1, cuda_device=0, instanced=instanced)
    hitcount, pdf = sim.create_pdf(_sources(ids), *PDF_SHAPE, event_ids=ids, max_steps=10)
    assert np.array_equal(hitcount, pdf.reshape(len(hitcount), -1).sum(axis=1))     # test/test_pdf.py:51-53
    out = hitcount.copy(), pdf.copy()
    sim.context.pop()
    return out


def _worker(rank, world, port, out):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import torch
    import torch.distributed as td
    from chroma_b200 import dist
    os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
    td.init_process_group('gloo', rank=rank, world_size=world)
    try:
        hitcount, pdf = _pdf_of(dist.deal_events(NEVENTS, world, rank))
        h, p = torch.from_numpy(hitcount.astype(np.int32)), torch.from_numpy(pdf.astype(np.int32))
        dist.merge_pdf_tensors(h, p)
        if rank == 0:
            np.savez(os.path.join(out, 'merged.npz'), hitcount=h.numpy(), pdf=p.numpy())
    finally:
        td.destroy_process_group()


def test_pdf_of_dealt_events_equals_single_process(tmp_path):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    merged = np.load(os.path.join(str(tmp_path), 'merged.npz'))
    hitcount, pdf = _pdf_of(list(range(NEVENTS)))
    assert pdf.sum() > 100                     # ~32 channels x 6 events, each hit channel counts once per event
    assert np.array_equal(merged['hitcount'], hitcount.astype(np.int32))
    assert np.array_equal(merged['pdf'], pdf.astype(np.int32))
    # a different order of the same events gives the same PDF too
    h2, p2 = _pdf_of([5, 0, 3, 1, 4, 2])
    assert np.array_equal(p2, pdf)


def _channels(ph, start, stop, comm):
    "run_daq channels of photons [start, stop) of ONE event (id_base = start); merged over the ranks when comm is given"
    import fixtures as F
    from chroma_b200 import gpu
    det = F.attach_bvh(F.small_detector(), __import__('oracle.oracle', fromlist=['x']).build_bvh)
    ctx = gpu.create_cuda_context(0)
    try:
        out = []
        for instanced in (False, True):
            gg = gpu.GPUDetector(det, instanced=instanced)
            gp = gpu.GPUPhotons(ph[start:stop], id_base=start)
            rng = gpu.get_rng_states(1, seed=5)
            gp.propagate(gg, rng, max_steps=10)
            rng.epoch = 1000                    # the event's run_daq epoch, the same on every rank
            d = gpu.GPUDaq(gg)
            d.begin_acquire(); d.acquire(gp, rng)
            if comm is not None:
                comm.reduce_channels(d)         # chroma_daq_pack + all-gather + chroma_daq_merge_gathered
            ch = d.end_acquire().get()
            out.append((ch.hit.copy(), ch.flags.copy(), ch.t.view(np.uint32).copy(), ch.q.view(np.uint32).copy()))
        return out
    finally:
        ctx.pop()


def _event_photons():
    import fixtures as F
    return F.isotropic_photons(400000, seed=41, wavelength_range=(400, 800))


def _merge_worker(rank, world, port, out):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from oracle import oracle
    oracle.load()
    from chroma_b200 import dist
    os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
    comm = dist.Communicator(world, rank, 0, backend='gloo')      # both ranks share cuda:0 here
    try:
        ph = _event_photons()
        start, stop = dist.shard_range(len(ph), world, rank)
        res = _channels(ph, start, stop, comm)
        if rank == 0:
            np.savez(os.path.join(out, 'merged_channels.npz'), **{'%d_%d' % (g, k): a for g, r in enumerate(res) for k, a in enumerate(r)})
    finally:
        comm.close()


def test_channels_of_one_event_over_two_ranks_equal_one_process(tmp_path, oracle_mod):
    """SURVEY 8(e) "one giant event": two ranks propagate halves of one bomb (streams keyed by the global photon
    id), run run_daq on their halves and merge the channel accumulators (Communicator.reduce_channels); the
    merged channels equal, bit for bit, the ones a single process gets from the whole bomb, on both geometries."""
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_merge_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    merged = np.load(os.path.join(str(tmp_path), 'merged_channels.npz'))
    ph = _event_photons()
    whole = _channels(ph, 0, len(ph), None)
    for g, r in enumerate(whole):
        assert r[0].sum() > 10
        for k, a in enumerate(r):
            assert np.array_equal(merged['%d_%d' % (g, k)], a), (g, k)
