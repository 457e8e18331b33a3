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
