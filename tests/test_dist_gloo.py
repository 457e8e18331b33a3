"""CPU tests of the multi-GPU merge logic (chroma_b200/dist.py) with world_size 2 over gloo:
photon sharding, event dealing, channel merge (min / sum / or) and PDF sum."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as td
import torch.multiprocessing as mp

from chroma_b200 import dist


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 10_000_001):
        for world in (1, 2, 3, 8):
            spans = [dist.shard_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def test_deal_events_round_robin():
    seen = sorted(e for r in range(8) for e in dist.deal_events(1000, 8, r))
    assert seen == list(range(1000))
    assert dist.deal_events(10, 4, 1) == [1, 5, 9]


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, nch, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    td.init_process_group('gloo', rank=rank, world_size=world)
    try:
        rs = np.random.RandomState(100 + rank)
        # per-rank accumulators exactly as run_daq leaves them (daq.cu:5-12,73-76): time as the
        # raw bits of a positive float (ordered like the float), charge ints, history words
        t = rs.uniform(1.0, 500.0, nch).astype(np.float32)
        t[rs.uniform(size=nch) < 0.4] = 1e9
        ti = torch.from_numpy(t.view(np.int32).copy())
        qi = torch.from_numpy(rs.randint(0, 1000, nch).astype(np.int32))
        fl = torch.from_numpy(rs.randint(0, 1 << 12, nch).astype(np.int32))
        mine = (ti.clone(), qi.clone(), fl.clone())
        packed = torch.stack([ti, qi, fl])
        dist.merge_channel_tensors(ti, qi, fl)
        dist.merge_channels_packed(packed)               # the one-collective form gives the same
        assert torch.equal(packed[0], ti) and torch.equal(packed[1], qi) and torch.equal(packed[2], fl)
        hc = torch.from_numpy(rs.randint(0, 5, nch).astype(np.int32))
        pdf = torch.from_numpy(rs.randint(0, 3, nch * 6).astype(np.int32))
        mine_pdf = (hc.clone(), pdf.clone())
        dist.merge_pdf_tensors(hc, pdf)
        # PDF evaluation: counts add, nearest tables keep the k smallest of both ranks
        k, nhit = 5, 40
        eh = torch.from_numpy(rs.randint(0, 50, nch).astype(np.int64))
        eb = torch.from_numpy(rs.randint(0, 4, nch).astype(np.int64))
        nn = np.sort(rs.uniform(0, 10, (nhit, k)).astype(np.float32), axis=1)
        nn[rs.uniform(size=nhit) < 0.3, 2:] = 1e9        # partly filled rows
        nn[0] = 1e9                                       # an empty row
        nn_t = torch.from_numpy(nn.copy())
        mine_eval = (eh.clone(), eb.clone(), nn_t.clone())
        dist.merge_pdf_eval_tensors(eh, eb, nn_t)
        comm = dist.Communicator(world, rank, rank)
        mx = comm.max_float(10.0 + rank)
        sm = comm.sum_int(rank + 1)
        comm.barrier()
        torch.save({'mine': mine, 'merged': (ti, qi, fl), 'mine_pdf': mine_pdf, 'pdf': (hc, pdf),
                    'mine_eval': mine_eval, 'eval': (eh, eb, nn_t),
                    'max': mx, 'sum': sm}, os.path.join(out, 'r%d.pt' % rank))
    finally:
        td.destroy_process_group()


def test_channel_and_pdf_merge_world2(tmp_path):
    world, nch = 2, 257
    mp.spawn(_worker, args=(world, _free_port(), nch, str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(os.path.join(str(tmp_path), 'r%d.pt' % k)) for k in range(world)]
    t0, q0, f0 = r[0]['mine']
    t1, q1, f1 = r[1]['mine']
    for k in range(world):
        ti, qi, fl = r[k]['merged']
        # earliest time: min of the floats == min of their bit patterns for positive floats
        ft = np.minimum(t0.numpy().view(np.float32), t1.numpy().view(np.float32))
        assert np.array_equal(ti.numpy().view(np.float32), ft)
        assert torch.equal(qi, q0 + q1)
        assert torch.equal(fl, f0 | f1)
        hc, pdf = r[k]['pdf']
        assert torch.equal(hc, r[0]['mine_pdf'][0] + r[1]['mine_pdf'][0])
        assert torch.equal(pdf, r[0]['mine_pdf'][1] + r[1]['mine_pdf'][1])
        assert r[k]['max'] == 11.0 and r[k]['sum'] == 3
        eh, eb, nn = r[k]['eval']
        assert torch.equal(eh, r[0]['mine_eval'][0] + r[1]['mine_eval'][0])
        assert torch.equal(eb, r[0]['mine_eval'][1] + r[1]['mine_eval'][1])
        both = np.concatenate([r[0]['mine_eval'][2].numpy(), r[1]['mine_eval'][2].numpy()], axis=1)
        assert np.array_equal(nn.numpy(), np.sort(both, axis=1)[:, :5])
        assert (nn.numpy()[0] == np.float32(1e9)).all()


def test_rank_seeds_and_event_epochs():
    "ADVICE r1: ranks that are dealt events must not draw identical streams"
    seeds = [dist.seed_for_rank(1, r) for r in range(8)]
    assert seeds[0] == 1 and len(set(seeds)) == 8 and all(0 <= s < 2 ** 64 for s in seeds)
    assert dist.seed_for_rank(12345, 3) != dist.seed_for_rank(12346, 3)
    # epochs of dealt events: propagate uses 2g, run_daq 2g + 1 -> no two events of any ranks share one
    used = set()
    for rank in range(4):
        for g in dist.deal_events(50, 4, rank):
            e = dist.event_epoch(g)
            assert e not in used and e + 1 not in used
            used.update((e, e + 1))
