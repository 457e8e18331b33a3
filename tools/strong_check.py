"""One event over N GPUs, checked: every rank propagates its contiguous share of ONE bomb (id_base = start of the
share), runs run_daq on it, the ranks merge their channel accumulators (Communicator.reduce_channels:
chroma_daq_pack + one all-gather + chroma_daq_merge_gathered) and rank 0 compares the merged channels with the
ones it gets from the whole bomb alone: hit pattern, history words, time and charge bits must be equal.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/strong_check.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))


def main():
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    import fixtures as F
    from chroma_b200 import dist, gpu
    from chroma_b200.loader import build
    comm = dist.Communicator(world, rank, local_rank) if world > 1 else None
    ctx = gpu.create_cuda_context(local_rank)
    det = F.small_detector()
    det.flatten()
    build(det, fetch_nodes=False)
    n = 2_000_000
    ph = F.isotropic_photons(n, seed=41, wavelength_range=(400, 800))
    results = {}
    for instanced in (False, True):
        gg = gpu.GPUDetector(det, instanced=instanced)

        def channels(start, stop, merge):
            gp = gpu.GPUPhotons(ph[start:stop], id_base=start)
            rng = gpu.get_rng_states(1, seed=5)
            gp.propagate(gg, rng, max_steps=10)
            rng.epoch = 1000
            d = gpu.GPUDaq(gg)
            d.begin_acquire(); d.acquire(gp, rng)
            if merge:
                comm.reduce_channels(d)
            return d.end_acquire().get()
        start, stop = dist.shard_range(n, world, rank)
        merged = channels(start, stop, comm is not None)
        if rank == 0:
            whole = channels(0, n, False)
            same = (np.array_equal(merged.hit, whole.hit) and np.array_equal(merged.flags, whole.flags) and
                    np.array_equal(merged.t.view(np.uint32), whole.t.view(np.uint32)) and
                    np.array_equal(merged.q.view(np.uint32), whole.q.view(np.uint32)))
            results['two-level' if instanced else 'flat'] = {'merged_equals_one_gpu': bool(same), 'channels_hit': int(whole.hit.sum())}
        if comm is not None:
            comm.barrier()
    if rank == 0:
        print(json.dumps({'check': 'one event over N GPUs: merged channels == one GPU', 'n_gpus': world, 'photons': n, **results}))
        assert all(v['merged_equals_one_gpu'] for v in results.values())
    if comm is not None:
        comm.barrier()
        comm.close()


if __name__ == '__main__':
    main()
