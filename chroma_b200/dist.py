"""Multi-GPU plumbing: photons shard over ranks (one process per GPU, geometry replicated);
``torch.distributed`` (NCCL over NVLink on GPUs, gloo on CPU for the tests) is used only to
merge per-channel DAQ results and PDF histograms.  The reference has no multi-GPU path
(SURVEY.md F7); the merge rules follow its per-channel atomics (chroma/cuda/daq.cu:73-76):
earliest time = min of the raw time bit patterns, charge = sum of the integer charges,
history = bitwise OR."""
import os

import numpy as np


def shard_range(n, world, rank):
    "Contiguous share [start, stop) of n items for `rank`: keeps direction-sorted photons coherent."
    base, extra = divmod(int(n), int(world))
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def deal_events(nevents, world, rank):
    "Round-robin event ids for `rank` (whole events per rank: the per-event DAQ stays local)."
    return list(range(rank, nevents, world))


def merge_channel_tensors(time_int, q_int, flags, group=None):
    """In-place merge of per-rank channel accumulators (torch int32 tensors, CPU or CUDA).
    NCCL has no bitwise OR, so the history words are all-gathered and OR-ed locally."""
    import torch
    import torch.distributed as td
    td.all_reduce(time_int, op=td.ReduceOp.MIN, group=group)
    td.all_reduce(q_int, op=td.ReduceOp.SUM, group=group)
    world = td.get_world_size(group)
    gathered = [torch.empty_like(flags) for _ in range(world)]
    td.all_gather(gathered, flags, group=group)
    acc = gathered[0]
    for g in gathered[1:]:
        acc = torch.bitwise_or(acc, g)
    flags.copy_(acc)


def merge_pdf_tensors(hitcount, pdf, group=None):
    import torch.distributed as td
    td.all_reduce(hitcount, op=td.ReduceOp.SUM, group=group)
    td.all_reduce(pdf, op=td.ReduceOp.SUM, group=group)


def _as_torch(device_array, dtype):
    "zero-copy torch view of a DeviceArray (same CUDA primary context)"
    import torch

    class _Iface(object):
        pass
    holder = _Iface()
    iface = dict(device_array.__cuda_array_interface__)
    iface['typestr'] = np.dtype(dtype).str
    holder.__cuda_array_interface__ = iface
    return torch.as_tensor(holder, device='cuda')


class Communicator(object):
    "One rank of a torchrun job (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* from the environment)."

    def __init__(self, world, rank, local_rank, backend=None):
        import torch
        import torch.distributed as td
        self.torch, self.td = torch, td
        self.world, self.rank, self.local_rank = world, rank, local_rank
        use_cuda = torch.cuda.is_available()
        if use_cuda:
            torch.cuda.set_device(local_rank)
        if not td.is_initialized():
            os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
            os.environ.setdefault('MASTER_PORT', '29511')
            td.init_process_group(backend or ('nccl' if use_cuda else 'gloo'), rank=rank,
                                  world_size=world)
        self.device = torch.device('cuda', local_rank) if use_cuda else torch.device('cpu')

    def close(self):
        if self.td.is_initialized():
            self.td.destroy_process_group()

    def barrier(self):
        if self.device.type == 'cuda':
            self.torch.cuda.synchronize()
        self.td.barrier()

    def max_float(self, x):
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.device)
        self.td.all_reduce(t, op=self.td.ReduceOp.MAX)
        return float(t.item())

    def sum_int(self, x):
        t = self.torch.tensor([int(x)], dtype=self.torch.int64, device=self.device)
        self.td.all_reduce(t, op=self.td.ReduceOp.SUM)
        return int(t.item())

    def reduce_channels(self, gpu_daq):
        """Merge one event's channel accumulators over the ranks that shared its photons.
        Call between GPUDaq.acquire() and end_acquire()."""
        gpu_daq.ctx.synchronize()
        t = _as_torch(gpu_daq.earliest_time_int_gpu, np.int32)
        q = _as_torch(gpu_daq.channel_q_int_gpu, np.int32)
        f = _as_torch(gpu_daq.channel_history_gpu, np.int32)
        merge_channel_tensors(t, q, f)
        self.torch.cuda.synchronize()

    def reduce_pdf(self, gpu_pdf):
        "Sum per-rank PDFs and hit counts (once per PDF, not per event)."
        gpu_pdf.ctx.synchronize()
        h = _as_torch(gpu_pdf.hitcount_gpu, np.int32)
        flat = gpu_pdf.pdf_gpu
        from chroma_b200.gpu.tools import DeviceArray
        view = DeviceArray(flat.size, np.uint32, ctx=flat.ctx, ptr=flat.ptr, base=flat)
        p = _as_torch(view, np.int32)
        merge_pdf_tensors(h, p)
        self.torch.cuda.synchronize()
