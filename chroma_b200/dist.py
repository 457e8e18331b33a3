"""Multi-GPU plumbing: photons shard over ranks (one process per GPU, geometry replicated);
``torch.distributed`` (NCCL over NVLink on GPUs, gloo on CPU for the tests) is used only to
merge per-channel DAQ results and PDF histograms.  The reference has no multi-GPU path
(SURVEY.md F7); the merge rules follow its per-channel atomics (chroma/cuda/daq.cu:73-76):
earliest time = min of the raw time bit patterns, charge = sum of the integer charges,
history = bitwise OR."""
import os

import numpy as np


def bind_to_gpu_numa(device_index):
    """Pin this process to the CPUs of the NUMA node its GPU hangs off, so that the page-locked
    upload buffers it allocates afterwards (first touch) are local to that GPU's PCIe root.
    Returns a short description, or None when the topology cannot be read (single node, no sysfs)."""
    try:
        import subprocess
        bus = subprocess.run(['nvidia-smi', '-i', str(device_index), '--query-gpu=pci.bus_id',
                              '--format=csv,noheader'], capture_output=True, text=True, timeout=20).stdout.strip()
        if not bus:
            return None
        dev = '/sys/bus/pci/devices/' + bus.lower()[-12:]     # 00000000:1B:00.0 -> 0000:1b:00.0
        with open(dev + '/numa_node') as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open('/sys/devices/system/node/node%d/cpulist' % node) as f:
            cpus = set()
            for part in f.read().strip().split(','):
                lo, _, hi = part.partition('-')
                cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return 'gpu %d (%s) -> numa node %d, %d cpus' % (device_index, bus, node, len(allowed))
    except Exception:
        return None


def shard_range(n, world, rank):
    "Contiguous share [start, stop) of n items for `rank`: keeps direction-sorted photons coherent."
    base, extra = divmod(int(n), int(world))
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def deal_events(nevents, world, rank):
    "Round-robin event ids for `rank` (whole events per rank: the per-event DAQ stays local)."
    return list(range(rank, nevents, world))


def event_epoch(global_event_index):
    """Call epoch that keys an event's random streams by its GLOBAL number: set ``rng.epoch`` to this
    before the event's propagate (which uses it) and run_daq (which uses the next one).  The Philox
    counter of every draw is (photon id, block, epoch, stream), so an event then gives the same photons
    and hits whichever rank runs it and whatever that rank ran before: PDFs summed over dealt events do
    not depend on the number of ranks, and no two ranks ever reuse a stream."""
    return (2 * int(global_event_index)) & 0xFFFFFFFF


def seed_for_rank(seed, rank):
    """A seed for `rank` whose streams are independent of every other rank's (splitmix64 of seed and
    rank; rank 0 keeps `seed`, so single-process results are unchanged).  For jobs that deal events to
    ranks without keying them by event number (event_epoch): with the same seed on every rank the
    ranks would draw IDENTICAL uniforms for photon i of their k-th launch."""
    seed = int(seed) & (2 ** 64 - 1)
    if int(rank) == 0:
        return seed
    z = (seed + 0x9E3779B97F4A7C15 * (int(rank) + 1)) & (2 ** 64 - 1)
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & (2 ** 64 - 1)
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & (2 ** 64 - 1)
    return z ^ (z >> 31)


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


def merge_channels_packed(packed, group=None):
    """The same merge with ONE collective: `packed` is an int32 tensor [3, nchannels] (time bits, charge
    ints, history words); every rank gathers all ranks' copies (348 KB per rank at 29,023 channels) and
    reduces them locally: min / sum / or.  In place."""
    import torch
    import torch.distributed as td
    world = td.get_world_size(group)
    gathered = torch.empty((world,) + tuple(packed.shape), dtype=packed.dtype, device=packed.device)
    td.all_gather(list(gathered.unbind(0)), packed.contiguous(), group=group)
    packed[0] = gathered[:, 0].min(dim=0).values
    packed[1] = gathered[:, 1].sum(dim=0, dtype=packed.dtype)
    acc = gathered[0, 2].clone()
    for r in range(1, world):
        acc |= gathered[r, 2]
    packed[2] = acc
    return packed


def merge_pdf_tensors(hitcount, pdf, group=None):
    import torch.distributed as td
    td.all_reduce(hitcount, op=td.ReduceOp.SUM, group=group)
    td.all_reduce(pdf, op=td.ReduceOp.SUM, group=group)


def merge_pdf_eval_tensors(hitcount, bincount, nearest, group=None):
    """Merge per-rank PDF-evaluation accumulators (GPUPDF.get_pdf_eval_arrays) of ranks that were
    dealt different Monte Carlo events: counts add; the k-nearest table of a hit channel becomes the
    k smallest of all ranks' entries (1e9 = empty sorts last).  `nearest` is [nhit, k] float32.
    The merged value equals the single-rank one: distances are only dropped once a rank's own in-bin
    count has reached k, and then the summed count selects the fixed-bin estimate anyway."""
    import torch
    import torch.distributed as td
    td.all_reduce(hitcount, op=td.ReduceOp.SUM, group=group)
    td.all_reduce(bincount, op=td.ReduceOp.SUM, group=group)
    if nearest.numel():
        world = td.get_world_size(group)
        gathered = [torch.empty_like(nearest) for _ in range(world)]
        td.all_gather(gathered, nearest.contiguous(), group=group)
        k = nearest.shape[1]
        nearest.copy_(torch.sort(torch.cat(gathered, dim=1), dim=1).values[:, :k])


def merge_kernel_pdf_tensors(hitcount, time_sum, charge_sum, group=None):
    "Sum per-rank kernel-estimate accumulators (GPUKernelPDF.get_kernel_sums / get_moments)."
    import torch.distributed as td
    for t in (hitcount, time_sum, charge_sum):
        td.all_reduce(t, op=td.ReduceOp.SUM, group=group)


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

    def rng_states(self, size, seed=1):
        "get_rng_states for this rank: streams independent of the other ranks' (seed_for_rank)"
        from chroma_b200 import gpu
        return gpu.get_rng_states(size, seed=seed_for_rank(seed, self.rank))

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
        Call between GPUDaq.acquire() and end_acquire().  Pack (chroma_daq_pack), ONE all-gather of the
        packed block, reduce (chroma_daq_merge_gathered: min / sum / or), all on the library's own stream
        (torch.cuda.ExternalStream): ordered after run_daq and before end_acquire without any host
        synchronisation.  merge_channels_packed above states the same rule on tensors (CPU tests)."""
        import ctypes as C
        import torch.distributed as td
        from chroma_b200 import _lib
        torch = self.torch
        n = int(gpu_daq.earliest_time_int_gpu.size)
        stream = torch.cuda.ExternalStream(gpu_daq.ctx.stream_ptr(), device=self.device)
        with torch.cuda.stream(stream):
            packed = torch.empty((3, n), dtype=torch.int32, device=self.device)
            gathered = torch.empty((self.world, 3, n), dtype=torch.int32, device=self.device)
            _lib.call('chroma_daq_pack', gpu_daq.handle, C.c_void_p(packed.data_ptr()))
            if td.get_backend() == 'nccl':
                td.all_gather_into_tensor(gathered, packed)
            else:
                # gloo gathers host tensors only (tests with several ranks on one GPU): through the host
                host = [torch.empty((3, n), dtype=torch.int32) for _ in range(self.world)]
                td.all_gather(host, packed.cpu())
                gathered.copy_(torch.stack(host))
            _lib.call('chroma_daq_merge_gathered', gpu_daq.handle, C.c_void_p(gathered.data_ptr()), self.world)

    def reduce_pdf_eval(self, gpu_pdf):
        """(hitcount, pdf value, pdf uncertainty) of GPUPDF.get_pdf_eval() over all ranks' Monte
        Carlo (events dealt with deal_events); a few KB per rank, moved as host arrays."""
        from chroma_b200.gpu.pdf import pdf_eval_values
        t = self.torch
        hitcount, bincount, nearest = gpu_pdf.get_pdf_eval_arrays()
        hc = t.from_numpy(hitcount.astype(np.int64)).to(self.device)
        bc = t.from_numpy(bincount.astype(np.int64)).to(self.device)
        nn = t.from_numpy(nearest).to(self.device)
        merge_pdf_eval_tensors(hc, bc, nn)
        return pdf_eval_values(gpu_pdf.event_hit, hc.cpu().numpy().astype(np.uint32),
                               bc.cpu().numpy().astype(np.uint32), nn.cpu().numpy(),
                               gpu_pdf.map_hit_offset_to_channel_id, gpu_pdf.min_twidth,
                               gpu_pdf.min_bin_content)

    def reduce_kernel_eval(self, gpu_kernel_pdf):
        "(hitcount, pdf value, zeros) of GPUKernelPDF.get_kernel_eval() over all ranks' Monte Carlo."
        t = self.torch
        hitcount, tsum, qsum = gpu_kernel_pdf.get_kernel_sums()
        hc = t.from_numpy(hitcount.astype(np.int64)).to(self.device)
        ts, qs = t.from_numpy(tsum).to(self.device), t.from_numpy(qsum).to(self.device)
        merge_kernel_pdf_tensors(hc, ts, qs)
        hitcount = hc.cpu().numpy().astype(np.uint32)
        tv = ts.cpu().numpy() / np.maximum(1, hitcount)
        qv = qs.cpu().numpy() / np.maximum(1, hitcount)
        value = tv if gpu_kernel_pdf.time_only else tv * qv
        return hitcount, value, np.zeros_like(value)

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
