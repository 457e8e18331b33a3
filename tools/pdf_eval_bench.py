"""GPU measurement of the PDF-evaluation row: GPUPDF.accumulate_pdf_eval (one fused launch) against
the reference's accumulate_bincount + sync + accumulate_nearest_neighbor_block (oracle/_ref/pdf.cubin)
on the same channel arrays, chroma/benchmark.py:157-233 shape (29,023 channels, ndaq = 128,
min_bin_content = 20, min_twidth 0.05).  usage: python tools/pdf_eval_bench.py [nch] [ndaq] [k] [passes]"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from chroma_b200 import gpu
from oracle import ref_gpu

sys.argv += [None] * 5
nch, ndaq, k, passes = int(sys.argv[1] or 29023), int(sys.argv[2] or 128), int(sys.argv[3] or 20), int(sys.argv[4] or 40)
ctx = gpu.create_cuda_context(0)
rctx = ref_gpu.RefContext(0)
rs = np.random.RandomState(1)
event_hit = rs.uniform(size=nch) < 0.5
event_time = rs.normal(60.0, 10.0, nch).astype(np.float32)
trange, width = (-0.5, 999.5), 0.05
batches = []
for _ in range(passes):
    t = rs.normal(60.0, 10.0, size=(ndaq, nch)).astype(np.float32)
    t[rs.uniform(size=(ndaq, nch)) > 0.5] = 1e9
    batches.append(t)

pdf = gpu.GPUPDF()
pdf.setup_pdf_eval(event_hit, event_time, np.zeros(nch), width, trange, 1.0, (-0.5, 20), min_bin_content=k)
chans = [gpu.GPUChannels(gpu.to_gpu(t.ravel()), gpu.to_gpu(np.zeros(1, np.float32)), gpu.to_gpu(np.zeros(1, np.uint32)),
                         ndaq=ndaq, stride=nch) for t in batches]
ours = []
for rep in range(3):
    pdf.clear_pdf_eval(); ctx.synchronize(); ctx.timer_begin()
    for c in chans:
        pdf.accumulate_pdf_eval(c)
    ours.append(ctx.timer_end())
hc, bc, nn = pdf.get_pdf_eval_arrays()

nhit = int(event_hit.sum())
d_hit, d_time = rctx.to_gpu(event_hit.astype(np.uint32)), rctx.to_gpu(event_time)
d_h2c = rctx.to_gpu(np.where(event_hit)[0].astype(np.uint32))
d_c2h = rctx.to_gpu(np.maximum(0, event_hit.cumsum() - 1).astype(np.uint32))
d_mc = [rctx.to_gpu(t.ravel()) for t in batches]
ref = []
for rep in range(3):
    d_hc, d_bc = rctx.to_gpu(np.zeros(nch, np.uint32)), rctx.to_gpu(np.zeros(nch, np.uint32))
    d_nn = rctx.to_gpu(np.full(nhit * k, 1e9, np.float32))
    d_q = rctx.alloc(4 * nhit * (ndaq + 1))
    rctx.synchronize(); rctx.timer_begin()
    for m in d_mc:
        rctx.fill32(d_q, 1, nhit * (ndaq + 1))           # gpu/pdf.py:298-299 work_queues.fill(1)
        rctx.launch('pdf', 'accumulate_bincount', nch // 64 + 1, 64, np.int32(nch), np.int32(ndaq), d_hit, d_time,
                    m, d_hc, d_bc, np.float32(width), np.float32(trange[0]), np.float32(trange[1]), np.int32(k),
                    d_c2h, d_q)
        rctx.synchronize()                                  # gpu/pdf.py:315
        rctx.launch('pdf', 'accumulate_nearest_neighbor_block', nhit, 64, np.int32(nhit), np.int32(ndaq), d_h2c,
                    d_q, d_time, m, d_nn, np.int32(k))
        rctx.synchronize()                                  # gpu/pdf.py:328
    ref.append(rctx.timer_end())
same = (np.array_equal(hc, rctx.get(d_hc, np.uint32, nch)) and np.array_equal(bc, rctx.get(d_bc, np.uint32, nch))
        and np.array_equal(nn.ravel().view(np.uint32), rctx.get(d_nn, np.float32, nhit * k).view(np.uint32)))
bytes_per_pass = 4.0 * ndaq * nch
print(json.dumps(dict(nch=nch, ndaq=ndaq, min_bin_content=k, passes=passes, identical=bool(same),
                      engine_ms_per_pass=min(ours) / passes, reference_ms_per_pass=min(ref) / passes,
                      speedup=min(ref) / min(ours), engine_gb_s=bytes_per_pass / (min(ours) / passes * 1e-3) / 1e9,
                      daq_replicas_per_s=ndaq / (min(ours) / passes * 1e-3))))
