"""GPU-box experiment: pageable upload rate (GPUPhotons(src) + synchronise) against copy threads / chunks, beside
the host's own memcpy rate (numpy copy of the same arrays with N threads)."""
import json, os, sys, time, threading
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
from chroma_b200 import gpu, event
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000000
ctx = gpu.create_cuda_context(0)
ph = bench.make_bomb(n, seed=100)
F = ('pos', 'dir', 'pol', 'wavelengths', 't', 'last_hit_triangles', 'flags', 'weights', 'evidx')
php = event.Photons(*[gpu.pinned_copy(getattr(ph, f)) for f in F])
nbytes = sum(getattr(ph, f).nbytes for f in F)
# host memcpy rate: pageable -> pinned with T python threads (numpy releases the GIL in copyto)
for T in (1, 2, 4, 8, 16):
    srcs = [getattr(ph, f).reshape(-1) for f in F]; dsts = [getattr(php, f).reshape(-1) for f in F]
    def work(k):
        for s, d in zip(srcs, dsts):
            m = len(s); a, b = m * k // T, m * (k + 1) // T
            np.copyto(d[a:b], s[a:b])
    best = 1e9
    for rep in range(3):
        th = [threading.Thread(target=work, args=(k,)) for k in range(T)]
        t0 = time.perf_counter(); [t.start() for t in th]; [t.join() for t in th]; best = min(best, time.perf_counter() - t0)
    print(json.dumps(dict(host_memcpy_threads=T, GBs=round(nbytes / best / 1e9, 1), ms=round(best * 1e3, 1))), flush=True)
def up(src):
    ctx.synchronize(); t0 = time.perf_counter(); g = gpu.GPUPhotons(src); ctx.synchronize(); dt = time.perf_counter() - t0
    g.get_flags_sum = None; del g
    return dt
for label, src in (('pinned', php), ('pageable', ph)):
    for threads, chunks in ((8, 16), (16, 16), (4, 16), (8, 8), (8, 32), (16, 32), (12, 64), (16, 64)):
        os.environ['CHROMA_UPLOAD_THREADS'] = str(threads); os.environ['CHROMA_UPLOAD_CHUNKS'] = str(chunks)
        ctx.reload_knobs()
        ts = [up(src) for _ in range(5)]
        print(json.dumps(dict(src=label, threads=threads, chunks=chunks, ms=round(float(np.median(ts[1:])) * 1e3, 2),
                              GBs=round(nbytes / float(np.median(ts[1:])) / 1e9, 1))), flush=True)
        if label == 'pinned' and (threads, chunks) == (8, 8): break
