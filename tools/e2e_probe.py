"""GPU experiment: where the end-to-end step spends its time (upload / propagate / daq)."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from chroma_b200 import demo, gpu, event, _lib
from chroma_b200.loader import build
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000000
ctx = gpu.create_cuda_context(0)
det = demo.detector(spiral_step=206.0); det.flatten(); build(det, fetch_nodes=False)
gg = gpu.GPUDetector(det); rng = gpu.get_rng_states(1, seed=1); daq = gpu.GPUDaq(gg)
ph = bench.make_bomb(n, seed=100)
F = ('pos', 'dir', 'pol', 'wavelengths', 't', 'last_hit_triangles', 'flags', 'weights', 'evidx')
php = event.Photons(*[gpu.pinned_copy(getattr(ph, f)) for f in F])
def T(): ctx.synchronize(); return time.perf_counter()
for label, src in (('pageable', ph), ('pinned', php)):
    for chunks in (1, 4, 8, 16):
        os.environ['CHROMA_UPLOAD_CHUNKS'] = str(chunks)
        rows = []
        for i in range(4):
            t0 = T(); g2 = gpu.GPUPhotons(src); t1 = T()
            g2.propagate(gg, rng, max_steps=10); t2 = T()
            daq.begin_acquire(); daq.acquire(g2, rng); ch = daq.end_acquire(); h = ch.get(); t3 = T()
            del g2; t4 = T()
            # pipelined: no sync between the calls
            t5 = T(); g2 = gpu.GPUPhotons(src); g2.propagate(gg, rng, max_steps=10)
            daq.begin_acquire(); daq.acquire(g2, rng); ch = daq.end_acquire(); h = ch.get(); t6 = T(); del g2
            rows.append([(t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, (t4-t3)*1e3, (t6-t5)*1e3])
        r = np.median(np.array(rows[1:]), axis=0)
        print(json.dumps(dict(src=label, chunks=chunks, upload_ms=round(r[0],2), prop_ms=round(r[1],2), daq_ms=round(r[2],2),
                              destroy_ms=round(r[3],2), pipelined_ms=round(r[4],2), h2d_GBs=round(n*60/r[0]/1e6,1))), flush=True)
