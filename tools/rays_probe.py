"""GPU experiment: the C2 ray shape of bench.py (1e7 isotropic rays from the centre of the 29,023-PMT detector,
directions of the bomb, L2 flushed) on its own, for an ncu capture of wide_intersect_kernel."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from chroma_b200 import demo, gpu
from chroma_b200.loader import build
import bench
rays = int(sys.argv[1]) if len(sys.argv) > 1 else 10000000
ctx = gpu.create_cuda_context(0)
det = demo.detector(spiral_step=206.0); det.flatten(); build(det, fetch_nodes=False)
gg = gpu.GPUDetector(det)
ph = bench.make_bomb(rays, seed=100)
rdir = gpu.to_gpu(np.ascontiguousarray(ph.dir, dtype=np.float32).ravel())
rpos = gpu.to_gpu(np.zeros(rays * 3, np.float32))
ms = []
for i in range(4):
    ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin()
    tri, dist = gpu.intersect(gg, rpos, rdir)
    ms.append(ctx.timer_end())
print(json.dumps(dict(rays=rays, ms=[round(m, 3) for m in ms], rays_per_s=rays / (np.median(ms[1:]) * 1e-3), hit=float((tri.get() >= 0).mean()))))
