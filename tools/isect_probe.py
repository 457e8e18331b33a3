"""GPU experiment: intersect timing of sorted centre rays under env settings."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
from chroma_b200 import demo, gpu
from chroma_b200.loader import build
from chroma_b200.tools import argsort_direction
import fixtures as F
ctx = gpu.create_cuda_context(0)
det = demo.detector(spiral_step=206.0); det.flatten(); build(det, fetch_nodes=False)
gg = gpu.GPUDetector(det)
pos, d = F.isotropic_rays(1000000, seed=3)
ds = d[argsort_direction(d)]
for name, dd in (('sorted', ds), ('unsorted', d)):
    dp, dv = gpu.to_gpu(pos.ravel()), gpu.to_gpu(dd.ravel())
    for setting in [{}, {'CHROMA_WIDE_LEAF_THRESH': 1}, {'CHROMA_WIDE_LEAF_THRESH': 33}, {'CHROMA_WIDE_ITERS': 1}, {'CHROMA_ISECT_REFILL': 32}, {'CHROMA_ISECT_REFILL': 1}]:
        for k, v in setting.items(): os.environ[k] = str(v)
        ms = []
        for i in range(7):
            ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin(); gpu.intersect(gg, dp, dv); ms.append(round(ctx.timer_end(), 2))
        print(json.dumps(dict(rays=name, setting=setting, ms=ms)), flush=True)
        for k in setting: os.environ.pop(k)
