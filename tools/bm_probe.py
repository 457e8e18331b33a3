"""One-off: per-iteration times of the chroma.benchmark intersect loop (looking for outliers)."""
import os, sys, time, gc
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from chroma_b200 import demo, gpu, sample, tools
from chroma_b200.loader import create_geometry_from_obj
ctx = gpu.create_cuda_context()
det = create_geometry_from_obj(demo.detector())
gg = gpu.GPUDetector(det)
n = 500000
for mode in ('discard', 'keep', 'discard_nogc'):
    if mode == 'discard_nogc':
        gc.disable()
    ts, ta, tb = [], [], []
    for i in range(100):
        pos = gpu.zeros(n * 3, np.float32)
        d = sample.uniform_sphere(n)
        d = gpu.to_gpu(np.ascontiguousarray(d[tools.argsort_direction(d)], dtype=np.float32).ravel())
        ctx.synchronize()
        t0 = time.time()
        if mode == 'keep':
            r = gpu.intersect(gg, pos, d)
        else:
            gpu.intersect(gg, pos, d)
        t1 = time.time()
        ctx.synchronize()
        t2 = time.time()
        ts.append((t2 - t0) * 1e3)
    ts = np.array(ts)
    print(mode, 'median %.3f mean %.3f max %.3f' % (np.median(ts), ts.mean(), ts.max()), 'slow:', [(i, round(x, 1)) for i, x in enumerate(ts) if x > 5])
