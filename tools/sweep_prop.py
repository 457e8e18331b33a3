"""Parameter sweep of the propagate kernel on one detector (dev tool, GPU box only)."""
import itertools, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from chroma_b200 import gpu, _lib

class A: pass
args = A(); args.spiral_step = float(os.environ.get('SWEEP_STEP', '350')); args.max_steps = 10
n = int(os.environ.get('SWEEP_N', '2000000'))
ctx = gpu.create_cuda_context(0)
det, _, _ = bench.build_detector(args, fetch_nodes=False)
ph = bench.make_bomb(n, seed=100)
rng = gpu.get_rng_states(1, seed=1)
settings = json.loads(os.environ.get('SWEEP', '[{}]'))
for s in settings:
    for k, v in s.items():
        os.environ[k] = str(v)
    gg = gpu.GPUDetector(det)
    pristine = gpu.GPUPhotons(ph); gp = gpu.GPUPhotons(ph)
    ms = []
    for i in range(4):
        _lib.call('chroma_photons_copy', gp.handle, pristine.handle)
        ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin()
        gp.propagate(gg, rng, max_steps=10)
        ms.append(ctx.timer_end())
    fl = gp.flags.get()
    print(json.dumps(dict(setting=s, ms=ms[1:], mphot_s=n / np.mean(ms[1:]) / 1e3, det=float(((fl & 4) != 0).mean()))), flush=True)
    del gg, gp, pristine
    for k in s:
        os.environ.pop(k, None)
