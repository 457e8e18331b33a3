"""GPU experiment: propagate timing under env settings (geometry rebuilt when a CHROMA_WIDE_* build
setting changes).  usage: python tools/prop_sweep.py settings.json [spiral_step] [nphotons] [reps]"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from chroma_b200 import demo, gpu, event, _lib
from chroma_b200.loader import build
import bench
sys.argv += [None] * 4
settings = json.load(open(sys.argv[1]))
step = float(sys.argv[2] or 206.0)
nph = int(sys.argv[3] or 4000000)
reps = int(sys.argv[4] or 5)
BUILD_KEYS = ('CHROMA_WIDE_LEAF', 'CHROMA_WIDE', 'CHROMA_WIDE_SAH', 'CHROMA_OPTICS_SMEM_FLOATS')
ctx = gpu.create_cuda_context(0)
det = demo.detector(spiral_step=step); det.flatten()
build(det, fetch_nodes=False)
ph = bench.make_bomb(nph, seed=100)
rng = gpu.get_rng_states(1, seed=1)
pristine = gpu.GPUPhotons(ph)
gp = gpu.GPUPhotons(ph)
gg, key = None, None
for setting in settings:
    for k, v in setting.items():
        if not k.startswith('_'): os.environ[k] = str(v)
    bk = tuple(os.environ.get(k) for k in BUILD_KEYS)
    if gg is None or bk != key:
        gg = None
        import gc; gc.collect()
        gg = gpu.GPUDetector(det); key = bk
    ms = []
    for i in range(reps + 1):
        _lib.call('chroma_photons_copy', gp.handle, pristine.handle)
        ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin()
        gp.propagate(gg, rng, max_steps=int(setting.get('_max_steps', 10)))
        ms.append(ctx.timer_end())
    fl = gp.get().flags
    info = gg.info()
    print(json.dumps(dict(setting=setting, med_ms=round(float(np.median(ms[1:])), 3), mphot_s=round(nph/np.median(ms[1:])/1e3, 2),
                          det=float(((fl & event.SURFACE_DETECT) != 0).mean()), groups=int(info['wide_groups']),
                          levels=int(info['wide_levels']))), flush=True)
    for k in setting:
        if not k.startswith('_'): os.environ.pop(k)
