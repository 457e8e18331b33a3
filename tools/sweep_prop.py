"""Parameter sweep of the propagate kernel on one detector (dev tool, GPU box only).
SWEEP = JSON list of {env var: value}; pseudo-keys: _reps (timed reps), _fixed_epoch (1: new RNG
handle per rep, so every rep propagates identical histories)."""
import gc, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from chroma_b200 import gpu, _lib

class A: pass
args = A(); args.spiral_step = float(os.environ.get('SWEEP_STEP', '350')); args.max_steps = int(os.environ.get('SWEEP_MAXSTEPS', '10'))
n = int(os.environ.get('SWEEP_N', '2000000'))
ctx = gpu.create_cuda_context(0)
det, _, _ = bench.build_detector(args, fetch_nodes=False)
ph = bench.make_bomb(n, seed=100)
settings = json.loads(os.environ.get('SWEEP', '[{}]'))
GEOM_KEYS = ('CHROMA_TOP_NODES', 'CHROMA_OPTICS_SMEM_FLOATS')
gg, gg_key = None, None
pristine = gpu.GPUPhotons(ph); gp = gpu.GPUPhotons(ph)
for s in settings:
    env = {k: v for k, v in s.items() if not k.startswith('_')}
    for k, v in env.items():
        os.environ[k] = str(v)
    key = tuple(os.environ.get(k) for k in GEOM_KEYS)
    if gg is None or key != gg_key:
        if gg is not None:
            _lib.call('chroma_geometry_destroy', gg.handle); gg.handle = None
            gg = None; gc.collect()
        gg, gg_key = gpu.GPUDetector(det), key
    reps = int(s.get('_reps', 6))
    rng = gpu.get_rng_states(1, seed=1)
    ms = []
    for i in range(reps):
        if s.get('_fixed_epoch'):
            rng = gpu.get_rng_states(1, seed=1)
        _lib.call('chroma_photons_copy', gp.handle, pristine.handle)
        ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin()
        gp.propagate(gg, rng, max_steps=args.max_steps)
        ms.append(ctx.timer_end())
    fl = gp.flags.get()
    t = np.array(ms[1:])
    print(json.dumps(dict(setting=s, med_ms=round(float(np.median(t)), 2), min_ms=round(float(t.min()), 2),
                          max_ms=round(float(t.max()), 2), outliers=[round(float(x), 1) for x in t[t > 1.15 * np.median(t)]],
                          mphot_s=round(n / np.median(t) / 1e3, 2), det=float(((fl & 4) != 0).mean()))), flush=True)
    for k in env:
        os.environ.pop(k, None)
