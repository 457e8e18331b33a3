"""GPU experiment: propagate timing under env settings (geometry rebuilt when a CHROMA_WIDE_* build
setting changes).  usage: python tools/prop_sweep.py settings.json|'[{"ENV": value, ...}, ...]' [spiral_step] [nphotons] [reps]
CHROMA_B200_LIB selects an A/B build of the library (python -m chroma_b200.csrc.build --variant ...)."""
import json, os, sys, zlib
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from chroma_b200 import demo, gpu, event, _lib
from chroma_b200.loader import build
import bench
sys.argv += [None] * 4
settings = json.loads(sys.argv[1]) if sys.argv[1].lstrip().startswith('[') else json.load(open(sys.argv[1]))
step = float(sys.argv[2] or 206.0)
nph = int(sys.argv[3] or 4000000)
reps = int(sys.argv[4] or 5)
BUILD_KEYS = ('CHROMA_WIDE_FAT', 'CHROMA_WIDE_LEAF', 'CHROMA_WIDE', 'CHROMA_WIDE_SAH', 'CHROMA_OPTICS_SMEM_FLOATS')
ctx = gpu.create_cuda_context(0)
det = demo.detector(spiral_step=step)
if not all(st.get('_instanced') for st in settings):      # the two-level geometry needs neither mesh nor BVH
    det.flatten()
    build(det, fetch_nodes=False)
ph = bench.make_bomb(nph, seed=100)
rng = gpu.get_rng_states(1, seed=1)
pristine = gpu.GPUPhotons(ph)
gp = gpu.GPUPhotons(ph)


def direction_order(d, bits=12):
    "host stand-in for a device sort: Morton order of the octahedral map of the direction"
    d = d / np.abs(d).sum(axis=1)[:, None]
    u, v = d[:, 0].copy(), d[:, 1].copy()
    neg = d[:, 2] < 0
    u[neg] = (1 - np.abs(d[neg, 1])) * np.sign(d[neg, 0])
    v[neg] = (1 - np.abs(d[neg, 0])) * np.sign(d[neg, 1])
    q = lambda x: np.clip(((x * 0.5 + 0.5) * (1 << bits)).astype(np.uint64), 0, (1 << bits) - 1)
    def spread(x):
        out = np.zeros_like(x)
        for b in range(bits):
            out |= ((x >> np.uint64(b)) & np.uint64(1)) << np.uint64(2 * b)
        return out
    return np.argsort(spread(q(u)) | (spread(q(v)) << np.uint64(1)), kind='stable')


order = direction_order(np.asarray(ph.dir, np.float64))
ph_sorted = ph[order]
pristine_sorted = gpu.GPUPhotons(ph_sorted)
gg, key = None, None
for setting in settings:
    for k, v in setting.items():
        if not k.startswith('_'): os.environ[k] = str(v)
    ctx.reload_knobs()
    bk = tuple(os.environ.get(k) for k in BUILD_KEYS) + (bool(setting.get('_instanced')),)
    if gg is None or bk != key:
        gg = None
        import gc; gc.collect()
        gg = gpu.GPUDetector(det, instanced=bool(setting.get('_instanced'))); key = bk
    ms = []
    for i in range(reps + 1):
        _lib.call('chroma_photons_copy', gp.handle, (pristine_sorted if setting.get('_sort') else pristine).handle)
        ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin()
        gp.propagate(gg, rng, max_steps=int(setting.get('_max_steps', 10)))
        ms.append(ctx.timer_end())
    fl = gp.get().flags
    info = gg.info()
    info.update(tree=gg.tree())
    print(json.dumps(dict(setting=setting, tree=info['tree']['kind'], tree_bytes=info['tree']['bytes'], med_ms=round(float(np.median(ms[1:])), 3), mphot_s=round(nph/np.median(ms[1:])/1e3, 2),
                          det=float(((fl & event.SURFACE_DETECT) != 0).mean()), flags_crc=zlib.crc32(np.ascontiguousarray(fl).tobytes()), groups=int(info['wide_groups']),
                          levels=int(info['wide_levels']), leaves=int(info['wide_leaves']))), flush=True)
    for k in setting:
        if not k.startswith('_'): os.environ.pop(k)
