"""GPU experiment: wide tree vs reference-format tree — exactness, visit counts, timing.
usage: python tools/wide_check.py [spiral_step] [nphotons]"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
from chroma_b200 import demo, gpu, event
from chroma_b200.loader import build
import fixtures as F
sys.argv += [None, None]
step = float(sys.argv[1] or 206.0)
nph = int(sys.argv[2] or 4000000)
ctx = gpu.create_cuda_context(0)
t0 = time.time()
det = demo.detector(spiral_step=step); det.flatten()
t1 = time.time()
build(det, fetch_nodes=False)
t2 = time.time()
gg = gpu.GPUDetector(det)
ctx.synchronize()
t3 = time.time()
print(json.dumps(dict(pmts=det.num_channels(), tris=len(det.mesh.triangles), mesh_s=t1-t0, bvh_s=t2-t1, geom_s=t3-t2, info={k:int(v) for k,v in gg.info().items()})))

def timed(fn, reps=5):
    fn(); ms=[]
    for _ in range(reps):
        ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin(); fn(); ms.append(ctx.timer_end())
    return float(np.median(ms))

def rays(kind, n, seed):
    pos, d = F.isotropic_rays(n, seed=seed, radius=(0.0 if kind == 'center' else 10000.0))
    if kind == 'center':
        from chroma_b200.tools import argsort_direction
        d = d[argsort_direction(d)]
    return pos, d

for kind in ('center', 'ball'):
    pos, d = rays(kind, 1000000, 3)
    dpos, ddir = gpu.to_gpu(pos.ravel()), gpu.to_gpu(d.ravel())
    res = {}
    for v in (2, 3):
        os.environ['CHROMA_ISECT_V'] = str(v)
        tri, dist = gpu.intersect(gg, dpos, ddir)
        res[v] = (tri.get(), dist.get())
        res['ms%d' % v] = timed(lambda: gpu.intersect(gg, dpos, ddir))
    bad = np.flatnonzero((res[2][0] != res[3][0]) | (res[2][1].view(np.uint32) != res[3][1].view(np.uint32)))
    os.environ['CHROMA_COUNT_WIDE'] = '0'
    _, _, c_ref = gpu.intersect(gg, dpos, ddir, counts=True)
    os.environ['CHROMA_COUNT_WIDE'] = '1'
    _, _, c_w = gpu.intersect(gg, dpos, ddir, counts=True)
    c_ref, c_w = c_ref.get(), c_w.get()
    out = dict(kind=kind, mismatches=int(len(bad)), ref_ms=res['ms2'], wide_ms=res['ms3'],
               ref_mrays=1e3/res['ms2'], wide_mrays=1e3/res['ms3'],
               ref_nodes=float(c_ref[:,0].mean()), ref_tris=float(c_ref[:,1].mean()),
               wide_groups=float(c_w[:,0].mean()), wide_tris=float(c_w[:,1].mean()),
               wide_groups_max=int(c_w[:,0].max()), hitfrac=float((res[3][0]>=0).mean()))
    if len(bad):
        b = bad[:5]
        out['first_bad'] = [(int(i), int(res[2][0][i]), float(res[2][1][i]), int(res[3][0][i]), float(res[3][1][i])) for i in b]
    print(json.dumps(out))
    # second bounce from the hit points
    ok = res[3][0] >= 0
    hp = (pos + d * res[3][1][:, None])[ok].astype(np.float32)
    _, d2 = F.isotropic_rays(len(hp), seed=9)
    last = res[3][0][ok].astype(np.int32)
    r = {}
    dp, dd, dl = gpu.to_gpu(hp.ravel()), gpu.to_gpu(d2.ravel()), gpu.to_gpu(last)
    for v in (2, 3):
        os.environ['CHROMA_ISECT_V'] = str(v)
        tri, dist = gpu.intersect(gg, dp, dd, dl)
        r[v] = (tri.get(), dist.get())
        r['ms%d' % v] = timed(lambda: gpu.intersect(gg, dp, dd, dl))
    bad = np.flatnonzero((r[2][0] != r[3][0]) | (r[2][1].view(np.uint32) != r[3][1].view(np.uint32)))
    os.environ['CHROMA_COUNT_WIDE'] = '0'
    _, _, c_ref = gpu.intersect(gg, dp, dd, dl, counts=True)
    os.environ['CHROMA_COUNT_WIDE'] = '1'
    _, _, c_w = gpu.intersect(gg, dp, dd, dl, counts=True)
    c_ref, c_w = c_ref.get(), c_w.get()
    print(json.dumps(dict(kind=kind + '-bounce', n=len(hp), mismatches=int(len(bad)), ref_ms=r['ms2'], wide_ms=r['ms3'],
                          ref_nodes=float(c_ref[:,0].mean()), ref_tris=float(c_ref[:,1].mean()),
                          wide_groups=float(c_w[:,0].mean()), wide_tris=float(c_w[:,1].mean()),
                          wide_groups_max=int(c_w[:,0].max()))))
os.environ.pop('CHROMA_ISECT_V')

# propagate
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
ph = bench.make_bomb(nph, seed=100)
rng = gpu.get_rng_states(1, seed=1)
from chroma_b200 import _lib
pristine = gpu.GPUPhotons(ph)
gp = gpu.GPUPhotons(ph)
for setting in [{'CHROMA_PROP_V': 2}, {'CHROMA_PROP_V': 4}, {'CHROMA_PROP_V': 4, 'CHROMA_WIDE_ITERS': 1},
                {'CHROMA_PROP_V': 4, 'CHROMA_WIDE_ITERS': 2}, {'CHROMA_PROP_V': 4, 'CHROMA_WIDE_ITERS': 8},
                {'CHROMA_PROP_V': 4, 'CHROMA_REFILL': 4}, {'CHROMA_PROP_V': 4, 'CHROMA_REFILL': 16}]:
    for k, v in setting.items(): os.environ[k] = str(v)
    ms = []
    for i in range(5):
        _lib.call('chroma_photons_copy', gp.handle, pristine.handle)
        ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin()
        gp.propagate(gg, rng, max_steps=10)
        ms.append(ctx.timer_end())
    fl = gp.get().flags
    print(json.dumps(dict(setting=setting, med_ms=float(np.median(ms[1:])), mphot_s=nph/np.median(ms[1:])/1e3,
                          det=float(((fl & event.SURFACE_DETECT) != 0).mean()),
                          nohit=float(((fl & event.NO_HIT) != 0).mean()))))
    for k in setting: os.environ.pop(k)
