"""Distribution of reference node visits per ray for 2nd-step rays in a demo detector (dev tool, GPU)."""
import os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from chroma_b200 import gpu, event

class A: pass
args = A(); args.spiral_step = float(os.environ.get('SWEEP_STEP', '206')); args.max_steps = 10
n = int(os.environ.get('SWEEP_N', '2000000'))
ctx = gpu.create_cuda_context(0)
det, _, _ = bench.build_detector(args, fetch_nodes=False)
gg = gpu.GPUDetector(det)
ph = bench.make_bomb(n, seed=100)
rng = gpu.get_rng_states(1, seed=1)
gp = gpu.GPUPhotons(ph)
TERM = event.NO_HIT | event.BULK_ABSORB | event.SURFACE_DETECT | event.SURFACE_ABSORB | event.NAN_ABORT
for step in range(1, 5):
    out = gp.get()
    live = (out.flags & TERM) == 0
    pos, dir, last = out.pos[live], out.dir[live], out.last_hit_triangles[live]
    dpos, ddir = gpu.to_gpu(pos.astype(np.float32).ravel()), gpu.to_gpu(dir.astype(np.float32).ravel())
    dlast = gpu.to_gpu(last.astype(np.int32))
    tri, dist, cnt = gpu.intersect(gg, dpos, ddir, dlast, counts=True)
    c = cnt.get().astype(np.int64); dist = dist.get()
    nn = c[:, 0]
    q = {p: float(np.percentile(nn, p)) for p in (50, 90, 99, 99.9, 99.99)}
    print(json.dumps(dict(step=step, rays=int(live.sum()), mean=float(nn.mean()), pct=q, max=int(nn.max()),
                          tri_mean=float(c[:, 1].mean()), tri_max=int(c[:, 1].max()),
                          frac_work_top1pct=float(np.sort(nn)[-len(nn) // 100:].sum() / nn.sum()))), flush=True)
    top = np.argsort(nn)[::-1][:8]
    for i in top:
        print('   nodes %d tris %d dir %s pos %s dist %.3f' % (nn[i], c[i, 1], dir[i], pos[i], dist[i]))
    gp.propagate(gg, rng, max_steps=1)
