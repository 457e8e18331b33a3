"""GPU measurement of the render row: GPURays.render (near-first search of the wide tree, list's last
distance as pruning bound) against the reference's render kernel (oracle/_ref/render.cubin) on the
same rays and geometry.  usage: python tools/render_bench.py [spiral_step] [width] [height] [alpha_depth]"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from chroma_b200 import demo, gpu
from chroma_b200.loader import build
from chroma_b200.tools import from_film
from oracle import ref_gpu

sys.argv += [None] * 5
step = float(sys.argv[1] or 900.0)
size = (int(sys.argv[2] or 800), int(sys.argv[3] or 600))
ad = int(sys.argv[4] or 10)
ctx = gpu.create_cuda_context(0)
det = demo.detector(spiral_step=step); det.flatten()
build(det, fetch_nodes=True)
gg = gpu.GPUDetector(det)
radius = float(np.abs(det.mesh.vertices).max())
pos, dir = from_film(position=(0, -2.5 * radius, 0), size=size)
if (dir[:, 1] < 0).mean() > 0.5:
    dir = -dir
pos, dir = pos.astype(np.float32), dir.astype(np.float32)
rays = gpu.GPURays(pos, dir, max_alpha_depth=ad)
pixels = gpu.empty(len(pos), np.uint32)
ours = []
for _ in range(4):
    ctx.synchronize(); ctx.timer_begin()
    rays.render(gg, pixels, alpha_depth=ad)
    ours.append(ctx.timer_end())
px = pixels.get()
out = dict(pmts=det.num_channels(), triangles=len(det.mesh.triangles), rays=len(pos), alpha_depth=ad,
           engine_ms=min(ours[1:]), engine_mrays_s=len(pos) / min(ours[1:]) / 1e3,
           hit_fraction=float((rays.dxlen.get() > 0).mean()), mean_list=float(rays.dxlen.get().mean()))
rctx = ref_gpu.RefContext(0)
rg = ref_gpu.RefGeometry(rctx, det)
rref = ref_gpu.RefRays(rctx, pos, dir, max_alpha_depth=ad)
px_ref = rref.render(rg, alpha_depth=ad)
out['identical_pixels'] = bool(np.array_equal(px, px_ref))
out['pixels_differing'] = int((px != px_ref).sum())
out['reference_ms'] = rref.time_render(rg, alpha_depth=ad)
out['speedup'] = out['reference_ms'] / out['engine_ms']
print(json.dumps(out))
if os.environ.get('RENDER_OUT'):
    np.save(os.environ['RENDER_OUT'], px.reshape(size[0], size[1]))
