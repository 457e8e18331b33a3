"""One propagate launch of the C4 shape (scintillator + WLS + dichroic detector of the parity tests,
380 nm bombs, max_steps=100): CUDA-event time per launch, for ncu captures of the physics-bound regime."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import fixtures as F
from chroma_b200 import event, gpu
from oracle import oracle as O
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
ctx = gpu.create_cuda_context(0)
det = F.attach_bvh(F.scintillator_detector(), O.build_bvh)
gg = gpu.GPUDetector(det)
rng = gpu.get_rng_states(1, seed=3)
for i in range(reps):
    gp = gpu.GPUPhotons(event.PhotonSource(n, wavelengths=380.0, seed=11, sequence=i))
    ctx.synchronize(); ctx.timer_begin()
    gp.propagate(gg, rng, max_steps=100)
    ms = ctx.timer_end()
    f = gp.flags.get()
    print('launch %d: %.2f ms, %.3g photons/s, reemit %.3f' % (i, ms, n / ms * 1e3, ((f >> 9) & 1).mean()), flush=True)
