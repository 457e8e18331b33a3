"""Device photon source (chroma_photons_generate) against the HBM roofline, next to what it replaces:
numpy sampling on the host + the H2D upload of GPUPhotons (chroma/benchmark.py:48-66 load_photons).

Algorithmic bytes per photon: 64 written (four 16-byte records), 0 read.  CUDA-event timing on the
context's stream, L2 flushed between launches.  usage: python tools/source_bench.py [--photons N]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--photons', type=int, default=10_000_000)
    ap.add_argument('--reps', type=int, default=10)
    ap.add_argument('--out', default=None)
    args = ap.parse_args()
    from chroma_b200 import event, gpu, sample
    from chroma_b200.transform import normalize
    ctx = gpu.create_cuda_context(0)
    n = args.photons
    peak = None
    try:
        peak = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json'))).get('hbm_gbs')
    except Exception:
        pass
    g = gpu.GPUPhotons.empty(n)
    ms = {}
    for name, kw in (('isotropic', {}), ('flashlight', dict(phi=0.4, direction=(0, 1, 1)))):
        src = event.PhotonSource(n, wavelengths=(400.0, 800.0), seed=1, **kw)
        for _ in range(3):
            g.load(0, src)
        t = []
        for _ in range(args.reps):
            ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin()
            g.load(0, src)
            t.append(ctx.timer_end())
        ms[name] = t
    # variants (store policy, CTAs per SM) and the write-only ceiling of this box (cudaMemset of the
    # same 640 MB)
    variants = {}
    src = event.PhotonSource(n, wavelengths=(400.0, 800.0), seed=1)
    for store in (0, 1):
        for per_sm in (4, 8, 16, 64, 100000):
            os.environ['CHROMA_GEN_STORE'] = str(store)
            os.environ['CHROMA_GEN_CTAS_PER_SM'] = str(per_sm)
            g.load(0, src)
            t = []
            for _ in range(5):
                ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin()
                g.load(0, src)
                t.append(ctx.timer_end())
            variants['store%d_ctas%d' % (store, per_sm)] = round(float(np.median(t)), 4)
    os.environ.pop('CHROMA_GEN_STORE'); os.environ.pop('CHROMA_GEN_CTAS_PER_SM')
    buf = gpu.empty(n * 16, np.float32)
    t = []
    for _ in range(6):
        ctx.flush_l2(); ctx.synchronize(); ctx.timer_begin()
        buf.fill(1.0)
        t.append(ctx.timer_end())
    memset_ms = float(np.median(t[1:]))
    del buf
    # the pair it replaces: host sampling, then the upload (pageable arrays, as a user holds them)
    t0 = time.perf_counter()
    d = sample.uniform_sphere(n)
    pol = normalize(np.cross(sample.uniform_sphere(n), d))
    ph = event.Photons(np.zeros((n, 3)), d, pol, np.random.uniform(400, 800, n))
    t_host = time.perf_counter() - t0
    up = []
    for _ in range(4):
        ctx.synchronize()
        t0 = time.perf_counter()
        h = gpu.GPUPhotons(ph)
        h.flags.owner  # keep alive
        ctx.synchronize()
        up.append((time.perf_counter() - t0) * 1e3)
        del h
    med = float(np.median(ms['isotropic']))
    gbs = 64.0 * n / (med * 1e-3) / 1e9
    line = {'kernel': 'generate_photons_kernel', 'photons': n, 'ms_isotropic': [round(x, 4) for x in ms['isotropic']],
            'ms_flashlight': [round(x, 4) for x in ms['flashlight']], 'photons_per_s': n / (med * 1e-3),
            'roofline': {'bound': 'hbm', 'achieved': gbs, 'peak': peak, 'unit': 'GB/s',
                         'frac': (gbs / peak) if peak else None, 'bytes_per_photon': 64},
            'variants_ms': variants, 'memset_640MB_ms': memset_ms, 'memset_gbs': 64.0 * n / (memset_ms * 1e-3) / 1e9,
            'host_sampling_s': t_host, 'upload_ms': [round(x, 2) for x in up],
            'host_path_photons_per_s': n / (t_host + float(np.median(up)) * 1e-3)}
    s = json.dumps(line)
    print(s)
    if args.out:
        open(args.out, 'w').write(s + '\n')
    ctx.pop()


if __name__ == '__main__':
    main()
