"""Run the REFERENCE's own CUDA kernels on the GPU box — TEST INFRASTRUCTURE ONLY.

`make -C oracle ref` compiles the untouched reference sources (chroma/cuda/*.cu under
/root/reference) into sm_100a cubins under oracle/_ref/ with the reference's nvcc options
(``--use_fast_math``, chroma/gpu/tools.py:12).  The reference launches them through PyCUDA
(chroma/gpu/*.py), which is not installed here, so this module drives the same kernels through
the CUDA *driver* API with ctypes and rebuilds the pointer-linked device structs the way
chroma/gpu/geometry.py:47-253 and chroma/gpu/detector.py:37-46 do (`make_gpu_struct`,
chroma/gpu/tools.py:146-168).  It is

  * the bit-exact oracle of the deterministic tier (first-hit triangle + distance), and
  * the "reference kernels JIT-built on the same B200" arm of the benchmarks,
    with the reference's own launch shapes (64 threads x <= 1024 blocks per chunk,
    one launch + one 4-byte D2H per step while the queue holds >= 8192 photons,
    chroma/gpu/photon.py:193-259).

Only tests/, __graft_entry__.smoke() and bench.py import this file.  Nothing in chroma_b200/
does.  /root/reference is NOT read at run time: only the prebuilt cubins are.
"""
import ctypes as C
import os
import struct

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, '_ref')

_cu = None


def available():
    return all(os.path.exists(os.path.join(REF_DIR, f + '.cubin'))
               for f in ('propagate', 'daq', 'pdf', 'bvh', 'intersect', 'render', 'transform'))


class CudaDriverError(RuntimeError):
    pass


def _check(status, what):
    if status != 0:
        msg = C.c_char_p()
        try:
            _cu.cuGetErrorString(status, C.byref(msg))
            text = msg.value.decode() if msg.value else '?'
        except Exception:
            text = '?'
        raise CudaDriverError('%s failed: CUDA driver error %d (%s)' % (what, status, text))


def driver():
    global _cu
    if _cu is None:
        _cu = C.CDLL('libcuda.so.1')
        _check(_cu.cuInit(0), 'cuInit')
    return _cu


class DeviceBuffer(object):
    "cuMemAlloc'd buffer (freed with the owning RefContext or explicitly)"

    def __init__(self, ctx, nbytes):
        self.ctx = ctx
        self.nbytes = int(max(nbytes, 16))
        p = C.c_uint64()
        _check(ctx.cu.cuMemAlloc_v2(C.byref(p), C.c_size_t(self.nbytes)), 'cuMemAlloc')
        self.ptr = p.value

    def free(self):
        if self.ptr:
            self.ctx.cu.cuMemFree_v2(C.c_uint64(self.ptr))
            self.ptr = 0


class RefContext(object):
    """The device's primary context (shared with libchroma_b200's runtime-API context) and the
    reference modules."""

    def __init__(self, device=0):
        self.cu = driver()
        dev = C.c_int()
        _check(self.cu.cuDeviceGet(C.byref(dev), device), 'cuDeviceGet')
        self.dev = dev
        self.ctx = C.c_void_p()
        _check(self.cu.cuDevicePrimaryCtxRetain(C.byref(self.ctx), dev), 'cuDevicePrimaryCtxRetain')
        _check(self.cu.cuCtxSetCurrent(self.ctx), 'cuCtxSetCurrent')
        self.modules, self.funcs, self.buffers = {}, {}, []
        self.ev0, self.ev1 = C.c_void_p(), C.c_void_p()
        _check(self.cu.cuEventCreate(C.byref(self.ev0), 0), 'cuEventCreate')
        _check(self.cu.cuEventCreate(C.byref(self.ev1), 0), 'cuEventCreate')
        self.launches = 0

    def bind(self):
        _check(self.cu.cuCtxSetCurrent(self.ctx), 'cuCtxSetCurrent')

    def function(self, module, name):
        key = (module, name)
        if key not in self.funcs:
            if module not in self.modules:
                path = os.path.join(REF_DIR, module + '.cubin')
                if not os.path.exists(path):
                    raise CudaDriverError('%s missing: run `make -C oracle ref` where /root/reference '
                                          'is mounted' % path)
                m = C.c_void_p()
                _check(self.cu.cuModuleLoad(C.byref(m), path.encode()), 'cuModuleLoad(%s)' % path)
                self.modules[module] = m
            f = C.c_void_p()
            _check(self.cu.cuModuleGetFunction(C.byref(f), self.modules[module], name.encode()),
                   'cuModuleGetFunction(%s)' % name)
            self.funcs[key] = f
        return self.funcs[key]

    # -- memory -------------------------------------------------------------------------
    def alloc(self, nbytes):
        b = DeviceBuffer(self, nbytes)
        self.buffers.append(b)
        return b

    def to_gpu(self, array):
        a = np.ascontiguousarray(array)
        b = self.alloc(a.nbytes)
        if a.nbytes:
            _check(self.cu.cuMemcpyHtoD_v2(C.c_uint64(b.ptr), a.ctypes.data_as(C.c_void_p),
                                           C.c_size_t(a.nbytes)), 'cuMemcpyHtoD')
        return b

    def set(self, buf, array, offset=0):
        a = np.ascontiguousarray(array)
        _check(self.cu.cuMemcpyHtoD_v2(C.c_uint64(buf.ptr + offset), a.ctypes.data_as(C.c_void_p),
                                       C.c_size_t(a.nbytes)), 'cuMemcpyHtoD')

    def get(self, buf, dtype, count, offset=0):
        out = np.empty(count, dtype=dtype)
        if out.nbytes:
            _check(self.cu.cuMemcpyDtoH_v2(out.ctypes.data_as(C.c_void_p), C.c_uint64(buf.ptr + offset),
                                           C.c_size_t(out.nbytes)), 'cuMemcpyDtoH')
        return out

    def fill32(self, buf, value, count, offset=0):
        _check(self.cu.cuMemsetD32_v2(C.c_uint64(buf.ptr + offset), C.c_uint32(value), C.c_size_t(count)),
               'cuMemsetD32')

    def struct(self, fmt_items):
        """make_gpu_struct (chroma/gpu/tools.py:146-168): members packed in order, pointers
        8-byte aligned; `fmt_items` is a list of (struct format char, value)."""
        blob = b''
        for fmt, v in fmt_items:
            if fmt == 'P':
                assert len(blob) % 8 == 0, 'pointer member not 8-byte aligned'
                blob += struct.pack('<Q', v.ptr if isinstance(v, DeviceBuffer) else int(v))
            else:
                blob += struct.pack('<' + fmt, v)
        blob += b'\0' * ((-len(blob)) % 8)
        return self.to_gpu(np.frombuffer(blob, dtype=np.uint8))

    # -- launches -----------------------------------------------------------------------
    def launch(self, module, name, grid, block, *args):
        f = self.function(module, name)
        holders = []
        for a in args:
            if isinstance(a, DeviceBuffer):
                holders.append(C.c_uint64(a.ptr))
            elif isinstance(a, (np.int32, int)) and not isinstance(a, bool):
                holders.append(C.c_int32(int(a)))
            elif isinstance(a, np.uint32):
                holders.append(C.c_uint32(int(a)))
            elif isinstance(a, (np.float32, float)):
                holders.append(C.c_float(float(a)))
            elif isinstance(a, np.uint64):
                holders.append(C.c_uint64(int(a)))
            elif isinstance(a, tuple) and len(a) == 3:      # float3 by value
                holders.append((C.c_float * 3)(*[float(x) for x in a]))
            else:
                raise TypeError('unsupported kernel argument %r' % (a,))
        arr = (C.c_void_p * len(holders))(*[C.cast(C.byref(h), C.c_void_p) for h in holders])
        _check(self.cu.cuLaunchKernel(f, int(grid), 1, 1, int(block), 1, 1, 0, None, arr, None),
               'cuLaunchKernel(%s)' % name)
        self.launches += 1

    def synchronize(self):
        _check(self.cu.cuCtxSynchronize(), 'cuCtxSynchronize')

    def timer_begin(self):
        _check(self.cu.cuEventRecord(self.ev0, None), 'cuEventRecord')

    def timer_end(self):
        _check(self.cu.cuEventRecord(self.ev1, None), 'cuEventRecord')
        _check(self.cu.cuEventSynchronize(self.ev1), 'cuEventSynchronize')
        ms = C.c_float()
        _check(self.cu.cuEventElapsedTime(C.byref(ms), self.ev0, self.ev1), 'cuEventElapsedTime')
        return ms.value

    def close(self):
        for b in self.buffers:
            b.free()
        self.buffers = []
        for m in self.modules.values():
            self.cu.cuModuleUnload(m)
        self.modules, self.funcs = {}, {}
        self.cu.cuDevicePrimaryCtxRelease(self.dev)


def chunk_iterator(nelements, nthreads_per_block=64, max_blocks=1024):
    "launch chunks of the reference (chroma/gpu/tools.py:98-119): (first, count, blocks)"
    first = 0
    while first < nelements:
        left = nelements - first
        blocks = min(max_blocks, -(-left // nthreads_per_block))
        n = min(left, blocks * nthreads_per_block)
        yield first, n, blocks
        first += n


def _interp(grid, prop):
    prop = np.asarray(prop)
    return np.interp(grid, prop[:, 0], prop[:, 1]).astype(np.float32)


class RefGeometry(object):
    """Device-side `Geometry` (+ `Detector`) structs of the reference
    (chroma/cuda/geometry_types.h:4-82, chroma/cuda/detector.h:4-23) for a flattened,
    BVH-carrying geometry object."""

    def __init__(self, rctx, geometry, wavelengths=None, times=None):
        self.rctx = c = rctx
        c.bind()
        if wavelengths is None:
            wavelengths = np.arange(60, 1000, 5).astype(np.float32)   # chroma/geometry.py:17
        if times is None:
            times = np.arange(0, 1000, 0.05)
        wstep = float(np.unique(np.diff(wavelengths)).item())
        tstep = 0.05 if times is None else float(times[1] - times[0])
        nw, nt = len(wavelengths), len(times)

        def ptr_array(bufs):
            if not bufs:
                return 0
            return c.to_gpu(np.array([b.ptr for b in bufs], dtype=np.uint64))

        mat_ptrs = []
        for m in geometry.unique_materials:
            if m is None:
                raise Exception('one or more triangles is missing a material.')
            ri = c.to_gpu(_interp(wavelengths, m.refractive_index))
            al = c.to_gpu(_interp(wavelengths, m.absorption_length))
            sl = c.to_gpu(_interp(wavelengths, m.scattering_length))
            prob = ptr_array([c.to_gpu(_interp(wavelengths, x)) for x in m.comp_reemission_prob])
            wcdf = ptr_array([c.to_gpu(_interp(wavelengths, x)) for x in m.comp_reemission_wvl_cdf])
            tcdf = ptr_array([c.to_gpu(_interp(times, x)) for x in m.comp_reemission_time_cdf])
            cabs = ptr_array([c.to_gpu(_interp(wavelengths, x)) for x in m.comp_absorption_length])
            # struct Material, geometry_types.h:4-20
            mat_ptrs.append(c.struct([('P', ri), ('P', al), ('P', sl), ('P', prob), ('P', wcdf),
                                      ('P', tcdf), ('P', cabs),
                                      ('I', len(m.comp_reemission_prob)), ('I', nw), ('f', wstep),
                                      ('f', float(wavelengths[0])), ('I', nt), ('f', tstep),
                                      ('f', float(times[0]))]))
        self.materials = ptr_array(mat_ptrs)

        surf_ptrs = []
        for s in geometry.unique_surfaces:
            if s is None:
                surf_ptrs.append(None)
                continue
            tabs = [c.to_gpu(_interp(wavelengths, getattr(s, name)))
                    for name in ('detect', 'absorb', 'reemit', 'reflect_diffuse', 'reflect_specular',
                                 'eta', 'k', 'reemission_cdf')]
            dich = 0
            if s.dichroic_props:
                p = s.dichroic_props
                ang = c.to_gpu(np.asarray(p.angles, dtype=np.float32))
                refl = ptr_array([c.to_gpu(_interp(wavelengths, x)) for x in p.dichroic_reflect])
                tran = ptr_array([c.to_gpu(_interp(wavelengths, x)) for x in p.dichroic_transmit])
                # struct DichroicProps, geometry_types.h:24-30
                dich = c.struct([('P', ang), ('P', refl), ('P', tran), ('I', len(p.angles))])
            # struct Surface, geometry_types.h:32-50
            surf_ptrs.append(c.struct([('P', t) for t in tabs] + [('P', dich)] +
                                      [('I', int(s.model)), ('I', nw), ('I', int(s.transmissive)),
                                       ('f', wstep), ('f', float(wavelengths[0])),
                                       ('f', float(s.thickness))]))
        self.surfaces = c.to_gpu(np.array([b.ptr if b is not None else 0 for b in surf_ptrs] or [0],
                                          dtype=np.uint64))

        mesh = geometry.mesh
        self.ntriangles = len(mesh.triangles)
        self.vertices = c.to_gpu(np.ascontiguousarray(mesh.vertices, dtype=np.float32))
        self.triangles = c.to_gpu(np.ascontiguousarray(mesh.triangles, dtype=np.uint32))
        codes = (((geometry.inner_material_index & 0xff) << 24) |
                 ((geometry.outer_material_index & 0xff) << 16) |
                 ((geometry.surface_index & 0xff) << 8)).astype(np.uint32)
        self.material_codes = c.to_gpu(codes)
        self.colors = c.to_gpu(np.asarray(geometry.colors).astype(np.uint32))
        self.solid_id_map = c.to_gpu(np.asarray(geometry.solid_id).astype(np.uint32))
        nodes = np.ascontiguousarray(geometry.bvh.nodes).view(np.uint32).reshape(-1, 4)
        self.nnodes = len(nodes)
        self.nodes = c.to_gpu(nodes)
        self.extra_nodes = c.alloc(16)
        wo = [float(x) for x in geometry.bvh.world_coords.world_origin]
        ws = float(geometry.bvh.world_coords.world_scale)
        # struct Geometry, geometry_types.h:69-82 (member order of gpu/geometry.py:243-253)
        self.gpudata = c.struct([('P', self.vertices), ('P', self.triangles), ('P', self.material_codes),
                                 ('P', self.colors), ('P', self.nodes), ('P', self.extra_nodes),
                                 ('P', self.materials), ('P', self.surfaces),
                                 ('f', wo[0]), ('f', wo[1]), ('f', wo[2]), ('f', ws),
                                 ('i', self.nnodes)])
        self.nchannels = 0
        if hasattr(geometry, 'num_channels'):
            self.nchannels = geometry.num_channels()
            s2c = c.to_gpu(np.asarray(geometry.solid_id_to_channel_index).astype(np.int32))
            tx = c.to_gpu(np.asarray(geometry.time_cdf[0]).astype(np.float32))
            ty = c.to_gpu(np.asarray(geometry.time_cdf[1]).astype(np.float32))
            qx = c.to_gpu(np.asarray(geometry.charge_cdf[0]).astype(np.float32))
            qy = c.to_gpu(np.asarray(geometry.charge_cdf[1]).astype(np.float32))
            # struct Detector, detector.h:4-23 (gpu/detector.py:37-46)
            self.detector_gpu = c.struct([('P', s2c), ('P', tx), ('P', ty), ('P', qx), ('P', qy),
                                          ('i', self.nchannels), ('i', len(geometry.time_cdf[0])),
                                          ('i', len(geometry.charge_cdf[0])),
                                          ('f', float(np.float32(geometry.charge_cdf[0][-1] / 2 ** 16)))])

    # -- deterministic tier --------------------------------------------------------------
    def intersect(self, pos, dir, last_hit=None, nthreads_per_block=64):
        """(triangle int32[n], distance float32[n]) from the reference's intersect_mesh
        (oracle/ref_intersect_wrapper.cu); distance is -1 where nothing is hit."""
        c = self.rctx
        c.bind()
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        dir = np.ascontiguousarray(dir, dtype=np.float32)
        n = len(pos)
        if n == 0:
            return np.empty(0, np.int32), np.empty(0, np.float32)
        dpos, ddir = c.to_gpu(pos), c.to_gpu(dir)
        dlast = c.to_gpu(np.ascontiguousarray(last_hit, dtype=np.int32)) if last_hit is not None else np.uint64(0)
        dtri, ddist = c.alloc(4 * n), c.to_gpu(np.full(n, -1.0, np.float32))
        c.launch('intersect', 'intersect_with_index', n // nthreads_per_block + 1, nthreads_per_block,
                 np.int32(n), dpos, ddir, dlast, self.gpudata, dtri, ddist)
        c.synchronize()
        tri, dist = c.get(dtri, np.int32, n), c.get(ddist, np.float32, n)
        for b in (dpos, ddir, dtri, ddist) + ((dlast,) if isinstance(dlast, DeviceBuffer) else ()):
            b.free()
        return tri, dist

    def time_distance_to_mesh(self, pos, dir, reps=5, nthreads_per_block=64):
        "chroma/benchmark.py:22-46: ms per `distance_to_mesh` launch (grid n//64+1 x 64), CUDA events"
        c = self.rctx
        c.bind()
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        dir = np.ascontiguousarray(dir, dtype=np.float32)
        n = len(pos)
        dpos, ddir, ddist = c.to_gpu(pos), c.to_gpu(dir), c.to_gpu(np.full(n, -1.0, np.float32))
        ms = []
        for i in range(reps + 1):
            c.synchronize()
            c.timer_begin()
            c.launch('propagate', 'distance_to_mesh', n // nthreads_per_block + 1, nthreads_per_block,
                     np.int32(n), dpos, ddir, self.gpudata, ddist)
            t = c.timer_end()
            if i > 0:
                ms.append(t)
        dist = c.get(ddist, np.float32, n)
        for b in (dpos, ddir, ddist):
            b.free()
        return ms, dist


class RefRng(object):
    "get_rng_states (chroma/gpu/tools.py:75-84): `size` XORWOW states of 48 bytes"

    def __init__(self, rctx, size=64 * 1024, seed=1, nthreads_per_block=64):
        self.size = size
        rctx.bind()
        self.states = rctx.alloc(48 * size)
        rctx.launch('propagate', 'init_rng', size // nthreads_per_block + 1, nthreads_per_block,
                    np.int32(size), self.states, np.uint64(seed), np.uint64(0))
        rctx.synchronize()


class RefPhotons(object):
    "GPUPhotons of the reference (chroma/gpu/photon.py:13-61): nine AoS device arrays"
    FIELDS = (('pos', np.float32, 3), ('dir', np.float32, 3), ('pol', np.float32, 3),
              ('wavelengths', np.float32, 1), ('t', np.float32, 1),
              ('last_hit_triangles', np.int32, 1), ('flags', np.uint32, 1),
              ('weights', np.float32, 1), ('evidx', np.uint32, 1))

    def __init__(self, rctx, photons):
        self.rctx = rctx
        rctx.bind()
        self.n = len(photons)
        self.host = {}
        for name, dtype, width in self.FIELDS:
            a = np.ascontiguousarray(getattr(photons, name), dtype=dtype)
            self.host[name] = a
            setattr(self, name, rctx.to_gpu(a))

    def reset(self):
        "re-upload the initial state (between timed repetitions)"
        for name, _, _ in self.FIELDS:
            self.rctx.set(getattr(self, name), self.host[name])

    def get(self):
        from chroma_b200 import event
        c, n = self.rctx, self.n
        g = {name: c.get(getattr(self, name), dtype, n * width).reshape((n, 3) if width == 3 else (n,))
             for name, dtype, width in self.FIELDS}
        return event.Photons(g['pos'], g['dir'], g['pol'], g['wavelengths'], g['t'],
                             g['last_hit_triangles'], g['flags'], g['weights'], g['evidx'])

    def free(self):
        for name, _, _ in self.FIELDS:
            getattr(self, name).free()

    def propagate(self, geom, rng, nthreads_per_block=64, max_blocks=1024, max_steps=10,
                  use_weights=False, scatter_first=0):
        """The reference's host step loop, chroma/gpu/photon.py:193-259: queue swap per step,
        one launch per <= 65,536-photon chunk, `nsteps`=1 per launch while the queue is
        >= 8192 photons, all remaining steps otherwise."""
        c = self.rctx
        c.bind()
        n = self.n
        assert rng.size >= nthreads_per_block * max_blocks
        queue = np.empty(n + 1, dtype=np.uint32)
        queue[0] = 0
        queue[1:] = np.arange(n, dtype=np.uint32)
        q_in = c.to_gpu(queue)
        q_out = c.to_gpu(np.concatenate([[1], np.zeros(n, np.uint32)]).astype(np.uint32))
        step, nphotons = 0, n
        while step < max_steps:
            nsteps = max_steps - step if (nphotons < nthreads_per_block * 16 * 8 or use_weights) else 1
            for first, count, blocks in chunk_iterator(nphotons, nthreads_per_block, max_blocks):
                c.launch('propagate', 'propagate', blocks, nthreads_per_block,
                         np.int32(first), np.int32(count), np.uint64(q_in.ptr + 4), q_out, rng.states,
                         self.pos, self.dir, self.wavelengths, self.pol, self.t, self.flags,
                         self.last_hit_triangles, self.weights, self.evidx,
                         np.int32(nsteps), np.int32(int(use_weights)), np.int32(scatter_first),
                         geom.gpudata)
            step += nsteps
            scatter_first = 0
            if step < max_steps:
                q_in, q_out = q_out, q_in
                c.set(q_out, np.ones(1, np.uint32))
                nphotons = int(c.get(q_in, np.uint32, 1)[0]) - 1
                if nphotons == 0:
                    break
        c.synchronize()
        q_in.free()
        q_out.free()


class RefDaq(object):
    """GPUDaq of the reference (chroma/gpu/daq.py:37-99): ndaq=1 -> run_daq (chroma/cuda/daq.cu:35-86),
    ndaq>1 -> run_daq_many (daq.cu:88-150, one block per photon, replica k at channel + k*stride)."""

    def __init__(self, rctx, geom, ndaq=1):
        self.rctx, self.geom, self.ndaq = rctx, geom, int(ndaq)
        self.stride = geom.nchannels
        n = self.n = geom.nchannels * self.ndaq
        self.time_int, self.time = rctx.alloc(4 * n), rctx.alloc(4 * n)
        self.q_int, self.q, self.hist = rctx.alloc(4 * n), rctx.alloc(4 * n), rctx.alloc(4 * n)

    def begin_acquire(self, nthreads_per_block=64):
        c, n = self.rctx, self.n
        c.launch('daq', 'reset_earliest_time_int', n // nthreads_per_block + 1, nthreads_per_block,
                 np.float32(1e9), np.int32(n), self.time_int)
        c.fill32(self.q_int, 0, n)
        c.fill32(self.q, 0, n)
        c.fill32(self.hist, 0, n)

    def acquire(self, photons, rng, nthreads_per_block=64, max_blocks=1024, weight=1.0):
        c = self.rctx
        if self.ndaq == 1:
            for first, count, blocks in chunk_iterator(photons.n, nthreads_per_block, max_blocks):
                c.launch('daq', 'run_daq', blocks, nthreads_per_block, rng.states, np.uint32(0x1 << 2),
                         np.int32(first), np.int32(count), photons.t, photons.flags,
                         photons.last_hit_triangles, photons.weights, self.geom.solid_id_map,
                         self.geom.detector_gpu, self.time_int, self.q_int, self.hist, np.float32(weight))
        else:
            assert rng.size >= nthreads_per_block * max_blocks
            for first, count, blocks in chunk_iterator(photons.n, 1, max_blocks):   # gpu/daq.py:78-79
                c.launch('daq', 'run_daq_many', blocks, nthreads_per_block, rng.states, np.uint32(0x1 << 2),
                         np.int32(first), np.int32(count), photons.t, photons.flags,
                         photons.last_hit_triangles, photons.weights, self.geom.solid_id_map,
                         self.geom.detector_gpu, self.time_int, self.q_int, self.hist,
                         np.int32(self.ndaq), np.int32(self.stride), np.float32(weight))
        c.synchronize()

    def end_acquire(self, nthreads_per_block=64):
        c, n = self.rctx, self.n
        c.launch('daq', 'convert_sortable_int_to_float', n // nthreads_per_block + 1, nthreads_per_block,
                 np.int32(n), self.time_int, self.time)
        c.launch('daq', 'convert_charge_int_to_float', n // nthreads_per_block + 1, nthreads_per_block,
                 self.geom.detector_gpu, self.q_int, self.q)
        c.synchronize()

    def get(self):
        from chroma_b200 import event
        c, n = self.rctx, self.n
        t = c.get(self.time, np.float32, n)
        return event.Channels(t < 1e8, t, c.get(self.q, np.float32, n), c.get(self.hist, np.uint32, n))

    def get_many(self):
        "(t, q, q_int * charge_unit is not applied beyond replica 0 by the reference: q_int returned too, flags), [ndaq, nchannels]"
        c, n, sh = self.rctx, self.n, (self.ndaq, self.stride)
        return (c.get(self.time, np.float32, n).reshape(sh), c.get(self.q, np.float32, n).reshape(sh),
                c.get(self.q_int, np.uint32, n).reshape(sh), c.get(self.hist, np.uint32, n).reshape(sh))


class RefRays(object):
    "chroma/gpu/render.py:8-73 GPURays on the reference's render.cu / transform.cu kernels"

    def __init__(self, rctx, pos, dir, max_alpha_depth=10, nblocks=64):
        self.rctx, self.nblocks, self.max_alpha_depth = rctx, nblocks, max_alpha_depth
        rctx.bind()
        pos = np.ascontiguousarray(pos, dtype=np.float32).reshape(-1, 3)
        dir = np.ascontiguousarray(dir, dtype=np.float32).reshape(-1, 3)
        self.n = len(pos)
        self.pos, self.dir = rctx.to_gpu(pos), rctx.to_gpu(dir)
        self.dx = rctx.alloc(4 * self.n * max_alpha_depth)
        self.color = rctx.alloc(16 * self.n * max_alpha_depth)
        self.dxlen = rctx.to_gpu(np.zeros(self.n, np.uint32))

    def _grid(self):
        return self.n // self.nblocks + 1

    def rotate(self, phi, n):
        c = self.rctx
        for buf in (self.pos, self.dir):
            c.launch('transform', 'rotate', self._grid(), self.nblocks, np.int32(self.n), buf,
                     np.float32(phi), tuple(n))
        c.synchronize()

    def rotate_around_point(self, phi, n, point):
        c = self.rctx
        c.launch('transform', 'rotate_around_point', self._grid(), self.nblocks, np.int32(self.n), self.pos,
                 np.float32(phi), tuple(n), tuple(point))
        c.launch('transform', 'rotate', self._grid(), self.nblocks, np.int32(self.n), self.dir,
                 np.float32(phi), tuple(n))
        c.synchronize()

    def translate(self, v):
        self.rctx.launch('transform', 'translate', self._grid(), self.nblocks, np.int32(self.n), self.pos, tuple(v))
        self.rctx.synchronize()

    def get(self):
        c = self.rctx
        return (c.get(self.pos, np.float32, 3 * self.n).reshape(-1, 3),
                c.get(self.dir, np.float32, 3 * self.n).reshape(-1, 3))

    def render(self, geom, alpha_depth=10, keep_last_render=False, bg_color=0):
        "pixels uint32[n] (gpu/render.py:43-66)"
        c = self.rctx
        c.bind()
        if not keep_last_render:
            c.fill32(self.dxlen, 0, self.n)
        pixels = c.alloc(4 * self.n)
        c.launch('render', 'render', self._grid(), self.nblocks, np.int32(self.n), self.pos, self.dir,
                 geom.gpudata, np.uint32(alpha_depth), pixels, self.dx, self.dxlen, self.color,
                 np.uint32(bg_color))
        c.synchronize()
        out = c.get(pixels, np.uint32, self.n)
        pixels.free()
        return out

    def lists(self, alpha_depth):
        "(dxlen[n], dx[n, alpha_depth], color[n, alpha_depth, 4]) of the last render"
        c = self.rctx
        return (c.get(self.dxlen, np.uint32, self.n),
                c.get(self.dx, np.float32, self.n * alpha_depth).reshape(self.n, alpha_depth),
                c.get(self.color, np.float32, self.n * alpha_depth * 4).reshape(self.n, alpha_depth, 4))

    def time_render(self, geom, alpha_depth=10, reps=3):
        c = self.rctx
        pixels = c.alloc(4 * self.n)
        best = None
        for _ in range(reps):
            c.fill32(self.dxlen, 0, self.n)
            c.synchronize()
            c.timer_begin()
            c.launch('render', 'render', self._grid(), self.nblocks, np.int32(self.n), self.pos, self.dir,
                     geom.gpudata, np.uint32(alpha_depth), pixels, self.dx, self.dxlen, self.color, np.uint32(0))
            ms = c.timer_end()
            best = ms if best is None else min(best, ms)
        pixels.free()
        return best


def make_leaves(rctx, mesh, world_origin, world_scale, nthreads_per_block=256):
    "chroma/gpu/bvh.py:41-78 create_leaf_nodes -> (uint32[n][4] leaf nodes, uint64[n] Morton codes)"
    c = rctx
    c.bind()
    tri = np.ascontiguousarray(mesh.triangles, dtype=np.uint32)
    vert = np.ascontiguousarray(mesh.vertices, dtype=np.float32)
    n = len(tri)
    dtri, dvert = c.to_gpu(tri), c.to_gpu(vert)
    dnodes, dcodes = c.alloc(16 * n), c.alloc(8 * n)
    wo = [float(x) for x in world_origin]
    # float3 by value = three consecutive floats in the parameter buffer (12 bytes, 4-aligned)
    f = c.function('bvh', 'make_leaves')
    args = [C.c_uint32(0), C.c_uint32(n), C.c_uint64(dtri.ptr), C.c_uint64(dvert.ptr),
            (C.c_float * 3)(*wo), C.c_float(float(world_scale)), C.c_uint64(dnodes.ptr),
            C.c_uint64(dcodes.ptr)]
    arr = (C.c_void_p * len(args))(*[C.cast(C.byref(a), C.c_void_p) for a in args])
    _check(c.cu.cuLaunchKernel(f, -(-n // nthreads_per_block), 1, 1, nthreads_per_block, 1, 1, 0, None,
                               arr, None), 'cuLaunchKernel(make_leaves)')
    c.synchronize()
    nodes, codes = c.get(dnodes, np.uint32, 4 * n).reshape(n, 4), c.get(dcodes, np.uint64, n)
    for b in (dtri, dvert, dnodes, dcodes):
        b.free()
    return nodes, codes


# -- the kernels chroma/bvh/grid.py drives through chroma/gpu/bvh.py, on the reference's bvh.cubin --
def merge_nodes_detailed(rctx, nodes, first_child, nchild, nthreads_per_block=256):
    "chroma/gpu/bvh.py:81-112 on make_parents_detailed (bvh.cu:269-308): uint32[nparent][4]"
    c = rctx
    c.bind()
    nodes = np.ascontiguousarray(nodes).view(np.uint32).reshape(-1, 4)
    dn = c.to_gpu(nodes)
    dfc = c.to_gpu(np.ascontiguousarray(first_child, dtype=np.int32))
    dnc = c.to_gpu(np.ascontiguousarray(nchild, dtype=np.int32))
    nparent = len(first_child)
    dp = c.alloc(16 * nparent)
    for first, count, blocks in chunk_iterator(nparent, nthreads_per_block, max_blocks=10000):
        c.launch('bvh', 'make_parents_detailed', blocks, nthreads_per_block, np.uint32(first),
                 np.uint32(count), dn, dp, dfc, dnc)
    c.synchronize()
    out = c.get(dp, np.uint32, 4 * nparent).reshape(nparent, 4)
    for b in (dn, dfc, dnc, dp):
        b.free()
    return out


def concatenate_layers(rctx, layers, nthreads_per_block=256):
    "chroma/gpu/bvh.py:239-267 on copy_and_offset (bvh.cu:364-384): (uint32[n][4], layer_bounds)"
    c = rctx
    c.bind()
    bounds = np.insert(np.cumsum([len(x) for x in layers]), 0, 0)
    dnodes = c.alloc(16 * int(bounds[-1]))
    for start, end, layer in zip(bounds[:-1], bounds[1:], layers):
        src = c.to_gpu(np.ascontiguousarray(layer).view(np.uint32).reshape(-1, 4))
        offset = 0 if end == bounds[-1] else int(end)
        for first, count, blocks in chunk_iterator(int(end - start), nthreads_per_block, max_blocks=10000):
            c.launch('bvh', 'copy_and_offset', blocks, nthreads_per_block, np.uint32(first), np.uint32(count),
                     np.uint32(offset), src, np.uint64(dnodes.ptr + 16 * int(start)))
        c.synchronize()
        src.free()
    out = c.get(dnodes, np.uint32, 4 * int(bounds[-1])).reshape(-1, 4)
    dnodes.free()
    return out, bounds


def collapse_chains(rctx, nodes, layer_bounds, nthreads_per_block=256):
    "chroma/gpu/bvh.py:114-130 on collapse_child (bvh.cu:530-543), grid (120,1) as the reference"
    c = rctx
    c.bind()
    nodes = np.ascontiguousarray(nodes).view(np.uint32).reshape(-1, 4)
    dn = c.to_gpu(nodes)
    bounds = list(zip(layer_bounds[:-1], layer_bounds[1:]))[:-1]
    bounds.reverse()
    for start, end in bounds:
        c.launch('bvh', 'collapse_child', 120, nthreads_per_block, np.uint32(start), np.uint32(end), dn)
    c.synchronize()
    out = c.get(dn, np.uint32, nodes.size).reshape(-1, 4)
    dn.free()
    return out
