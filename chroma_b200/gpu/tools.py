"""Context, device arrays and launch helpers (API of ``chroma/gpu/tools.py``), backed by
the C-ABI library instead of PyCUDA."""
import ctypes as C

import numpy as np

from chroma_b200 import _lib

# pycuda.gpuarray.vec look-alikes (structured dtypes with the CUDA vector layout)
float3 = np.dtype([('x', np.float32), ('y', np.float32), ('z', np.float32)])
uint3 = np.dtype([('x', np.uint32), ('y', np.uint32), ('z', np.uint32)])
uint4 = np.dtype([('x', np.uint32), ('y', np.uint32), ('z', np.uint32), ('w', np.uint32)])
float4 = np.dtype([('x', np.float32), ('y', np.float32), ('z', np.float32), ('w', np.float32)])


class vec(object):
    float3 = float3
    uint3 = uint3
    uint4 = uint4

    @staticmethod
    def make_float3(x, y, z):
        return np.array((x, y, z), dtype=float3)


_current = []


class Context(object):
    "One CUDA device + stream (chroma/gpu/tools.py:121-142 returns a pycuda context)."

    def __init__(self, device_id=None):
        handle = C.c_void_p()
        _lib.call('chroma_ctx_create', -1 if device_id is None else int(device_id),
                  C.byref(handle))
        self.handle = handle
        self.device_id = device_id
        _current.append(self)

    def synchronize(self):
        _lib.call('chroma_ctx_synchronize', self.handle)

    def info(self):
        out = (C.c_uint64 * 6)()
        _lib.call('chroma_ctx_info', self.handle, out)
        return dict(free=out[0], total=out[1], sm_count=out[2], l2_bytes=out[3],
                    l2_persist_max=out[4], cc=out[5])

    def launch_count(self):
        out = C.c_uint64()
        _lib.call('chroma_ctx_launch_count', self.handle, C.byref(out))
        return out.value

    def timer_begin(self):
        _lib.call('chroma_ctx_timer_begin', self.handle)

    def timer_end(self):
        ms = C.c_float()
        _lib.call('chroma_ctx_timer_end', self.handle, C.byref(ms))
        return ms.value

    def flush_l2(self):
        _lib.call('chroma_ctx_flush_l2', self.handle)

    def stream_ptr(self):
        "cudaStream_t of this context as an integer (torch.cuda.ExternalStream, NCCL ...)"
        p = C.c_void_p()
        _lib.call('chroma_ctx_stream', self.handle, C.byref(p))
        return int(p.value or 0)

    def reload_knobs(self):
        """Re-read the CHROMA_* tuning variables (they are read once, when the context is created)."""
        _lib.call('chroma_ctx_reload_knobs', self.handle)

    def pop(self):
        "Detach from the context stack and release the device resources."
        if self in _current:
            _current.remove(self)
        if self.handle is not None:
            _lib.call('chroma_ctx_destroy', self.handle)
            self.handle = None

    # pycuda spells it this way in a few callers
    detach = pop


def create_cuda_context(device_id=None):
    "Initialise and return a context on `device_id` (default device if None)."
    return Context(device_id)


def current_context():
    if not _current:
        raise _lib.ChromaError(-3, 'no context: call chroma_b200.gpu.create_cuda_context() first')
    return _current[-1]


class DeviceArray(object):
    """Minimal stand-in for pycuda.gpuarray.GPUArray: device pointer + shape + dtype with
    get/set/fill, 1-D slicing into non-owning views and __cuda_array_interface__."""

    def __init__(self, shape, dtype, ctx=None, ptr=None, base=None):
        self.ctx = ctx or current_context()
        self.shape = (int(shape),) if np.isscalar(shape) else tuple(int(s) for s in shape)
        self.dtype = np.dtype(dtype)
        self.size = int(np.prod(self.shape)) if len(self.shape) else 1
        self.nbytes = self.size * self.dtype.itemsize
        self.base = base
        if ptr is None:
            p = C.c_void_p()
            _lib.call('chroma_malloc', self.ctx.handle, self.nbytes, C.byref(p))
            self.ptr = p.value
            self._owner = True
        else:
            self.ptr = int(ptr) if ptr else 0
            self._owner = False

    @property
    def gpudata(self):
        return self.ptr

    def __len__(self):
        return self.shape[0] if self.shape else 1

    def get(self):
        out = np.empty(self.shape, dtype=self.dtype)
        if self.nbytes:
            _lib.call('chroma_memcpy_d2h', self.ctx.handle, _lib.ptr(out), C.c_void_p(self.ptr),
                      self.nbytes)
        return out

    def set(self, array):
        array = np.ascontiguousarray(array, dtype=self.dtype)
        if array.size != self.size:
            raise ValueError('size mismatch in DeviceArray.set')
        if self.nbytes:
            _lib.call('chroma_memcpy_h2d', self.ctx.handle, C.c_void_p(self.ptr), _lib.ptr(array),
                      self.nbytes)
        return self

    def fill(self, value):
        if self.dtype.itemsize != 4:
            raise TypeError('fill() supports 4-byte element types only')
        word = np.array([value], dtype=self.dtype).view(np.uint32)[0]
        _lib.call('chroma_memset32', self.ctx.handle, C.c_void_p(self.ptr), int(word), self.size)
        return self

    def __getitem__(self, key):
        if not isinstance(key, slice):
            raise TypeError('DeviceArray supports slice indexing only')
        start, stop, step = key.indices(len(self))
        if step != 1:
            raise ValueError('DeviceArray slices must be contiguous')
        n = max(0, stop - start)
        row = self.dtype.itemsize * int(np.prod(self.shape[1:])) if len(self.shape) > 1 \
            else self.dtype.itemsize
        return DeviceArray((n,) + self.shape[1:], self.dtype, ctx=self.ctx,
                           ptr=self.ptr + start * row, base=self)

    @property
    def __cuda_array_interface__(self):
        if self.dtype.names:
            raise TypeError('vector dtypes have no __cuda_array_interface__')
        return dict(shape=self.shape, typestr=self.dtype.str, data=(self.ptr, False), version=2,
                    strides=None)

    def __del__(self):
        try:
            if self._owner and self.ptr and self.ctx.handle is not None:
                _lib.call('chroma_free', self.ctx.handle, C.c_void_p(self.ptr))
        except Exception:
            pass


class _PinnedBlock(object):
    "page-locked host allocation (chroma_host_alloc); freed when the last array view dies"

    def __init__(self, nbytes):
        p = C.c_void_p()
        _lib.call('chroma_host_alloc', max(int(nbytes), 16), C.byref(p))
        self.ptr, self.nbytes = p.value, int(nbytes)
        self.buf = (C.c_char * max(int(nbytes), 16)).from_address(self.ptr)

    def __del__(self):
        try:
            if self.ptr:
                _lib.call('chroma_host_free', C.c_void_p(self.ptr))
                self.ptr = None
        except Exception:
            pass


def pinned_empty(shape, dtype):
    """numpy array in page-locked host memory.  Photons built from such arrays are uploaded by
    DMA at full PCIe rate and the upload is asynchronous: do not modify them between
    ``GPUPhotons(photons)`` and the next synchronising call (``propagate``, ``get`` ...)."""
    dtype = np.dtype(dtype)
    shape = (int(shape),) if np.isscalar(shape) else tuple(int(x) for x in shape)
    n = int(np.prod(shape)) if len(shape) else 1
    block = _PinnedBlock(n * dtype.itemsize)
    arr = np.frombuffer(block.buf, dtype=dtype, count=n).reshape(shape)
    _PINNED[id(block)] = block
    import weakref
    weakref.finalize(arr, _PINNED.pop, id(block), None)
    return arr


_PINNED = {}


def pinned_copy(array):
    "page-locked copy of `array` (see pinned_empty)"
    array = np.asarray(array)
    out = pinned_empty(array.shape, array.dtype)
    out[...] = array
    return out


def to_gpu(array, ctx=None):
    array = np.ascontiguousarray(array)
    return DeviceArray(array.shape, array.dtype, ctx=ctx).set(array)


def empty(shape, dtype, ctx=None):
    return DeviceArray(shape, dtype, ctx=ctx)


def zeros(shape, dtype, ctx=None):
    return DeviceArray(shape, dtype, ctx=ctx).fill(0)


class RNGStates(object):
    """Handle returned by get_rng_states.  The reference allocates `size` XORWOW states
    (chroma/gpu/tools.py:75-84); the engine's generator is counter-based, so this holds a
    seed and a call counter only."""

    def __init__(self, size, seed, ctx=None):
        self.ctx = ctx or current_context()
        self.size = int(size)
        self.seed = int(seed)
        h = C.c_void_p()
        _lib.call('chroma_rng_create', self.ctx.handle, self.size, self.seed & (2 ** 64 - 1),
                  C.byref(h))
        self.handle = h

    @property
    def epoch(self):
        "call counter: counter word 2 of the streams the next propagate / acquire call will use"
        e = C.c_uint32()
        _lib.call('chroma_rng_get_epoch', self.handle, C.byref(e))
        return int(e.value)

    @epoch.setter
    def epoch(self, value):
        _lib.call('chroma_rng_set_epoch', self.handle, int(value) & 0xFFFFFFFF)

    def fill_uniform(self, n):
        "n floats in (0, 1] on the device (test hook; cf. random.h:72-83 fill_uniform)."
        out = DeviceArray(n, np.float32, ctx=self.ctx)
        _lib.call('chroma_rng_fill_uniform', self.handle, n, C.c_void_p(out.ptr))
        return out

    def __del__(self):
        try:
            if self.handle is not None:
                _lib.call('chroma_rng_destroy', self.handle)
        except Exception:
            pass


def get_rng_states(size, seed=1):
    "Return `size` number of CUDA random number generator states."
    return RNGStates(size, seed)


def to_float3(arr):
    "(N,3) array -> float3 structured array."
    arr = np.ascontiguousarray(arr, dtype=np.float32)
    return arr.view(float3)[:, 0]


def to_uint3(arr):
    "(N,3) array -> uint3 structured array."
    arr = np.ascontiguousarray(arr, dtype=np.uint32)
    return arr.view(uint3)[:, 0]


def chunk_iterator(nelements, nthreads_per_block=64, max_blocks=1024):
    """Yield (first_index, elements_this_iteration, nblocks_this_iteration) covering
    `nelements` with at most `max_blocks` blocks per pass (chroma/gpu/tools.py:98-119).

        >>> list(chunk_iterator(300, 32, 2))
        [(0, 64, 2), (64, 64, 2), (128, 64, 2), (192, 64, 2), (256, 44, 2)]
    """
    first = 0
    while first < nelements:
        left = nelements - first
        blocks = min(max_blocks, -(-left // nthreads_per_block))
        this_round = min(left, blocks * nthreads_per_block)
        yield (first, this_round, blocks)
        first += this_round


def format_size(size):
    for limit, unit, div in ((1e3, ' ', 1.0), (1e6, 'K', 1e3), (1e9, 'M', 1e6)):
        if size < limit:
            return '%.1f%s' % (size / div, unit)
    return '%.1f%s' % (size / 1e9, 'G')


def format_array(name, array):
    return '%-15s %6s %6s' % (name, format_size(len(array)), format_size(array.nbytes))


_NO_JIT = ('chroma_b200 does not JIT-compile CUDA source: kernels live in libchroma_b200.so. '
           'Use the GPU* classes, or chroma_b200.gpu.intersect() for distance_to_mesh.')


def get_cu_module(name, options=None, include_source_directory=True):
    raise NotImplementedError(_NO_JIT)


def get_cu_source(name):
    raise NotImplementedError(_NO_JIT)


class GPUFuncs(object):
    def __init__(self, module):
        raise NotImplementedError(_NO_JIT)


cuda_options = ('--use_fast_math',)
