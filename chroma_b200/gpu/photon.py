"""GPUPhotons: photons resident on the device and the propagate step loop
(API of ``chroma/gpu/photon.py:13-381``)."""
import ctypes as C
import sys

import numpy as np

from chroma_b200 import _lib, event
from chroma_b200.gpu.tools import DeviceArray, current_context
from chroma_b200.tools import profile_if_possible

# field name -> (position in chroma_photons_download's argument list, dtype, columns)
_FIELD_SPEC = {
    'pos': (0, np.float32, 3), 'dir': (1, np.float32, 3), 'pol': (2, np.float32, 3),
    'wavelengths': (3, np.float32, 1), 't': (4, np.float32, 1),
    'last_hit_triangles': (5, np.int32, 1), 'flags': (6, np.uint32, 1),
    'weights': (7, np.float32, 1), 'evidx': (8, np.uint32, 1), 'channel': (9, np.int32, 1),
}


class PhotonField(object):
    """One attribute of a GPUPhotons (``.pos``, ``.flags`` ...).  The device keeps photons as
    packed 16-byte records, so this is a (container, field, window) view that behaves like
    the GPUArray the reference exposes: ``get()``, ``set()``, ``fill()``, ``size``, ``len()``, slicing,
    and ``to_device()`` / ``gpudata`` for a contiguous device copy to hand to other kernels."""

    def __init__(self, owner, name, offset, n):
        self.owner, self.name, self.offset, self.n = owner, name, int(offset), int(n)
        _, self.dtype, self.cols = _FIELD_SPEC[name]
        self.size = self.n
        self.shape = (self.n,)

    def __len__(self):
        return self.n

    def get(self):
        pos, dtype, cols = _FIELD_SPEC[self.name]
        out = np.empty((self.n, 3) if cols == 3 else self.n, dtype=dtype)
        args = [None] * 10
        args[pos] = _lib.ptr(out)
        _lib.call('chroma_photons_download', self.owner.handle, self.offset, self.n, *args)
        return out

    def to_device(self):
        """The field as a contiguous DeviceArray ((n,3) float32 for pos/dir/pol): a device-side COPY of
        the packed records, for callers that pass ``gpu_photons.pos`` to their own kernels
        (``.gpudata`` / ``__cuda_array_interface__``).  Write it back with ``set()``."""
        pos, dtype, cols = _FIELD_SPEC[self.name]
        out = DeviceArray((self.n, 3) if cols == 3 else self.n, dtype, ctx=self.owner.ctx)
        _lib.call('chroma_photons_field_get', self.owner.handle, self.offset, self.n, pos, C.c_void_p(out.ptr))
        return out

    @property
    def gpudata(self):
        "device pointer of a contiguous snapshot of the field (kept alive by this view)"
        self._snapshot = self.to_device()
        return self._snapshot.ptr

    def set(self, array):
        "overwrite the field from a host array or a DeviceArray of the field's shape (GPUArray.set)"
        pos, dtype, cols = _FIELD_SPEC[self.name]
        if not isinstance(array, DeviceArray):
            host = np.ascontiguousarray(array, dtype=dtype)
            if host.size != self.n * cols:
                raise ValueError('size mismatch in PhotonField.set')
            array = DeviceArray(host.shape, dtype, ctx=self.owner.ctx).set(host)
        elif array.size != self.n * cols:
            raise ValueError('size mismatch in PhotonField.set')
        _lib.call('chroma_photons_field_set', self.owner.handle, self.offset, self.n, pos, C.c_void_p(array.ptr))
        return self

    def fill(self, value):
        pos, dtype, cols = _FIELD_SPEC[self.name]
        return self.set(np.full((self.n, 3) if cols == 3 else self.n, value, dtype=dtype))

    def __getitem__(self, key):
        if not isinstance(key, slice):
            raise TypeError('PhotonField supports slice indexing only')
        start, stop, step = key.indices(self.n)
        if step != 1:
            raise ValueError('PhotonField slices must be contiguous')
        return PhotonField(self.owner, self.name, self.offset + start, max(0, stop - start))


class _Handle(object):
    "Owns a chroma_photons container."

    def __init__(self, ctx, handle, capacity):
        self.ctx, self.handle, self.capacity = ctx, handle, capacity
        self.host_arrays = []   # page-locked upload sources are read asynchronously: kept alive here

    def __del__(self):
        try:
            if self.handle is not None and self.ctx.handle is not None:
                _lib.call('chroma_photons_destroy', self.handle)
                self.handle = None
        except Exception:
            pass


def _new_container(ctx, capacity, id_base=0):
    h = C.c_void_p()
    _lib.call('chroma_photons_create', ctx.handle, int(capacity), int(id_base), C.byref(h))
    return _Handle(ctx, h, int(capacity))


def _wrap_result(ctx, raw):
    cap = C.c_uint64()
    _lib.call('chroma_photons_size', raw, C.byref(cap))
    return _Handle(ctx, raw, cap.value)


def generate_source(owner, offset, n, source):
    """Sample `n` photons of `source` (event.PhotonSource) into slots [offset, offset+n) of a
    container (chroma_photons_generate).  A photon's random stream is keyed by (seed, sequence,
    id_base + slot): a caller that splits one source over several GPUs gives each container the
    `id_base` of the first photon it holds and gets the same photons as a single GPU would."""
    desc = _lib.SourceDesc()
    desc.pos[:] = [float(x) for x in source.pos]
    desc.rot[:] = [float(x) for x in np.asarray(source.rotation(), dtype=np.float32).ravel()]
    desc.cos_half_angle = source.cos_half_angle
    desc.wavelength_lo, desc.wavelength_hi = source.wavelength_lo, source.wavelength_hi
    desc.t0, desc.weight = source.t0, source.weight
    desc.evidx, desc.sequence = int(source.evidx), source.sequence & 0xFFFFFFFF
    _lib.call('chroma_photons_generate', owner.handle, int(offset), int(n), C.byref(desc),
              int(source.seed))


def _is_broadcast(a):
    "a numpy view that repeats one row (np.broadcast_to): the whole field is one value"
    return isinstance(a, np.ndarray) and a.ndim >= 1 and len(a) > 0 and a.strides[0] == 0


def upload_photons(owner, offset, photons, copy_flags=True, copy_triangles=True, copy_weights=True, evidx=None,
                   borrow=False):
    """H2D copy of an event.Photons into slots [offset, offset+n) of `owner`.  Fields that are one value
    for the whole range are passed BY VALUE and not uploaded (chroma_photons_upload_arrays): `pos` / `t`
    given as np.broadcast_to views (a photon bomb has one vertex and one start time), the defaults of
    copy_*=False (chroma/gpu/photon.py:36-45), any other field given as a broadcast view, and `evidx`
    when given as a number (one event per window, chroma/sim.py:62-72).  Returns the arrays that must stay alive while a
    page-locked source is still being read.  `borrow=True`: the caller promises not to modify PAGEABLE source
    arrays either until the next synchronising call; they are then staged by a worker thread while the caller
    goes on to launch propagate (chroma_photon_arrays.borrow_sources)."""
    n = len(photons)
    a = _lib.PhotonArrays()
    keep = []

    def arr(x, dtype):
        x = _lib.carray(x, dtype)
        keep.append(x)
        return _lib.ptr(x)
    if _is_broadcast(photons.pos):
        a.pos, a.pos_value = None, (C.c_float * 3)(*[float(v) for v in np.asarray(photons.pos[0])])
    else:
        a.pos = arr(photons.pos, np.float32)
    a.dir, a.pol = arr(photons.dir, np.float32), arr(photons.pol, np.float32)
    a.wavelengths = arr(photons.wavelengths, np.float32)
    if _is_broadcast(photons.t):
        a.t, a.t_value = None, float(photons.t[0])
    else:
        a.t = arr(photons.t, np.float32)
    a.weight_value, a.last_hit_value, a.flags_value = 1.0, -1, 0
    if copy_weights and not _is_broadcast(photons.weights):
        a.weights = arr(photons.weights, np.float32)
    elif copy_weights:
        a.weight_value = float(photons.weights[0])
    if copy_triangles and not _is_broadcast(photons.last_hit_triangles):
        a.last_hit = arr(photons.last_hit_triangles, np.int32)
    elif copy_triangles:
        a.last_hit_value = int(photons.last_hit_triangles[0])
    # no scan of the arrays here: looking at 1e7 flags costs more host time than their 40 MB cost on PCIe
    if copy_flags:
        if _is_broadcast(photons.flags):
            a.flags_value = int(photons.flags[0])
        else:
            a.flags = arr(photons.flags, np.uint32)
    if evidx is not None:
        a.evidx_value = int(evidx)
    elif _is_broadcast(photons.evidx):
        a.evidx_value = int(photons.evidx[0])
    else:
        a.evidx = arr(photons.evidx, np.uint32)
    a.borrow_sources = 1 if borrow else 0
    _lib.call('chroma_photons_upload_arrays', owner.handle, int(offset), int(n), C.byref(a))
    return keep


class GPUPhotons(object):
    def __init__(self, photons, ncopies=1, copy_flags=True, copy_triangles=True,
                 copy_weights=True, id_base=0, borrow=False):
        """Load ``photons`` (chroma_b200.event.Photons) onto the GPU, replicated `ncopies`
        times (copy k occupies [k*N, (k+1)*N)).  copy_* = False keeps the defaults
        last_hit_triangles=-1 / flags=0 / weights=1 (chroma/gpu/photon.py:36-45).
        ``id_base`` is the global index of photon 0: it keys the counter-based RNG so results
        do not depend on how photons are split over GPUs.  ``borrow=True``: the caller leaves the
        host arrays untouched until the next synchronising call (.get(), Context.synchronize()); pageable
        arrays are then staged in the background and propagate() can start at once."""
        self.ctx = current_context()
        nphotons = len(photons)
        self._owner = _new_container(self.ctx, nphotons * ncopies, id_base)
        self._offset, self._n = 0, nphotons * ncopies
        if isinstance(photons, event.PhotonSource):
            # sampled on the device: no host arrays, no H2D copy
            if nphotons:
                generate_source(self._owner, 0, nphotons, photons)
            if ncopies > 1 and nphotons:
                _lib.call('chroma_photons_duplicate', self._owner.handle, nphotons, ncopies)
        elif nphotons:
            # page-locked sources are read asynchronously: keep them alive
            self._host_arrays = upload_photons(self._owner, 0, photons, copy_flags, copy_triangles, copy_weights,
                                               borrow=borrow)
            if ncopies > 1:
                _lib.call('chroma_photons_duplicate', self._owner.handle, nphotons, ncopies)
        self.true_nphotons = nphotons
        self.ncopies = ncopies
        self._make_fields()

    @classmethod
    def empty(cls, nphotons, id_base=0):
        "A container of `nphotons` unset photons, to be filled window by window with load()."
        ctx = current_context()
        return cls._from_owner(ctx, _new_container(ctx, int(nphotons), id_base), 0, int(nphotons))

    def load(self, offset, photons, evidx=None, borrow=False):
        """Fill slots [offset, offset+len(photons)) from an event.Photons (H2D copy; last_hit_triangles -1
        and weights 1 as chroma/sim.py:70-72 loads a batch, flags kept) or an event.PhotonSource
        (sampled on the device).  `evidx`: event index of the whole window, passed by value instead of
        the photons' own evidx array (which is then neither read nor modified)."""
        n = len(photons)
        assert 0 <= offset and offset + n <= self._n
        if n == 0:
            return
        if isinstance(photons, event.PhotonSource):
            if evidx is not None:
                photons.evidx = int(evidx)
            generate_source(self._owner, self._offset + offset, n, photons)
            return
        keep = upload_photons(self._owner, self._offset + offset, photons, copy_flags=True,
                              copy_triangles=False, copy_weights=False, evidx=evidx, borrow=borrow)
        # page-locked sources are read asynchronously: keep them alive with the container
        self._owner.host_arrays.append(keep)

    # -- plumbing ---------------------------------------------------------------------
    @property
    def handle(self):
        return self._owner.handle

    def _make_fields(self):
        for name in ('pos', 'dir', 'pol', 'wavelengths', 't', 'last_hit_triangles', 'flags',
                     'weights', 'evidx'):
            setattr(self, name, PhotonField(self._owner, name, self._offset, self._n))

    @classmethod
    def _from_owner(cls, ctx, owner, offset, n):
        self = cls.__new__(cls)
        self.ctx, self._owner, self._offset, self._n = ctx, owner, offset, n
        self.true_nphotons, self.ncopies = n, 1
        self._make_fields()
        return self

    def __len__(self):
        return self._n

    # -- host copies ------------------------------------------------------------------
    def _download(self, offset, n, with_channel=False):
        out = [np.empty((n, 3), np.float32), np.empty((n, 3), np.float32),
               np.empty((n, 3), np.float32), np.empty(n, np.float32), np.empty(n, np.float32),
               np.empty(n, np.int32), np.empty(n, np.uint32), np.empty(n, np.float32),
               np.empty(n, np.uint32), np.empty(n, np.int32) if with_channel else None]
        _lib.call('chroma_photons_download', self._owner.handle, offset, n,
                  *[_lib.ptr(a) for a in out])
        return event.Photons(*out)

    def get(self):
        return self._download(self._offset, self._n)

    def get_flat_hits(self, gpu_detector, target_flag=(0x1 << 2), nthreads_per_block=64,
                      max_blocks=1024, start_photon=None, nphotons=None, no_map=False):
        """Photons with `target_flag` set that ended on a triangle whose solid maps to a
        channel; ``.channel`` holds the channel index.  Order: ascending photon index (the
        reference's order is unspecified, chroma/cuda/propagate.cu:199)."""
        start_photon = 0 if start_photon is None else start_photon
        if nphotons is None:
            nphotons = self._n - start_photon
        raw = C.c_void_p()
        _lib.call('chroma_photons_flat_hits', self._owner.handle, gpu_detector.handle,
                  int(target_flag), self._offset + start_photon, nphotons, C.byref(raw))
        owner = _wrap_result(self.ctx, raw)
        hits = GPUPhotons._from_owner(self.ctx, owner, 0, owner.capacity)
        return hits._download(0, owner.capacity, with_channel=True)

    def get_hits(self, *args, **kwargs):
        "Map channel -> Photons detected on it."
        flat_hits = self.get_flat_hits(*args, **kwargs)
        return {int(ch): flat_hits[flat_hits.channel == ch] for ch in np.unique(flat_hits.channel)}

    def iterate_copies(self):
        "Yield a GPUPhotonsSlice per replica."
        for i in range(self.ncopies):
            w = slice(self.true_nphotons * i, self.true_nphotons * (i + 1))
            yield GPUPhotonsSlice(pos=self.pos[w], dir=self.dir[w], pol=self.pol[w],
                                  wavelengths=self.wavelengths[w], t=self.t[w],
                                  last_hit_triangles=self.last_hit_triangles[w],
                                  flags=self.flags[w], weights=self.weights[w],
                                  evidx=self.evidx[w])

    # -- the step loop ----------------------------------------------------------------
    @profile_if_possible
    def propagate(self, gpu_geometry, rng_states, nthreads_per_block=64, max_blocks=1024,
                  max_steps=10, use_weights=False, scatter_first=0, track=False):
        """Propagate photons to termination or `max_steps`, whichever comes first.  May be
        called repeatedly (terminal photons are skipped).  `nthreads_per_block` and
        `max_blocks` are accepted for compatibility: the engine runs one persistent launch
        sized to the device.  With ``track=True`` returns (step_photon_ids, step_photons) as
        chroma/gpu/photon.py:218-238 does."""
        if not track:
            self._launch(gpu_geometry, rng_states, max_steps, use_weights, scatter_first)
            return None
        ids = np.arange(self._n, dtype=np.uint32)
        if self.ncopies > 1:  # clones next to each other (chroma/gpu/photon.py:211-212)
            ids = ids.reshape(self.ncopies, self.true_nphotons).T.ravel().copy()
        step_photon_ids = [ids]
        step_photons = [self.copy_queue(ids, len(ids)).get()]
        for _ in range(max_steps):
            self._launch(gpu_geometry, rng_states, 1, use_weights, scatter_first)
            scatter_first = 0
            step_photon_ids.append(ids)
            step_photons.append(self.copy_queue(ids, len(ids)).get())
            ids = self._live_ids()
            if len(ids) == 0:
                break
        return step_photon_ids, step_photons

    def _launch(self, gpu_geometry, rng_states, max_steps, use_weights, scatter_first):
        nan = C.c_uint64()
        _lib.call('chroma_photons_propagate', self._owner.handle, gpu_geometry.handle,
                  rng_states.handle, self._offset, self._n, int(max_steps), int(bool(use_weights)),
                  int(scatter_first), C.byref(nan))
        if nan.value:
            print('WARNING: ABORTED PHOTONS', file=sys.stderr)

    def _live_ids(self):
        ids = np.empty(self._n, dtype=np.uint32)
        count = C.c_uint64(self._n)
        _lib.call('chroma_photons_live_ids', self._owner.handle, self._offset, self._n,
                  _lib.ptr(ids), C.byref(count))
        return ids[:count.value] - np.uint32(self._offset)

    @profile_if_possible
    def copy_queue(self, queue_gpu, nphotons, nthreads_per_block=64, max_blocks=1024,
                   start_photon=0):
        "Gather photons by the first `nphotons` indices of `queue_gpu` (array or DeviceArray)."
        ids = queue_gpu.get() if isinstance(queue_gpu, DeviceArray) else np.asarray(queue_gpu)
        ids = np.ascontiguousarray(ids[start_photon:start_photon + nphotons], dtype=np.uint32) \
            + np.uint32(self._offset)
        raw = C.c_void_p()
        _lib.call('chroma_photons_gather', self._owner.handle, _lib.ptr(ids), len(ids), C.byref(raw))
        owner = _wrap_result(self.ctx, raw)
        return GPUPhotons._from_owner(self.ctx, owner, 0, owner.capacity)

    @profile_if_possible
    def select(self, target_flag, nthreads_per_block=64, max_blocks=1024, start_photon=None,
               nphotons=None):
        "New GPU photon list with only the photons that have `target_flag` set."
        start_photon = 0 if start_photon is None else start_photon
        if nphotons is None:
            nphotons = self._n - start_photon
        raw = C.c_void_p()
        _lib.call('chroma_photons_select', self._owner.handle, int(target_flag),
                  self._offset + start_photon, nphotons, C.byref(raw))
        owner = _wrap_result(self.ctx, raw)
        return GPUPhotons._from_owner(self.ctx, owner, 0, owner.capacity)


class GPUPhotonsSlice(GPUPhotons):
    """A window onto another GPUPhotons (what iterate_copies() yields); it does not own the
    device memory.  Built from PhotonField views, like the reference builds it from GPUArray
    slices."""

    def __init__(self, pos, dir, pol, wavelengths, t, last_hit_triangles, flags, weights, evidx):
        self.ctx = pos.owner.ctx
        self._owner, self._offset, self._n = pos.owner, pos.offset, pos.n
        self.pos, self.dir, self.pol, self.wavelengths, self.t = pos, dir, pol, wavelengths, t
        self.last_hit_triangles, self.flags, self.weights, self.evidx = \
            last_hit_triangles, flags, weights, evidx
        self.true_nphotons = len(pos)
        self.ncopies = 1
