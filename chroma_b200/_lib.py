"""ctypes binding of libchroma_b200.so (prototypes of include/chroma_b200.h).

There is no fallback: if the shared library is missing or a call fails, a
:class:`ChromaError` is raised.  ``CHROMA_B200_LIB`` overrides the library path.
"""
import atexit
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('CHROMA_B200_LIB', os.path.join(_HERE, 'libchroma_b200.so'))


class ChromaError(RuntimeError):
    "A libchroma_b200 call failed (status, message from chroma_last_error())."

    def __init__(self, status, message):
        RuntimeError.__init__(self, 'libchroma_b200 error %d: %s' % (status, message))
        self.status = status


class SourceDesc(C.Structure):
    "mirror of chroma_source"
    _fields_ = [
        ('pos', C.c_float * 3), ('rot', C.c_float * 9), ('cos_half_angle', C.c_float),
        ('wavelength_lo', C.c_float), ('wavelength_hi', C.c_float), ('t0', C.c_float),
        ('weight', C.c_float), ('evidx', C.c_uint32), ('sequence', C.c_uint32),
    ]


class GeometryDesc(C.Structure):
    "mirror of chroma_geometry_desc"
    _fields_ = [
        ('vertices', C.c_void_p), ('triangles', C.c_void_p), ('material_codes', C.c_void_p),
        ('colors', C.c_void_p), ('solid_id', C.c_void_p),
        ('nvertices', C.c_uint64), ('ntriangles', C.c_uint64),
        ('nodes', C.c_void_p), ('nnodes', C.c_uint64), ('bvh', C.c_void_p),
        ('world_origin', C.c_float * 3), ('world_scale', C.c_float),
        ('nwavelengths', C.c_uint32), ('wavelength_start', C.c_float), ('wavelength_step', C.c_float),
        ('ntimes', C.c_uint32), ('time_start', C.c_float), ('time_step', C.c_float),
        ('nmaterials', C.c_uint32), ('material_props', C.c_void_p),
        ('material_num_comp', C.c_void_p), ('ncomp_total', C.c_uint32),
        ('comp_reemission_prob', C.c_void_p), ('comp_reemission_wvl_cdf', C.c_void_p),
        ('comp_absorption_length', C.c_void_p), ('comp_reemission_time_cdf', C.c_void_p),
        ('nsurfaces', C.c_uint32), ('surface_present', C.c_void_p), ('surface_props', C.c_void_p),
        ('surface_model', C.c_void_p), ('surface_transmissive', C.c_void_p),
        ('surface_thickness', C.c_void_p),
        ('dichroic_nangles', C.c_void_p), ('dichroic_angles', C.c_void_p),
        ('dichroic_reflect', C.c_void_p), ('dichroic_transmit', C.c_void_p),
    ]


class PhotonArrays(C.Structure):
    "mirror of chroma_photon_arrays"
    _fields_ = [
        ('pos', C.c_void_p), ('dir', C.c_void_p), ('pol', C.c_void_p), ('wavelengths', C.c_void_p),
        ('t', C.c_void_p), ('weights', C.c_void_p), ('last_hit', C.c_void_p), ('flags', C.c_void_p),
        ('evidx', C.c_void_p), ('pos_value', C.c_float * 3), ('t_value', C.c_float),
        ('weight_value', C.c_float), ('flags_value', C.c_uint32), ('evidx_value', C.c_uint32),
        ('last_hit_value', C.c_int32), ('borrow_sources', C.c_uint32),
    ]


class InstancesDesc(C.Structure):
    "mirror of chroma_instances_desc"
    _fields_ = [
        ('ntemplates', C.c_uint32), ('vertices', C.c_void_p), ('triangles', C.c_void_p),
        ('material_codes', C.c_void_p), ('colors', C.c_void_p), ('vertex_offset', C.c_void_p),
        ('triangle_offset', C.c_void_p), ('ninstances', C.c_uint64), ('instance_template', C.c_void_p),
        ('rotations', C.c_void_p), ('displacements', C.c_void_p),
    ]


_P = C.c_void_p
_PP = C.POINTER(C.c_void_p)
_U64 = C.c_uint64
_U32 = C.c_uint32
_PU64 = C.POINTER(C.c_uint64)

# name -> argtypes (all functions return int except the two string getters)
_PROTOTYPES = {
    'chroma_ctx_create': [C.c_int, _PP],
    'chroma_ctx_destroy': [_P],
    'chroma_ctx_synchronize': [_P],
    'chroma_ctx_reload_knobs': [_P],
    'chroma_ctx_stream': [_P, _PP],
    'chroma_ctx_info': [_P, _PU64],
    'chroma_ctx_launch_count': [_P, _PU64],
    'chroma_ctx_timer_begin': [_P],
    'chroma_ctx_timer_end': [_P, C.POINTER(C.c_float)],
    'chroma_ctx_flush_l2': [_P],
    'chroma_malloc': [_P, C.c_size_t, _PP],
    'chroma_free': [_P, _P],
    'chroma_memcpy_h2d': [_P, _P, _P, C.c_size_t],
    'chroma_memcpy_d2h': [_P, _P, _P, C.c_size_t],
    'chroma_memcpy_d2d': [_P, _P, _P, C.c_size_t],
    'chroma_memset32': [_P, _P, _U32, C.c_size_t],
    'chroma_host_alloc': [C.c_size_t, _PP],
    'chroma_host_free': [_P],
    'chroma_geometry_create': [_P, C.POINTER(GeometryDesc), _PP],
    'chroma_geometry_create_instanced': [_P, C.POINTER(GeometryDesc), C.POINTER(InstancesDesc), _PP],
    'chroma_geometry_tree': [_P, _PU64],
    'chroma_geometry_destroy': [_P],
    'chroma_geometry_array': [_P, C.c_int, _PP, _PU64],
    'chroma_geometry_release_mesh': [_P],
    'chroma_geometry_info': [_P, _PU64],
    'chroma_detector_attach': [_P, _P, _U64, _U32, _P, _P, _U32, _P, _P, _U32],
    'chroma_intersect_device': [_P, _U64, _P, _P, _P, _P, _P],
    'chroma_intersect_host': [_P, _U64, _P, _P, _P, _P, _P],
    'chroma_intersect_host_static': [_P, _U64, _P, _P, _P, _P, _P],
    'chroma_intersect_count_device': [_P, _U64, _P, _P, _P, _P, _P, _P],
    'chroma_rng_create': [_P, _U64, _U64, _PP],
    'chroma_rng_destroy': [_P],
    'chroma_rng_get_epoch': [_P, C.POINTER(C.c_uint32)],
    'chroma_rng_set_epoch': [_P, _U32],
    'chroma_rng_fill_uniform': [_P, _U64, _P],
    'chroma_photons_create': [_P, _U64, _U64, _PP],
    'chroma_photons_destroy': [_P],
    'chroma_photons_size': [_P, _PU64],
    'chroma_photons_upload': [_P, _U64, _U64, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    'chroma_photons_upload_arrays': [_P, _U64, _U64, C.POINTER(PhotonArrays)],
    'chroma_photons_download': [_P, _U64, _U64, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    'chroma_photons_field_get': [_P, _U64, _U64, C.c_int, _P],
    'chroma_photons_field_set': [_P, _U64, _U64, C.c_int, _P],
    'chroma_photons_duplicate': [_P, _U64, _U32],
    'chroma_photons_generate': [_P, _U64, _U64, _P, _U64],
    'chroma_photons_copy': [_P, _P],
    'chroma_photons_propagate': [_P, _P, _P, _U64, _U64, C.c_int, C.c_int, C.c_int, _PU64],
    'chroma_photons_select': [_P, _U32, _U64, _U64, _PP],
    'chroma_photons_flat_hits': [_P, _P, _U32, _U64, _U64, _PP],
    'chroma_photons_gather': [_P, _P, _U64, _PP],
    'chroma_photons_live_ids': [_P, _U64, _U64, _P, _PU64],
    'chroma_daq_create': [_P, _U32, _PP],
    'chroma_daq_destroy': [_P],
    'chroma_daq_begin_acquire': [_P],
    'chroma_daq_acquire': [_P, _P, _P, _U64, _U64, C.c_float],
    'chroma_daq_end_acquire': [_P],
    'chroma_daq_get': [_P, _P, _P, _P],
    'chroma_daq_arrays': [_P, _PP, _PP, _PP, _PU64],
    'chroma_daq_pack': [_P, _P],
    'chroma_daq_merge_gathered': [_P, _P, _U32],
    'chroma_daq_int_arrays': [_P, _PP, _PP, _PU64],
    'chroma_pdf_create': [_P, _U32, _U32, C.c_float, C.c_float, _U32, C.c_float, C.c_float, _PP],
    'chroma_pdf_destroy': [_P],
    'chroma_pdf_clear': [_P],
    'chroma_pdf_add': [_P, _P],
    'chroma_pdf_add_device': [_P, _P, _P],
    'chroma_pdf_get': [_P, _P, _P],
    'chroma_pdf_arrays': [_P, _PP, _PP, _PU64],
    'chroma_pdf_eval_create': [_P, _U32, _P, _P, C.c_float, C.c_float, C.c_float, _U32, _PP],
    'chroma_pdf_eval_destroy': [_P],
    'chroma_pdf_eval_clear': [_P],
    'chroma_pdf_eval_accumulate': [_P, _P, _U32],
    'chroma_pdf_eval_get': [_P, _P, _P, _P, C.POINTER(C.c_uint32)],
    'chroma_kernel_pdf_create': [_P, _U32, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, _PP],
    'chroma_kernel_pdf_destroy': [_P],
    'chroma_kernel_pdf_clear_moments': [_P],
    'chroma_kernel_pdf_accumulate_moments': [_P, _P, _P],
    'chroma_kernel_pdf_get_moments': [_P, _P, _P, _P, _P, _P],
    'chroma_kernel_pdf_setup_kernel': [_P, _P, _P, _P, _P, _P],
    'chroma_kernel_pdf_clear_kernel': [_P],
    'chroma_kernel_pdf_accumulate_kernel': [_P, _P, _P],
    'chroma_kernel_pdf_get_kernel': [_P, _P, _P, _P],
    'chroma_rays_create': [_P, _U64, _P, _P, _U32, _PP],
    'chroma_rays_destroy': [_P],
    'chroma_rays_arrays': [_P, _PP, _PP, _PP, _PP, _PP, C.POINTER(C.c_uint64)],
    'chroma_rays_translate': [_P, _P],
    'chroma_rays_rotate': [_P, C.c_float, _P],
    'chroma_rays_rotate_around_point': [_P, C.c_float, _P, _P],
    'chroma_rays_render': [_P, _P, _U32, C.c_int, _U32, _P],
    'chroma_bvh_build': [_P, _P, _U64, _P, _U64, C.c_int, _PP],
    'chroma_bvh_build_from_leaves': [_P, _P, _P, _U64, _P, C.c_float, C.c_int, _PP],
    'chroma_bvh_destroy': [_P],
    'chroma_bvh_info': [_P, _PU64, C.POINTER(C.c_float), C.POINTER(C.c_float)],
    'chroma_bvh_get': [_P, _P, _P],
    'chroma_bvh_make_leaves': [_P, _P, _U64, _P, _U64, C.POINTER(C.c_float), C.c_float, _P, _P],
    'chroma_bvh_merge_nodes_detailed': [_P, _P, _U64, _P, _P, _U64, _P],
    'chroma_bvh_collapse_chains': [_P, _P, _U64, _P, _U64],
}

EXPORTED_SYMBOLS = sorted(list(_PROTOTYPES) + ['chroma_last_error', 'chroma_version'])

_lib = None


def load():
    "Load the shared library once; raise if it is not there (no fallback)."
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ChromaError(-1, 'shared library %s not found; build it with '
                              '`python -m chroma_b200.csrc.build`' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in _PROTOTYPES.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = C.c_int
    lib.chroma_last_error.restype = C.c_char_p
    lib.chroma_last_error.argtypes = []
    lib.chroma_version.restype = C.c_char_p
    lib.chroma_version.argtypes = []
    _lib = lib
    return lib


def has_ab_kernels():
    "True for a variant build that still carries round 1's A/B kernels (CHROMA_PROP_V < 4, CHROMA_ISECT_V=2)"
    return load().chroma_version().decode().endswith('+ab')


_shutting_down = False


def _mark_shutdown():
    global _shutting_down
    _shutting_down = True


atexit.register(_mark_shutdown)


def call(name, *args):
    "Call `name` and raise ChromaError on a non-zero status."
    if _shutting_down and (name.endswith('_destroy') or name == 'chroma_free'):
        return  # interpreter exit: the CUDA runtime may already be gone; the OS reclaims
    lib = load()
    status = getattr(lib, name)(*args)
    if status != 0:
        raise ChromaError(status, lib.chroma_last_error().decode('utf-8', 'replace'))


def ptr(array):
    "void* of a C-contiguous numpy array (None -> NULL)."
    if array is None:
        return None
    assert array.flags['C_CONTIGUOUS']
    return array.ctypes.data_as(C.c_void_p)


def carray(a, dtype):
    "C-contiguous array of `dtype` (copy only when needed)."
    return np.ascontiguousarray(a, dtype=dtype)
