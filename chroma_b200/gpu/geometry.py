"""GPUGeometry: upload a flattened Geometry (+ its BVH) to the device
(API of ``chroma/gpu/geometry.py:13-259``)."""
import ctypes as C

import numpy as np

from chroma_b200 import _lib
from chroma_b200.geometry import standard_wavelengths
from chroma_b200.gpu.tools import DeviceArray, current_context, float3, uint3, uint4, \
    format_size
from chroma_b200.log import logger


def _uniform_step(grid, what):
    steps = np.unique(np.diff(grid))
    if len(steps) != 1:
        raise ValueError('%s must be equally spaced apart.' % what)
    return float(steps[0])


def _interp(grid, prop):
    # linear interpolation is essential: the propagation code relies on interpolated
    # probabilities still summing to one (chroma/gpu/geometry.py:41-45)
    prop = np.asarray(prop)
    return np.interp(grid, prop[:, 0], prop[:, 1]).astype(np.float32)


def pack_optics(geometry, wavelengths, times):
    "Flatten Material/Surface objects into the arrays of chroma_geometry_desc."
    nw, nt = len(wavelengths), len(times)
    mats = geometry.unique_materials
    mat_props = np.zeros((len(mats), 3, nw), dtype=np.float32)
    num_comp = np.zeros(len(mats), dtype=np.uint32)
    c_prob, c_wcdf, c_abs, c_tcdf = [], [], [], []
    for i, m in enumerate(mats):
        if m is None:
            raise Exception('one or more triangles is missing a material.')
        mat_props[i, 0] = _interp(wavelengths, m.refractive_index)
        mat_props[i, 1] = _interp(wavelengths, m.absorption_length)
        mat_props[i, 2] = _interp(wavelengths, m.scattering_length)
        n = len(m.comp_reemission_prob)
        assert n == len(m.comp_reemission_wvl_cdf), 'component arrays must be same length'
        assert n == len(m.comp_reemission_time_cdf), 'component arrays must be same length'
        assert n == len(m.comp_absorption_length), 'component arrays must be same length'
        num_comp[i] = n
        for k in range(n):
            c_prob.append(_interp(wavelengths, m.comp_reemission_prob[k]))
            c_wcdf.append(_interp(wavelengths, m.comp_reemission_wvl_cdf[k]))
            c_abs.append(_interp(wavelengths, m.comp_absorption_length[k]))
            c_tcdf.append(_interp(times, m.comp_reemission_time_cdf[k]))
    ncomp = int(num_comp.sum())

    def stack(rows, width):
        return np.ascontiguousarray(rows, dtype=np.float32).reshape(len(rows), width) \
            if rows else np.zeros((0, width), dtype=np.float32)

    surfs = geometry.unique_surfaces
    ns = len(surfs)
    present = np.zeros(ns, dtype=np.uint8)
    s_props = np.zeros((ns, 8, nw), dtype=np.float32)
    s_model = np.zeros(ns, dtype=np.uint32)
    s_trans = np.zeros(ns, dtype=np.uint32)
    s_thick = np.zeros(ns, dtype=np.float32)
    d_nang = np.zeros(ns, dtype=np.uint32)
    d_ang, d_ref, d_tra = [], [], []
    names = ('detect', 'absorb', 'reemit', 'reflect_diffuse', 'reflect_specular', 'eta', 'k',
             'reemission_cdf')
    for i, s in enumerate(surfs):
        if s is None:
            continue
        present[i] = 1
        for k, name in enumerate(names):
            s_props[i, k] = _interp(wavelengths, getattr(s, name))
        s_model[i] = s.model
        s_trans[i] = s.transmissive
        s_thick[i] = s.thickness
        if s.dichroic_props:
            p = s.dichroic_props
            d_nang[i] = len(p.angles)
            d_ang.append(np.asarray(p.angles, dtype=np.float32))
            d_ref.append(np.stack([_interp(wavelengths, r) for r in p.dichroic_reflect]))
            d_tra.append(np.stack([_interp(wavelengths, r) for r in p.dichroic_transmit]))
    return dict(
        material_props=mat_props, material_num_comp=num_comp, ncomp_total=ncomp,
        comp_reemission_prob=stack(c_prob, nw), comp_reemission_wvl_cdf=stack(c_wcdf, nw),
        comp_absorption_length=stack(c_abs, nw), comp_reemission_time_cdf=stack(c_tcdf, nt),
        surface_present=present, surface_props=s_props, surface_model=s_model,
        surface_transmissive=s_trans, surface_thickness=s_thick, dichroic_nangles=d_nang,
        dichroic_angles=np.concatenate(d_ang).astype(np.float32) if d_ang else np.zeros(0, np.float32),
        dichroic_reflect=np.concatenate(d_ref).astype(np.float32) if d_ref else np.zeros((0, nw), np.float32),
        dichroic_transmit=np.concatenate(d_tra).astype(np.float32) if d_tra else np.zeros((0, nw), np.float32))


def pack_instances(geometry):
    """Arrays of chroma_instances_desc from Geometry.placements(): every distinct Solid once, in its own
    frame, and one (template, rotation, displacement) per placed solid.  Nothing is flattened."""
    templates, of_solid, rot, dis = geometry.placements()
    mlookup, slookup = geometry.material_tables()
    verts, tris, codes, colors = [], [], [], []
    voff, toff = [0], [0]
    for s in templates:
        m = s.mesh
        verts.append(np.asarray(m.vertices, dtype=np.float32))
        tris.append(np.asarray(m.triangles, dtype=np.uint32))
        inner = np.fromiter((mlookup[id(x)] for x in s.inner_material), dtype=np.int64, count=len(m.triangles))
        outer = np.fromiter((mlookup[id(x)] for x in s.outer_material), dtype=np.int64, count=len(m.triangles))
        surf = np.fromiter((-1 if x is None else slookup[id(x)] for x in s.surface), dtype=np.int64,
                           count=len(m.triangles))
        codes.append((((inner & 0xff) << 24) | ((outer & 0xff) << 16) | ((surf & 0xff) << 8)).astype(np.uint32))
        colors.append(np.asarray(s.color, dtype=np.uint32))
        voff.append(voff[-1] + len(m.vertices))
        toff.append(toff[-1] + len(m.triangles))
    return dict(
        ntemplates=len(templates), vertices=np.ascontiguousarray(np.concatenate(verts)),
        triangles=np.ascontiguousarray(np.concatenate(tris)), material_codes=np.concatenate(codes),
        colors=np.concatenate(colors), vertex_offset=np.asarray(voff, dtype=np.uint64),
        triangle_offset=np.asarray(toff, dtype=np.uint64), ninstances=len(of_solid),
        instance_template=np.ascontiguousarray(of_solid, dtype=np.uint32),
        rotations=np.ascontiguousarray(rot, dtype=np.float32).reshape(-1, 9),
        displacements=np.ascontiguousarray(dis, dtype=np.float32).reshape(-1, 3))


class GPUGeometry(object):
    """``instanced=True`` (not in the reference) builds the two-level geometry instead: every distinct
    Solid is uploaded once and the detector is a tree over its placements (Geometry.placements());
    the mesh is never flattened on the host and no BVH is needed.  Triangle and solid ids are those
    of the flattened mesh.  Results agree with the flat geometry statistically, not bit for bit."""

    def __init__(self, geometry, wavelengths=None, times=None, print_usage=False,
                 min_free_gpu_mem=300e6, instanced=False):
        self.ctx = current_context()
        self.instanced = bool(instanced)
        if wavelengths is None:
            wavelengths = standard_wavelengths
        wavelengths = np.asarray(wavelengths, dtype=np.float32)
        wavelength_step = _uniform_step(wavelengths, 'wavelengths')
        if times is None:
            time_step = 0.05
            times = np.arange(0, 1000, time_step)
        else:
            time_step = _uniform_step(times, 'times')
        if self.instanced:
            self._init_instanced(geometry, wavelengths, wavelength_step, times, time_step, print_usage)
            return
        if not hasattr(geometry, 'mesh'):
            geometry.flatten()
        if geometry.bvh is None:
            raise ValueError('geometry has no BVH: build one with chroma_b200.bvh.'
                             'make_recursive_grid_bvh(geometry.mesh) or loader.build()')

        opt = pack_optics(geometry, wavelengths, times)
        material_codes = (((geometry.inner_material_index & 0xff) << 24) |
                          ((geometry.outer_material_index & 0xff) << 16) |
                          ((geometry.surface_index & 0xff) << 8)).astype(np.uint32)
        keep = dict(opt)
        keep['vertices'] = _lib.carray(geometry.mesh.vertices, np.float32)
        keep['triangles'] = _lib.carray(geometry.mesh.triangles, np.uint32)
        keep['material_codes'] = material_codes
        keep['colors'] = _lib.carray(geometry.colors, np.uint32)
        keep['solid_id'] = _lib.carray(geometry.solid_id, np.uint32)
        bvh = geometry.bvh
        d = _lib.GeometryDesc()
        d.nvertices = len(keep['vertices'])
        d.ntriangles = len(keep['triangles'])
        device_bvh = getattr(bvh, 'device_handle', None)
        if device_bvh is not None:
            d.bvh = device_bvh
            d.nodes, d.nnodes = None, 0
        else:
            keep['nodes'] = np.ascontiguousarray(bvh.nodes).view(np.uint32).reshape(-1, 4)
            d.nodes = _lib.ptr(keep['nodes'])
            d.nnodes = len(keep['nodes'])
        d.world_origin = (C.c_float * 3)(*[float(x) for x in bvh.world_coords.world_origin])
        d.world_scale = float(bvh.world_coords.world_scale)
        d.nwavelengths, d.wavelength_start, d.wavelength_step = \
            len(wavelengths), float(wavelengths[0]), wavelength_step
        d.ntimes, d.time_start, d.time_step = len(times), float(times[0]), time_step
        d.nmaterials = len(geometry.unique_materials)
        d.nsurfaces = len(geometry.unique_surfaces)
        d.ncomp_total = opt['ncomp_total']
        for name in ('vertices', 'triangles', 'material_codes', 'colors', 'solid_id',
                     'material_props', 'material_num_comp', 'comp_reemission_prob',
                     'comp_reemission_wvl_cdf', 'comp_absorption_length',
                     'comp_reemission_time_cdf', 'surface_present', 'surface_props',
                     'surface_model', 'surface_transmissive', 'surface_thickness',
                     'dichroic_nangles', 'dichroic_angles', 'dichroic_reflect',
                     'dichroic_transmit'):
            arr = keep[name]
            setattr(d, name, _lib.ptr(arr) if arr.size else None)
        handle = C.c_void_p()
        _lib.call('chroma_geometry_create', self.ctx.handle, C.byref(d), C.byref(handle))
        self.handle = handle

        # public attributes of the reference class.  The engine searches its own triangle records, so the
        # device copies of the mesh arrays (5 GB at 29,023 PMTs) and the reference-format node array
        # (3.9 GB) only ever served these attributes: they are released / not kept, and an attribute is
        # made the first time it is read (__getattr__): the mesh arrays are uploaded from the host
        # geometry, .nodes is rebuilt on the device from the engine's copy.
        self._material_codes_host = material_codes
        _lib.call('chroma_geometry_release_mesh', self.handle)
        self.extra_nodes = DeviceArray(0, uint4, ctx=self.ctx, ptr=0)  # never needed on B200
        self.colors = self._array(4, np.uint32)
        self.world_origin = np.array(tuple(bvh.world_coords.world_origin), dtype=float3)
        self.world_scale = np.float32(bvh.world_coords.world_scale)
        self.gpudata = self.handle
        self.geometry = geometry
        if print_usage:
            self.print_device_usage()
        logger.info(self.device_usage_str())

    def _init_instanced(self, geometry, wavelengths, wavelength_step, times, time_step, print_usage):
        inst = pack_instances(geometry)
        opt = pack_optics(geometry, wavelengths, times)
        d = _lib.GeometryDesc()
        d.nwavelengths, d.wavelength_start, d.wavelength_step = \
            len(wavelengths), float(wavelengths[0]), wavelength_step
        d.ntimes, d.time_start, d.time_step = len(times), float(times[0]), time_step
        d.nmaterials = len(geometry.unique_materials)
        d.nsurfaces = len(geometry.unique_surfaces)
        d.ncomp_total = opt['ncomp_total']
        for name in ('material_props', 'material_num_comp', 'comp_reemission_prob',
                     'comp_reemission_wvl_cdf', 'comp_absorption_length', 'comp_reemission_time_cdf',
                     'surface_present', 'surface_props', 'surface_model', 'surface_transmissive',
                     'surface_thickness', 'dichroic_nangles', 'dichroic_angles', 'dichroic_reflect',
                     'dichroic_transmit'):
            arr = opt[name]
            setattr(d, name, _lib.ptr(arr) if arr.size else None)
        i = _lib.InstancesDesc()
        i.ntemplates, i.ninstances = inst['ntemplates'], inst['ninstances']
        for name in ('vertices', 'triangles', 'material_codes', 'colors', 'vertex_offset', 'triangle_offset',
                     'instance_template', 'rotations', 'displacements'):
            setattr(i, name, _lib.ptr(inst[name]))
        handle = C.c_void_p()
        _lib.call('chroma_geometry_create_instanced', self.ctx.handle, C.byref(d), C.byref(i), C.byref(handle))
        self.handle = handle
        self.gpudata = self.handle
        self.geometry = geometry
        self.extra_nodes = DeviceArray(0, uint4, ctx=self.ctx, ptr=0)
        self.nodes = DeviceArray(0, uint4, ctx=self.ctx, ptr=0)   # no reference-format BVH in this mode
        tree = self.tree()
        self.world_origin = None
        self.world_scale = None
        if print_usage:
            self.print_device_usage()
        logger.info('two-level geometry: %d placements of %d templates, %d triangle records, %s on the device'
                    % (tree['placements'], tree['templates'], tree['triangle_slots'], format_size(tree['bytes'])))

    # the flat arrays of the reference class are assembled on the device the first time they are read
    def __getattr__(self, name):
        lazy = {'vertices': (1, float3), 'triangles': (2, uint3), 'material_codes': (3, np.uint32),
                'colors': (4, np.uint32), 'solid_id_map': (5, np.uint32)}
        if name in lazy and self.__dict__.get('instanced'):
            arr = self._array(*lazy[name])               # assembled on the device from the placements
        elif name == 'nodes' and 'handle' in self.__dict__:
            arr = self._array(0, uint4)                  # reference-format nodes rebuilt on the device
        elif name in lazy and 'geometry' in self.__dict__ and self.__dict__['geometry'] is not None:
            from chroma_b200.gpu.tools import to_gpu
            g = self.__dict__['geometry']
            host = {'vertices': lambda: _lib.carray(g.mesh.vertices, np.float32).view(float3)[:, 0],
                    'triangles': lambda: _lib.carray(g.mesh.triangles, np.uint32).view(uint3)[:, 0],
                    'material_codes': lambda: self.__dict__['_material_codes_host'],
                    'solid_id_map': lambda: _lib.carray(g.solid_id, np.uint32)}[name]()
            arr = to_gpu(host, ctx=self.__dict__['ctx'])
        else:
            raise AttributeError(name)
        self.__dict__[name] = arr
        return arr

    def tree(self):
        "which search structure is live (chroma_geometry_tree)"
        out = (C.c_uint64 * 8)()
        _lib.call('chroma_geometry_tree', self.handle, out)
        kind = {0: 'reference-format tree only', 1: 'wide tree', 2: 'two-level tree'}[int(out[0])]
        return dict(kind=kind, kind_code=int(out[0]), placements=int(out[1]), templates=int(out[2]),
                    triangle_slots=int(out[3]), top_groups=int(out[4]), groups=int(out[5]), bytes=int(out[6]))

    def _array(self, which, dtype):
        p, n = C.c_void_p(), C.c_uint64()
        _lib.call('chroma_geometry_array', self.handle, which, C.byref(p), C.byref(n))
        return DeviceArray(n.value, dtype, ctx=self.ctx, ptr=p.value or 0, base=self)

    def info(self):
        out = (C.c_uint64 * 8)()
        _lib.call('chroma_geometry_info', self.handle, out)
        return dict(nnodes=out[0], ntriangles=out[1], nleaves=out[2], device_bytes=out[3],
                    optics_smem_floats=out[4], wide_groups=out[5], wide_levels=out[6],
                    wide_leaves=out[7])

    def device_usage_str(self):
        mem = self.ctx.info()
        s = 'device usage:\n' + '-' * 10 + '\n'
        nn = self.info()['nnodes']
        s += '%-15s %6s %6s\n' % ('nodes', format_size(nn), format_size(nn * 16))
        s += '%-15s %6s %6s\n' % ('total', '', format_size(self.info()['device_bytes']))
        s += '-' * 10 + '\n'
        s += '%-15s %6s %6s\n' % ('device total', '', format_size(mem['total']))
        s += '%-15s %6s %6s\n' % ('device used', '', format_size(mem['total'] - mem['free']))
        s += '%-15s %6s %6s\n' % ('device free', '', format_size(mem['free']))
        return s

    def print_device_usage(self):
        print(self.device_usage_str())
        print()

    def reset_colors(self):
        self.colors.set(self.geometry.colors.astype(np.uint32))

    def __del__(self):
        try:
            if getattr(self, 'handle', None) is not None and self.ctx.handle is not None:
                _lib.call('chroma_geometry_destroy', self.handle)
                self.handle = None
        except Exception:
            pass
