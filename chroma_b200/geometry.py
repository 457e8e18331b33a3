"""Host-side geometry schema consumed by :class:`chroma_b200.gpu.GPUGeometry`.

Same names and attributes as the reference's ``chroma/geometry.py`` (Mesh :19-110,
Solid :115-219, Material :221-249, DichroicProps :257-261, Surface :263-295,
Geometry :297-392) so user code building detectors keeps working; the bodies are a fresh,
vectorised numpy implementation (and run under numpy 2).
"""
from hashlib import md5 as _md5

import numpy as np

# all material/surface properties are resampled on this grid before upload
# (chroma/geometry.py:17)
standard_wavelengths = np.arange(60, 1000, 5).astype(np.float32)


class Mesh(object):
    "Triangle mesh: ``vertices`` float32 (nv,3), ``triangles`` int32 (nt,3)."

    def __init__(self, vertices, triangles, remove_duplicate_vertices=False, round=True,
                 remove_null_triangles=True):
        vertices = np.asarray(vertices, dtype=np.float32)
        triangles = np.asarray(triangles, dtype=np.int32)
        if vertices.ndim != 2 or vertices.shape[1] != 3:
            raise ValueError('shape mismatch')
        if triangles.ndim != 2 or triangles.shape[1] != 3:
            raise ValueError('shape mismatch')
        if (triangles < 0).any():
            raise ValueError('indices in `triangles` must be positive.')
        if len(triangles) and (triangles >= len(vertices)).any():
            raise ValueError('indices in `triangles` must be less than the '
                             'length of the vertex array.')
        self.vertices = vertices.round(decimals=12) if round else vertices
        self.triangles = triangles
        if remove_duplicate_vertices:
            self.remove_duplicate_vertices()
        if remove_null_triangles:
            self.remove_null_triangles()

    def get_triangle_centers(self):
        return np.mean(self.assemble(), axis=1)

    def get_bounds(self):
        return np.min(self.vertices, axis=0), np.max(self.vertices, axis=0)

    def remove_duplicate_vertices(self):
        unique, inverse = np.unique(self.vertices, axis=0, return_inverse=True)
        self.vertices = np.ascontiguousarray(unique, dtype=np.float32)
        self.triangles = inverse.ravel()[self.triangles].astype(self.triangles.dtype)

    def remove_null_triangles(self):
        "Drop triangles that reference the same vertex twice; returns the keep mask."
        if len(self.triangles) == 0:
            return
        t = self.triangles
        mask = (t[:, 0] != t[:, 1]) & (t[:, 1] != t[:, 2]) & (t[:, 0] != t[:, 2])
        self.triangles = t[mask]
        return mask

    def assemble(self, key=slice(None), group=True):
        idx = self.triangles[key] if group else self.triangles[key].flatten()
        return self.vertices[idx]

    def __add__(self, other):
        return Mesh(np.concatenate((self.vertices, other.vertices)),
                    np.concatenate((self.triangles, other.triangles + len(self.vertices))))

    def md5(self):
        checksum = _md5(np.ascontiguousarray(self.vertices))
        checksum.update(np.ascontiguousarray(self.triangles))
        return checksum.hexdigest()


def _unique_objects(seq):
    "Unique python objects of an object array, in first-seen order (None allowed)."
    seen, out = set(), []
    for x in seq:
        if id(x) not in seen:
            seen.add(id(x))
            out.append(x)
    return out


def _per_triangle(value, n, dtype=object):
    if isinstance(value, np.ndarray) or isinstance(value, (list, tuple)):
        if len(value) != n:
            raise ValueError('shape mismatch')
        arr = np.empty(n, dtype=dtype)
        arr[:] = list(value) if dtype is object else value
        return arr
    arr = np.empty(n, dtype=dtype)
    arr[:] = value
    return arr


class Solid(object):
    "Attaches inner/outer materials, a surface and a colour to every triangle of a Mesh."

    def __init__(self, mesh, inner_material=None, outer_material=None, surface=None,
                 color=0x33ffffff, material1=None, material2=None):
        if material1 is not None or material2 is not None:  # deprecated names
            inner_material, outer_material = material1, material2
        n = len(mesh.triangles)
        self.mesh = mesh
        self.inner_material = _per_triangle(inner_material, n)
        self.outer_material = _per_triangle(outer_material, n)
        self.surface = _per_triangle(surface, n)
        self.color = _per_triangle(color, n, dtype=np.uint32)
        self.unique_materials = _unique_objects(
            np.concatenate([self.inner_material, self.outer_material]))
        self.unique_surfaces = _unique_objects(self.surface)

    def __add__(self, other):
        return Solid(self.mesh + other.mesh,
                     np.concatenate((self.inner_material, other.inner_material)),
                     np.concatenate((self.outer_material, other.outer_material)),
                     np.concatenate((self.surface, other.surface)),
                     np.concatenate((self.color, other.color)))


def _table(value, wavelengths):
    if np.iterable(value):
        if len(value) != len(wavelengths):
            raise ValueError('shape mismatch')
    else:
        value = np.tile(value, len(wavelengths))
    return np.array(list(zip(wavelengths, value)), dtype=np.float32)


class Material(object):
    "Bulk optical properties; each property is an (n,2) array of (wavelength, value)."

    def __init__(self, name='none'):
        self.name = name
        self.refractive_index = None
        self.absorption_length = None
        self.scattering_length = None
        self.scintillation_spectrum = None
        self.scintillation_light_yield = None
        self.scintillation_rise_time = None
        self.scintillation_waveform = None
        self.scintillation_mod = None
        self.comp_reemission_prob = []
        self.comp_reemission_wvl_cdf = []
        self.comp_reemission_times = []
        self.comp_reemission_time_cdf = []
        self.comp_absorption_length = []
        self.density = 0.0
        self.composition = {}

    def set(self, name, value, wavelengths=standard_wavelengths):
        self.__dict__[name] = _table(value, wavelengths)

    def __repr__(self):
        return '<Material %s>' % self.name


vacuum = Material('vacuum')
vacuum.set('refractive_index', 1.0)
vacuum.set('absorption_length', 1e6)
vacuum.set('scattering_length', 1e6)


class DichroicProps(object):
    def __init__(self, angles, reflect, transmit):
        self.angles = np.asarray(angles)                 # [angle]
        self.dichroic_reflect = np.asarray(reflect)      # [angle][point] (wavelength, value)
        self.dichroic_transmit = np.asarray(transmit)    # [angle][point]


class Surface(object):
    "Surface optical properties; ``model`` selects default / complex / wls / dichroic."

    def __init__(self, name='none', model=0):
        self.name = name
        self.model = model
        for prop in ('detect', 'absorb', 'reemit', 'reflect_diffuse', 'reflect_specular',
                     'eta', 'k', 'reemission_cdf'):
            self.set(prop, 0)
        self.dichroic_props = None
        self.thickness = 0.0
        self.transmissive = 0

    def set(self, name, value, wavelengths=standard_wavelengths):
        if (np.asarray(value) < 0.0).any():
            raise Exception('all probabilities must be >= 0.0')
        self.__dict__[name] = _table(value, wavelengths)

    def __repr__(self):
        return '<Surface %s>' % self.name


class Geometry(object):
    "A list of placed Solids, flattened into one mesh with per-triangle property indices."

    def __init__(self, detector_material=None):
        self.detector_material = detector_material
        self.solids = []
        self.solid_rotations = []
        self.solid_displacements = []
        self.bvh = None

    def add_solid(self, solid, rotation=None, displacement=None):
        rotation = np.identity(3) if rotation is None else np.asarray(rotation, dtype=np.float32)
        if rotation.shape != (3, 3):
            raise ValueError('rotation matrix has the wrong shape.')
        displacement = np.zeros(3) if displacement is None else \
            np.asarray(displacement, dtype=np.float32)
        if displacement.shape != (3,):
            raise ValueError('displacement vector has the wrong shape.')
        self.solid_rotations.append(rotation.astype(np.float32))
        self.solid_displacements.append(displacement)
        self.solids.append(solid)
        return len(self.solids) - 1

    def placements(self):
        """The detector as placed templates, without flattening: (templates, template_of_solid,
        rotations, displacements) with `templates` the distinct Solid objects in order of first use,
        `template_of_solid` uint32[nsolids], `rotations` float32[nsolids, 3, 3] (world = R v + d, the
        matrices handed to add_solid) and `displacements` float32[nsolids, 3].  Input of the
        two-level geometry (GPUGeometry(..., instanced=True))."""
        pre = getattr(self, '_placements', None)
        if pre is not None:
            return pre
        templates, index, of_solid = [], {}, np.empty(len(self.solids), dtype=np.uint32)
        for i, solid in enumerate(self.solids):
            k = index.get(id(solid))
            if k is None:
                k = index[id(solid)] = len(templates)
                templates.append(solid)
            of_solid[i] = k
        rot = np.ascontiguousarray(self.solid_rotations, dtype=np.float32).reshape(-1, 3, 3)
        dis = np.ascontiguousarray(self.solid_displacements, dtype=np.float32).reshape(-1, 3)
        return templates, of_solid, rot, dis

    def material_tables(self):
        """unique_materials / unique_surfaces (what flatten() derives from the solids) without building
        the mesh; returns ({id(material): index}, {id(surface): index})."""
        if not hasattr(self, 'unique_materials'):
            templates = self.placements()[0]
            self.unique_materials = _unique_objects([m for s in templates for m in s.unique_materials])
            self.unique_surfaces = _unique_objects([x for s in templates for x in s.unique_surfaces])
        return ({id(m): i for i, m in enumerate(self.unique_materials)},
                {id(x): i for i, x in enumerate(self.unique_surfaces)})

    def flatten(self):
        "Build mesh / colors / solid_id / material and surface index arrays (no BVH)."
        if hasattr(self, 'mesh'):
            return
        deferred = getattr(self, '_deferred_flatten', None)
        if deferred is not None:      # demo.detector(fast=True): vectorised assembly, done on first need
            self._deferred_flatten = None
            deferred()
            return
        nv = np.cumsum([0] + [len(s.mesh.vertices) for s in self.solids])
        nt = np.cumsum([0] + [len(s.mesh.triangles) for s in self.solids])
        vertices = np.empty((nv[-1], 3), dtype=np.float32)
        triangles = np.empty((nt[-1], 3), dtype=np.uint32)
        for i, solid in enumerate(self.solids):
            vertices[nv[i]:nv[i + 1]] = \
                np.inner(solid.mesh.vertices, self.solid_rotations[i]) + self.solid_displacements[i]
            triangles[nt[i]:nt[i + 1]] = solid.mesh.triangles + nv[i]
        self.mesh = Mesh(vertices, triangles, remove_duplicate_vertices=True,
                         remove_null_triangles=False)
        self.colors = np.concatenate([s.color for s in self.solids])
        self.solid_id = np.concatenate(
            [np.full(len(s.mesh.triangles), i, dtype=np.uint32) for i, s in enumerate(self.solids)])

        self.unique_materials = _unique_objects(
            [m for s in self.solids for m in s.unique_materials])
        mlookup = {id(m): i for i, m in enumerate(self.unique_materials)}
        self.unique_surfaces = _unique_objects(
            [x for s in self.solids for x in s.unique_surfaces])
        slookup = {id(x): i for i, x in enumerate(self.unique_surfaces)}

        def indices(arrays, lookup):
            out = np.empty(nt[-1], dtype=np.int32)
            cache = {}  # the same Solid (e.g. one PMT model) is usually placed many times
            for i, a in enumerate(arrays):
                idx = cache.get(id(a))
                if idx is None:
                    idx = np.fromiter((lookup[id(x)] for x in a), dtype=np.int32, count=len(a))
                    cache[id(a)] = idx
                out[nt[i]:nt[i + 1]] = idx
            return out

        self.inner_material_index = indices([s.inner_material for s in self.solids], mlookup)
        self.outer_material_index = indices([s.outer_material for s in self.solids], mlookup)
        self.surface_index = indices([s.surface for s in self.solids], slookup)
        if id(None) in slookup:
            self.surface_index[self.surface_index == slookup[id(None)]] = -1
