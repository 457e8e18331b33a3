"""Event containers: the host layout contract of the engine (``chroma/event.py``).

``Photons`` holds nine per-photon arrays (+ ``channel`` on hit lists), ``Channels`` the
per-channel DAQ result; the history bits equal ``chroma/cuda/photon.h:49-64``.
"""
import numpy as np

# photon history bits (chroma/event.py:5-17)
NO_HIT = 0x1 << 0
BULK_ABSORB = 0x1 << 1
SURFACE_DETECT = 0x1 << 2
SURFACE_ABSORB = 0x1 << 3
RAYLEIGH_SCATTER = 0x1 << 4
REFLECT_DIFFUSE = 0x1 << 5
REFLECT_SPECULAR = 0x1 << 6
SURFACE_REEMIT = 0x1 << 7
SURFACE_TRANSMIT = 0x1 << 8
BULK_REEMIT = 0x1 << 9
CHERENKOV = 0x1 << 10
SCINTILLATION = 0x1 << 11
NAN_ABORT = 0x1 << 31

_FIELDS = ('pos', 'dir', 'pol', 'wavelengths', 't', 'last_hit_triangles', 'flags', 'weights',
           'evidx', 'channel')


class Vertex(object):
    "A particle vertex; only carried through (the generator is outside this engine)."

    def __init__(self, particle_name, pos, dir, ke, t0=0.0, pol=None, steps=None, children=None,
                 trackid=-1, pdgcode=None):
        self.particle_name = particle_name
        self.pos = pos
        self.dir = dir
        self.pol = pol
        self.ke = ke
        self.t0 = t0
        self.steps = steps
        self.children = children
        self.trackid = trackid
        self.pdgcode = pdgcode


class Photons(object):
    def __init__(self, pos=np.empty((0, 3)), dir=np.empty((0, 3)), pol=np.empty((0, 3)),
                 wavelengths=np.empty((0)), t=None, last_hit_triangles=None, flags=None,
                 weights=None, evidx=None, channel=None):
        """n photons: pos/dir/pol float32 (n,3) [mm, unit, unit]; wavelengths float32 n [nm];
        t float32 n [ns]; last_hit_triangles int32 n (-1 = none); flags uint32 n (history
        bits); weights float32 n; evidx uint32 n (event index inside a batch); channel
        uint32 n (hit lists only)."""
        self.pos = np.asarray(pos, dtype=np.float32)
        self.dir = np.asarray(dir, dtype=np.float32)
        self.pol = np.asarray(pol, dtype=np.float32)
        self.wavelengths = np.asarray(wavelengths, dtype=np.float32)
        n = len(self.pos)

        def default(value, fill, dtype):
            if value is None:
                return np.full(n, fill, dtype=dtype)
            return np.asarray(value, dtype=dtype)

        self.t = default(t, 0, np.float32)
        self.last_hit_triangles = default(last_hit_triangles, -1, np.int32)
        self.flags = default(flags, 0, np.uint32)
        self.weights = default(weights, 1, np.float32)
        self.evidx = default(evidx, 0, np.uint32)
        self.channel = default(channel, 0, np.uint32)

    @staticmethod
    def join(photon_list, concatenate=True):
        "Concatenate (or, with concatenate=False, stack scalar photons of) many Photons."
        op = np.concatenate if concatenate else np.asarray
        return Photons(*[op([getattr(p, f) for p in photon_list]) for f in _FIELDS])

    def __add__(self, other):
        return Photons.join([self, other])

    def __len__(self):
        return len(self.pos)

    def __str__(self):
        if len(self.pos) == 1:
            return ('Photon(pos=%s,dir=%s,pol=%s,wavelength=%s,t=%s,last_hit_triangle=%s,'
                    'flag=%s,weight=%s)' % (self.pos[0], self.dir[0], self.pol[0],
                                            self.wavelengths[0], self.t[0],
                                            self.last_hit_triangles[0], self.flags[0],
                                            self.weights[0]))
        return 'Photons[%d]' % len(self.pos)

    __repr__ = __str__

    def __getitem__(self, key):
        return Photons(*[getattr(self, f)[key] for f in _FIELDS])

    def reduced(self, reduction_factor=1.0):
        n = len(self)
        choice = np.random.permutation(n)[:int(n * reduction_factor)]
        return self[choice]


class PhotonSource(object):
    """Description of a photon bomb that is sampled on the GPU instead of being built on the host
    and uploaded.  Accepted wherever a :class:`Photons` starts an event:
    ``gpu.GPUPhotons(source)``, ``Simulation.simulate([source, ...])``, ``create_pdf`` ...

    The distribution is the host recipe of the reference: directions from
    ``sample.uniform_sphere`` (``phi=None``) or ``sample.flashlight(phi, direction)``
    (chroma/sample.py:4-57), polarisation ``normalize(cross(uniform_sphere, dir))`` and
    wavelengths ``U(lo, hi)`` (chroma/benchmark.py:75-83), a single wavelength when
    `wavelengths` is a number (doc/source/example_detector.py:116-121).  `seed` and `sequence`
    key the counter-based generator together with the photon's global index (container
    ``id_base`` + slot), so sharding a source over GPUs does not change its photons."""

    def __init__(self, nphotons, pos=(0, 0, 0), wavelengths=(400.0, 800.0), direction=(0, 0, 1),
                 phi=None, t0=0.0, weight=1.0, seed=0, sequence=0):
        self.nphotons = int(nphotons)
        self.pos = np.asarray(pos, dtype=np.float32).reshape(3)
        if np.isscalar(wavelengths):
            wavelengths = (wavelengths, wavelengths)
        self.wavelength_lo, self.wavelength_hi = float(wavelengths[0]), float(wavelengths[1])
        self.direction = np.asarray(direction, dtype=np.float64).reshape(3)
        self.phi = None if phi is None else float(phi)
        self.t0, self.weight = float(t0), float(weight)
        self.seed, self.sequence = int(seed), int(sequence)
        self.evidx = 0

    def __len__(self):
        return self.nphotons

    @property
    def cos_half_angle(self):
        return -1.0 if self.phi is None else float(np.clip(np.cos(self.phi), -1.0, 1.0))

    def rotation(self):
        """3x3 matrix taking the cone frame (axis = z) to `direction`: the rotation
        ``flashlight`` applies, angle -arccos(d_z/|d|) about z x d (chroma/sample.py:38-45,
        chroma/transform.py:34-43), written as a matrix."""
        d = self.direction
        if self.phi is None or np.equal(d, (0, 0, 1)).all():
            return np.eye(3)
        axis = np.cross((0, 0, 1), d)
        norm = np.linalg.norm(axis)
        if norm == 0:      # direction = -z: the reference divides by zero here; turn about x
            return np.diag([1.0, -1.0, -1.0])
        n = axis / norm
        angle = -np.arccos(np.dot(d, (0, 0, 1)) / np.linalg.norm(d))
        c, s = np.cos(angle), np.sin(angle)
        # rotate(x) = x c + n (x.n)(1-c) + (x cross n) s
        cross = np.array([[0.0, n[2], -n[1]], [-n[2], 0.0, n[0]], [n[1], -n[0], 0.0]])
        return c * np.eye(3) + (1 - c) * np.outer(n, n) + s * cross


class Channels(object):
    def __init__(self, hit, t, q, flags=None, evidx=None):
        """All channels of the detector: hit bool n, t float32 n [ns], q float32 n."""
        self.hit = hit
        self.t = t
        self.q = q
        self.flags = flags
        self.evidx = evidx

    def hit_channels(self, return_flags=False):
        if return_flags:
            return self.hit.nonzero()[0], self.t[self.hit], self.q[self.hit], self.flags[self.hit]
        return self.hit.nonzero()[0], self.t[self.hit], self.q[self.hit]


class Event(object):
    def __init__(self, id=0, vertices=None, photons_beg=None, photons_end=None,
                 photon_tracks=None, photon_parent_trackids=None, hits=None, flat_hits=None,
                 channels=None):
        self.id = id
        self.nphotons = None
        self.vertices = vertices if vertices is not None else []
        self.photons_beg = photons_beg
        self.photons_end = photons_end
        self.photon_tracks = photon_tracks
        self.photon_parent_trackids = photon_parent_trackids
        self.hits = hits
        self.flat_hits = flat_hits
        self.channels = channels
