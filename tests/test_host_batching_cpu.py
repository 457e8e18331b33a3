"""Host logic of the photon containers without a device: the sequence of C-ABI calls GPUPhotons.empty /
load / the device photon source issue (recorded through a stand-in for _lib.call), and the source
descriptor they pass.  The calls themselves are exercised on the GPU by tests/test_gpu_source.py."""
import ctypes as C

import numpy as np
import pytest

import fixtures as F
from chroma_b200 import _lib, event
from chroma_b200.gpu import photon


class FakeCtx(object):
    handle = object()

    def synchronize(self):
        pass


@pytest.fixture
def calls(monkeypatch):
    log = []

    def fake_call(name, *args):
        rec = [name]
        for a in args:
            obj = getattr(a, '_obj', None)             # C.byref(x) keeps x in _obj
            if name == 'chroma_photons_create' and isinstance(obj, C.c_void_p):
                obj.value = 0x1000 + len(log)
            if isinstance(obj, _lib.SourceDesc):
                rec.append(dict(pos=list(obj.pos), rot=list(obj.rot), cos=obj.cos_half_angle,
                                wl=(obj.wavelength_lo, obj.wavelength_hi), t0=obj.t0, weight=obj.weight,
                                evidx=obj.evidx, sequence=obj.sequence))
            else:
                rec.append(a)
        log.append(rec)

    monkeypatch.setattr(_lib, 'call', fake_call)
    monkeypatch.setattr(photon, 'current_context', lambda: FakeCtx())
    return log


def test_empty_then_windowed_loads(calls):
    g = photon.GPUPhotons.empty(10, id_base=7)
    assert len(g) == 10 and calls[0][0] == 'chroma_photons_create' and calls[0][2:4] == [10, 7]
    src = event.PhotonSource(4, pos=(1, 2, 3), wavelengths=420.0, phi=0.25, direction=(0, 1, 0), t0=5.0,
                             weight=0.5, seed=99, sequence=3)
    src.evidx = 2
    g.load(0, src)
    name, handle, offset, n, desc, seed = calls[1]
    assert (name, offset, n, seed) == ('chroma_photons_generate', 0, 4, 99)
    assert desc['pos'] == [1.0, 2.0, 3.0] and desc['wl'] == (420.0, 420.0)
    assert abs(desc['cos'] - np.cos(0.25)) < 1e-7 and desc['t0'] == 5.0 and desc['weight'] == 0.5
    assert desc['evidx'] == 2 and desc['sequence'] == 3
    rot = np.array(desc['rot']).reshape(3, 3)
    np.testing.assert_allclose(rot @ np.array([0, 0, 1.0]), [0, 1, 0], atol=1e-6)     # cone axis -> direction
    np.testing.assert_allclose(rot @ rot.T, np.eye(3), atol=1e-6)
    host = F.isotropic_photons(6, seed=1)
    host.evidx[:] = 1
    g.load(4, host)
    rec = calls[2]
    assert rec[0] == 'chroma_photons_upload' and rec[2:4] == [4, 6]
    ptrs = rec[4:]
    assert len(ptrs) == 9 and all(p is not None for p in ptrs[:5]) and ptrs[8] is not None
    assert ptrs[5] is None and ptrs[6] is None and ptrs[7] is None      # last_hit, flags, weights: defaults
    assert len(g._owner.host_arrays) == 1                                # sources kept alive with the container
    g.load(10, event.PhotonSource(0))                                    # empty window: nothing issued
    assert len(calls) == 3
    with pytest.raises(AssertionError):
        g.load(8, event.PhotonSource(3))                                 # past the end


def test_source_constructor_and_copies(calls):
    src = event.PhotonSource(5, seed=4)
    g = photon.GPUPhotons(src, ncopies=3, id_base=100)
    names = [c[0] for c in calls]
    assert names == ['chroma_photons_create', 'chroma_photons_generate', 'chroma_photons_duplicate']
    assert calls[0][2:4] == [15, 100] and calls[1][2:4] == [0, 5] and calls[2][2:4] == [5, 3]
    assert calls[1][4]['cos'] == -1.0 and calls[1][4]['rot'] == list(np.eye(3).ravel())
    assert len(g) == 15 and g.true_nphotons == 5 and g.ncopies == 3
    del calls[:]
    empty = photon.GPUPhotons(event.PhotonSource(0))
    assert [c[0] for c in calls] == ['chroma_photons_create'] and len(empty) == 0


def test_simulation_batches_load_events_into_windows(calls):
    "Simulation.simulate: batching, per-event windows, evidx = index in the batch, one propagate per batch"
    import types
    from chroma_b200.sim import Simulation

    class HostOnlySimulation(Simulation):
        def __init__(self):
            self.detector = object()                      # a plain geometry: no channels, no DAQ
            self.nthreads_per_block, self.max_blocks = 64, 1024
            self.photon_tracking = False
            self.gpu_geometry = types.SimpleNamespace(handle='GEOM')
            self.rng_states = types.SimpleNamespace(handle='RNG')

    sim = HostOnlySimulation()
    a = event.PhotonSource(300, seed=1, sequence=10)
    b = F.isotropic_photons(200, seed=2)
    c = event.PhotonSource(100, seed=1, sequence=11)
    d = F.isotropic_photons(50, seed=3)
    evs = list(sim.simulate([a, b, c, d], keep_hits=False, keep_flat_hits=False, max_steps=7, photons_per_batch=600,
                            evid_start=40))
    assert [ev.id for ev in evs] == [40, 41, 42, 43] and [ev.nphotons for ev in evs] == [300, 200, 100, 50]
    assert all(ev.photons_beg is None for ev in evs)      # keep_photons_beg=False
    ops = [(r[0],) + tuple(r[2:4]) for r in calls if r[0] != 'chroma_photons_destroy']
    assert ops == [('chroma_photons_create', 600, 0),      # batch 1: a, b, c (600 photons reach the batch size)
                   ('chroma_photons_generate', 0, 300),
                   ('chroma_photons_upload', 300, 200),
                   ('chroma_photons_generate', 500, 100),
                   ('chroma_photons_propagate', 'GEOM', 'RNG'),
                   ('chroma_photons_create', 50, 0),       # batch 2: d
                   ('chroma_photons_upload', 0, 50),
                   ('chroma_photons_propagate', 'GEOM', 'RNG')]
    gen = [r for r in calls if r[0] == 'chroma_photons_generate']
    assert [g[4]['evidx'] for g in gen] == [0, 2] and [g[4]['sequence'] for g in gen] == [10, 11]
    assert (b.evidx == 1).all() and (d.evidx == 0).all()
    prop = [r for r in calls if r[0] == 'chroma_photons_propagate']
    assert prop[0][4:9] == [0, 600, 7, 0, 0] and prop[1][4:9] == [0, 50, 7, 0, 0]
