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
            if isinstance(obj, _lib.PhotonArrays):
                rec.append(dict(ptrs={f: getattr(obj, f) for f in ('pos', 'dir', 'pol', 'wavelengths', 't', 'weights',
                                                                   'last_hit', 'flags', 'evidx')},
                                pos_value=list(obj.pos_value), t_value=obj.t_value, weight_value=obj.weight_value,
                                flags_value=obj.flags_value, evidx_value=obj.evidx_value,
                                last_hit_value=obj.last_hit_value, borrow=obj.borrow_sources))
            elif isinstance(obj, _lib.SourceDesc):
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
    assert rec[0] == 'chroma_photons_upload_arrays' and rec[2:4] == [4, 6]
    a = rec[4]
    assert all(a['ptrs'][f] is not None for f in ('pos', 'dir', 'pol', 'wavelengths', 't'))
    # last_hit, weights: defaults (chroma/sim.py:70-72); flags and the photons' own evidx are uploaded
    assert all(a['ptrs'][f] is None for f in ('last_hit', 'weights')) and a['ptrs']['flags'] is not None
    assert a['ptrs']['evidx'] is not None and (a['last_hit_value'], a['weight_value']) == (-1, 1.0)
    assert len(g._owner.host_arrays) == 1                                # sources kept alive with the container
    assert a['borrow'] == 0                                              # a plain load: the source may be reused at once
    # generator flags are kept (copy_flags stays True in chroma/sim.py:72); a bomb's vertex and start time
    # given as broadcast views are passed by value; an explicit evidx overrides the photons' own
    bomb = F.isotropic_photons(6, seed=2)
    bomb.flags[:] = event.CHERENKOV
    bomb.pos = np.broadcast_to(np.array([1, 2, 3], np.float32), (6, 3))
    bomb.t = np.broadcast_to(np.float32(100.0), (6,))
    g.load(4, bomb, evidx=5)
    a = calls[3][4]
    assert a['ptrs']['flags'] is not None and a['ptrs']['pos'] is None and a['ptrs']['t'] is None
    assert a['pos_value'] == [1.0, 2.0, 3.0] and a['t_value'] == 100.0 and a['evidx_value'] == 5
    del calls[3]
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
                   ('chroma_photons_upload_arrays', 300, 200),
                   ('chroma_photons_generate', 500, 100),
                   ('chroma_photons_propagate', 'GEOM', 'RNG'),
                   ('chroma_photons_create', 50, 0),       # batch 2: d
                   ('chroma_photons_upload_arrays', 0, 50),
                   ('chroma_photons_propagate', 'GEOM', 'RNG')]
    gen = [r for r in calls if r[0] == 'chroma_photons_generate']
    assert [g[4]['evidx'] for g in gen] == [0, 2] and [g[4]['sequence'] for g in gen] == [10, 11]
    up = [r for r in calls if r[0] == 'chroma_photons_upload_arrays']
    assert [u[4]['evidx_value'] for u in up] == [1, 0] and all(u[4]['ptrs']['evidx'] is None for u in up)
    # Simulation lends the events' arrays to the engine until its propagate() returns (borrow_sources)
    assert all(u[4]['borrow'] == 1 for u in up)
    assert (b.evidx == 0).all() and (d.evidx == 0).all()          # the events' own arrays are not written
    prop = [r for r in calls if r[0] == 'chroma_photons_propagate']
    assert prop[0][4:9] == [0, 600, 7, 0, 0] and prop[1][4:9] == [0, 50, 7, 0, 0]
