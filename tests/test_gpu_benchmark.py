"""The chroma.benchmark mirror (chroma_b200/benchmark.py) runs every loop of chroma/benchmark.py on a
small detector and returns finite, positive rates."""
import numpy as np
import pytest

import fixtures as F
from chroma_b200 import benchmark, event, gpu

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def gpu_detector(ctx, oracle_mod):
    det = F.attach_bvh(F.small_detector(), oracle_mod.build_bvh)
    return gpu.GPUDetector(det)


def ok(rate):
    mean, std = rate
    return np.isfinite(mean) and mean > 0 and np.isfinite(std) and std >= 0


def test_ray_load_generate_propagate(gpu_detector):
    assert ok(benchmark.intersect(gpu_detector, number=3, nphotons=20000))
    assert ok(benchmark.load_photons(number=3, nphotons=20000))
    assert ok(benchmark.generate_photons(number=3, nphotons=20000))
    assert ok(benchmark.propagate(gpu_detector, number=3, nphotons=20000))


def test_pdf_and_pdf_eval(gpu_detector):
    def events(n):
        return benchmark.bomb_events(n, photons_per_event=5000, seed=2)
    assert ok(benchmark.pdf(gpu_detector, npdfs=3, nevents=4, nreps=2, ndaq=2, events=events))
    assert ok(benchmark.pdf_eval(gpu_detector, npdfs=3, nevents=3, nreps=2, ndaq=64, events=events))
    # host-built events work in the same loops
    host = [event.Event(photons_beg=F.isotropic_photons(3000, seed=k, wavelength_range=(400, 800))) for k in range(3)]
    assert ok(benchmark.pdf(gpu_detector, npdfs=2, nevents=3, nreps=1, events=lambda n: iter(host[:n])))


def test_rate_helper():
    assert benchmark._rate(10, [1.0, 1.0]) == (10.0, 0.0)
    assert np.isnan(benchmark._rate(10, [])[0])
    assert 'e' in benchmark.rate_to_str((1.5e8, 2e6)) or '+/-' in benchmark.rate_to_str((1.5e8, 2e6))
