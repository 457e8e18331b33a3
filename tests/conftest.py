import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def ctx():
    "One CUDA context for the whole GPU session."
    from chroma_b200 import gpu
    c = gpu.create_cuda_context()
    yield c
    c.pop()


@pytest.fixture(scope='session')
def oracle_mod():
    from oracle import oracle
    oracle.load()
    return oracle
