"""Input-shaping helpers used by the benchmarks (API of ``chroma/tools.py``)."""
import time

import numpy as np

from chroma_b200.transform import normalize


def profile_if_possible(func):
    "kernprof's @profile when running under it, identity otherwise (chroma/tools.py:54-59)."
    try:
        return profile(func)  # noqa: F821  (injected by kernprof)
    except NameError:
        return func


def timeit(func):
    def f(*args, **kwargs):
        t0 = time.time()
        ret = func(*args, **kwargs)
        print('%s elapsed in %s().' % (time.time() - t0, func.__name__))
        return ret
    return f


def filled_array(value, shape, dtype):
    a = np.empty(shape=shape, dtype=dtype)
    a.fill(value)
    return a


def argsort_direction(dir):
    """Indices that sort unit vectors by a Morton code of (theta, phi), 16 bits each
    (chroma/tools.py:175-193): neighbouring rays end up in the same warp."""
    bits = 16
    maxint = 2 ** bits - 1
    dir = np.asarray(dir)
    theta = (np.arccos(np.clip(dir[:, 2], -1, 1)) / np.pi * maxint).astype(np.uint32)
    phi = ((np.arctan2(dir[:, 1], dir[:, 0]) / np.pi / 2.0 + 0.5) * maxint).astype(np.uint32)
    morton = np.zeros(len(dir), dtype=np.uint32)
    for i in range(bits):
        morton |= ((theta & np.uint32(1 << i)) << np.uint32(i)) | \
                  ((phi & np.uint32(1 << i)) << np.uint32(i + 1))
    return np.argsort(morton, kind='stable')


def from_film(position=(0, 0, 0), axis1=(0, 0, 1), axis2=(1, 0, 0), size=(800, 600),
              width=35.0, focal_length=18.0):
    """Rays through a `size` pixel film with focal point at `position`
    (chroma/tools.py:195-228).  Returns (positions, unit directions)."""
    height = width * (size[1] / float(size[0]))
    axis1 = normalize(axis1)
    axis2 = normalize(axis2)
    dx0 = width / size[0]
    dx1 = height / size[1]
    yy, xx = np.meshgrid(np.arange(size[1]), np.arange(size[0]))
    n = size[0] * size[1]
    grid = -np.tile(axis2, (n, 1)) * xx.ravel()[:, np.newaxis] * dx0 + \
        np.tile(axis1, (n, 1)) * yy.ravel()[:, np.newaxis] * dx1
    grid += axis2 * width / 2 - axis1 * height / 2
    grid -= np.cross(axis1, axis2) * focal_length
    return np.tile(position, (n, 1)), normalize(-grid)
