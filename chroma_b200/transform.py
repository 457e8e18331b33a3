"""Vector helpers (API of the reference's ``chroma/transform.py``)."""
import numpy as np


def norm(x):
    "Euclidean norm along the last axis."
    x = np.asarray(x)
    return np.sqrt((x * x).sum(-1))


def normalize(x):
    "Unit vectors in the direction of `x` (one vector or an (n,3) array)."
    x = np.atleast_2d(np.asarray(x, dtype=float))
    return (x / norm(x)[:, np.newaxis]).squeeze()


def get_perp(x):
    "An arbitrary vector perpendicular to `x`."
    a = np.zeros(3)
    a[np.argmin(abs(np.asarray(x)))] = 1
    return np.cross(a, x)


def make_rotation_matrix(phi, n):
    """Matrix rotating points by `phi` counter-clockwise about axis `n` when looking towards
    +infinity (Rodrigues' formula, same convention as chroma/transform.py:10-22: points are
    transformed as ``np.inner(points, matrix)``)."""
    n = normalize(n)
    c, s = np.cos(phi), np.sin(phi)
    skew = np.array([[0, n[2], -n[1]], [-n[2], 0, n[0]], [n[1], -n[0], 0]])
    return c * np.identity(3) + (1 - c) * np.outer(n, n) + s * skew


def rotate(x, phi, n):
    "Rotate points `x` by angle(s) `phi` about axis `n` (chroma/transform.py:34-43)."
    n = normalize(n)
    x = np.atleast_2d(x)
    phi = np.atleast_1d(phi)
    c, s = np.cos(phi)[:, np.newaxis], np.sin(phi)[:, np.newaxis]
    return (x * c + n * np.dot(x, n)[:, np.newaxis] * (1 - c) + np.cross(x, n) * s).squeeze()


def rotate_matrix(x, phi, n):
    return np.inner(np.asarray(x), make_rotation_matrix(phi, n))
