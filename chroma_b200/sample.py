"""Random direction samplers (API of the reference's ``chroma/sample.py``)."""
import numpy as np

from chroma_b200.transform import rotate


def uniform_sphere(size=None, dtype=np.double):
    "Isotropic unit vectors; draws (theta, u) from numpy's global RNG like the reference."
    theta = np.random.uniform(0.0, 2 * np.pi, size)
    u = np.random.uniform(-1.0, 1.0, size)
    c = np.sqrt(1 - u ** 2)
    if size is None:
        return np.array([c * np.cos(theta), c * np.sin(theta), u])
    points = np.empty((size, 3), dtype)
    points[:, 0] = c * np.cos(theta)
    points[:, 1] = c * np.sin(theta)
    points[:, 2] = u
    return points


def flashlight(phi=np.pi / 4, direction=(0, 0, 1), size=None, dtype=np.double):
    "Unit vectors uniform in the cone of half-angle `phi` around `direction`."
    theta = np.random.uniform(0.0, 2 * np.pi, size)
    u = np.random.uniform(np.cos(phi), 1, size)
    c = np.sqrt(1 - u ** 2)
    if np.equal(direction, (0, 0, 1)).all():
        axis, angle = (0, 0, 1), 0.0
    else:
        axis = np.cross((0, 0, 1), direction)
        angle = -np.arccos(np.dot(direction, (0, 0, 1)) / np.linalg.norm(direction))
    if size is None:
        return rotate(np.array([c * np.cos(theta), c * np.sin(theta), u]), angle, axis)
    points = np.empty((size, 3), dtype)
    points[:, 0] = c * np.cos(theta)
    points[:, 1] = c * np.sin(theta)
    points[:, 2] = u
    return rotate(points, angle, axis)
