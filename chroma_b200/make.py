"""Primitive mesh builders (API of the reference's ``chroma/make.py``): profiles are
traced counter-clockwise so triangle normals point out of the solid."""
import numpy as np

from chroma_b200.geometry import Mesh
from chroma_b200.transform import rotate


def mesh_grid(grid):
    """Triangulate a (rows, cols) grid of vertex ids that wraps around in the column
    direction: two triangles per cell, wound (a, b, b') and (a, b', a') where b is the
    next row and ' the next column."""
    grid = np.asarray(grid)
    a = grid[:-1]
    b = grid[1:]
    a2 = np.roll(a, -1, axis=1)
    b2 = np.roll(b, -1, axis=1)
    first = np.stack([a.ravel(), b.ravel(), b2.ravel()], axis=1)
    second = np.stack([a.ravel(), b2.ravel(), a2.ravel()], axis=1)
    return np.concatenate([first, second])


def linear_extrude(x1, y1, height, x2=None, y2=None, center=None, endcaps=True):
    """Extrude the polygon (x1, y1) along z by `height`, optionally tapering to (x2, y2)."""
    x1, y1 = np.asarray(x1, dtype=float), np.asarray(y1, dtype=float)
    if len(x1) != len(y1):
        raise Exception('`x` and `y` arrays must have the same length.')
    x2 = x1 if x2 is None else np.asarray(x2, dtype=float)
    y2 = y1 if y2 is None else np.asarray(y2, dtype=float)
    if len(x2) != len(y2) or len(x2) != len(x1):
        raise Exception('`x` and `y` arrays must have the same length.')
    n = len(x1)
    rings = [np.column_stack([x1, y1, np.full(n, -height / 2.0)]),
             np.column_stack([x2, y2, np.full(n, height / 2.0)])]
    if endcaps:
        rings = [np.column_stack([np.zeros(n), np.zeros(n), np.full(n, -height / 2.0)])] + rings + \
                [np.column_stack([np.zeros(n), np.zeros(n), np.full(n, height / 2.0)])]
    k = len(rings)
    # vertex id of ring r, polygon point i is i*k + r
    vertices = np.stack(rings, axis=1).reshape(n * k, 3)
    if center is not None:
        vertices = vertices + np.asarray(center, dtype=float)
    grid = np.arange(n * k).reshape(n, k).T[::-1]
    return Mesh(vertices, mesh_grid(grid), remove_duplicate_vertices=True)


def rotate_extrude(x, y, nsteps=64):
    """Revolve the profile (x, y) about the y axis in `nsteps` steps."""
    if len(x) != len(y):
        raise Exception('`x` and `y` arrays must have the same length.')
    points = np.array([x, y, np.zeros(len(x))], dtype=float).T
    steps = np.linspace(0, 2 * np.pi, nsteps, endpoint=False)
    vertices = np.vstack([np.atleast_2d(rotate(points, angle, (0, -1, 0))) for angle in steps])
    grid = np.arange(len(vertices)).reshape(len(steps), len(points)).T[::-1]
    return Mesh(vertices, mesh_grid(grid), remove_duplicate_vertices=True)


def box(dx, dy, dz, center=(0, 0, 0)):
    return linear_extrude([-dx / 2.0, dx / 2.0, dx / 2.0, -dx / 2.0],
                          [-dy / 2.0, -dy / 2.0, dy / 2.0, dy / 2.0], height=dz, center=center)


def cube(size, height=None, center=(0, 0, 0)):
    return box(size, size, size, center=center)


def cylinder(radius, height, radius2=None, nsteps=64):
    if radius2 is None:
        radius2 = radius
    return rotate_extrude([0, radius, radius2, 0],
                          [-height / 2.0, -height / 2.0, height / 2.0, height / 2.0], nsteps)


def sphere(radius, nsteps=64):
    a = np.linspace(-np.pi / 2, np.pi / 2, nsteps)
    return rotate_extrude(radius * np.cos(a), radius * np.sin(a), nsteps)


def torus(radius, offset, nsteps=64, circle_steps=None):
    if circle_steps is None:
        circle_steps = nsteps
    a = np.linspace(0, 2 * np.pi, circle_steps)
    return rotate_extrude(radius * np.cos(a) + offset, radius * np.sin(a), nsteps)


def convex_polygon(x, y):
    vertices = np.column_stack((x, y, np.zeros_like(x)))
    triangles = np.empty(shape=(len(vertices) - 2, 3), dtype=np.int32)
    triangles[:, 0] = 0
    triangles[:, 1] = np.arange(1, len(vertices) - 1)
    triangles[:, 2] = np.arange(2, len(vertices))
    return Mesh(vertices=vertices, triangles=triangles)
