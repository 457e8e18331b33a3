"""CPU test of the oracle's restatement of the ray-cast render (chroma/cuda/render.cu:11-181)
against a hand-computed pixel: a ray through a cube from outside crosses two faces head-on."""
import numpy as np
import pytest

import fixtures as F


@pytest.fixture(scope='module')
def cube(oracle_mod):
    geo = F.attach_bvh(F.cube_geometry(1000.0), oracle_mod.build_bvh)
    return geo, oracle_mod.Oracle(geo)


def blend(layers, bg, alpha_depth):
    "render.cu:143-178 in float32"
    f = np.float32
    scale, rgb = f(1.0), [f(0.0)] * 3
    for (r, g, b, a) in layers:
        rgb = [f(c + f(f(scale * f(v)) * f(a))) for c, v in zip(rgb, (r, g, b))]
        scale = f(scale * f(f(1.0) - f(a)))
    alpha = f(((bg >> 24) & 0xff) / 255.0)
    for i, sh in enumerate((16, 8, 0)):
        rgb[i] = f(rgb[i] + f(f(scale * f((bg >> sh) & 0xff)) * alpha))
    scale = f(scale * f(f(1.0) - alpha))
    a = int(np.floor(f(255) * f(f(1.0) - scale))) if len(layers) < alpha_depth else 255
    out = [int(np.floor(f(c / f(f(1.0) - scale)))) for c in rgb]
    return a << 24 | out[0] << 16 | out[1] << 8 | out[2]


def test_render_two_faces_head_on(cube):
    geo, O = cube
    geo.colors[:] = 0x33336699              # alpha byte 0x33 -> opacity (255-51)/255 = 0.8
    O.keep['colors'][:] = geo.colors
    pos = np.array([[-2000.0, 10.0, 20.0], [-2000.0, 5000.0, 0.0]], np.float32)
    dir = np.array([[1.0, 0.0, 0.0], [1.0, 0.0, 0.0]], np.float32)
    px, (dx, dxlen, color) = O.render(pos, dir, alpha_depth=10, bg_color=0x00000000)
    assert list(dxlen) == [2, 0]
    np.testing.assert_allclose(dx[0, :2], [1500.0, 2500.0], rtol=1e-6)
    np.testing.assert_allclose(color[0, :2], [[0x33, 0x66, 0x99, 0.8]] * 2, rtol=1e-6)
    assert px[1] == 0                                      # missed: background
    assert px[0] == blend([(0x33, 0x66, 0x99, np.float32(0.8))] * 2, 0, 10)
    # alpha_depth 1 keeps the nearer face only and reports full alpha
    px1, (dx1, dxlen1, _) = O.render(pos, dir, alpha_depth=1, bg_color=0x00000000)
    assert dxlen1[0] == 1 and dx1[0, 0] == dx[0, 0] and (px1[0] >> 24) == 255
    # a background with alpha is blended behind the faces and returned where nothing is hit
    px2, _ = O.render(pos, dir, alpha_depth=10, bg_color=0xFF204060)
    assert px2[1] == 0xFF204060
    assert px2[0] == blend([(0x33, 0x66, 0x99, np.float32(0.8))] * 2, 0xFF204060, 10)


def test_render_keep_last_render_merges_lists(cube):
    "a second pass over the same geometry with the lists kept doubles every entry, new in front"
    geo, O = cube
    pos = np.array([[-2000.0, 10.0, 20.0]], np.float32)
    dir = np.array([[1.0, 0.0, 0.0]], np.float32)
    _, state = O.render(pos, dir, alpha_depth=3)
    _, (dx, dxlen, _) = O.render(pos, dir, alpha_depth=3, state=state)
    assert dxlen[0] == 3
    np.testing.assert_allclose(dx[0], [1500.0, 1500.0, 2500.0], rtol=1e-6)
