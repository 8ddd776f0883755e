"""CPU: the rasteriser oracle (oracle/render.py) against properties of the scene the reference draws
(environment.py:297-412, rendering.py:223-250).  The reference's own pixels cannot be produced here (pyglet / OpenGL:
unavailable on this box), so these are known-answer checks of the restated rules, not fixtures."""
import numpy as np

from oracle import render as R


def test_unit_circle_is_make_circle():
    u = R.unit_circle(30)
    assert u.shape == (30, 2) and u[0, 0] == 1.0 and u[0, 1] == 0.0
    np.testing.assert_allclose((u ** 2).sum(1), 1.0, rtol=1e-15)


def test_empty_scene_is_white():
    img = R.rasterise([], width=64, height=48)
    assert img.shape == (48, 64, 3) and img.dtype == np.uint8 and (img == 255).all()


def test_opaque_landmark_colour_outline_and_orientation():
    # one opaque entity in the upper-right quadrant: world (+1, +1) -> column 3W/4, row H/4 (row 0 is the top)
    W = H = 200
    prims = R.primitives(np.array([[1.0, 1.0]]), [0.2], np.array([[0.25, 0.5, 0.75, 1.0]], np.float32))
    img = R.rasterise(prims, width=W, height=H)
    cx, cy = 150, 50
    assert tuple(img[cy, cx]) == (64, 128, 191)                       # round(255 c)
    assert (img[150, 50] == 255).all()                                # mirrored position stays background
    # outline: colour and alpha halved (rendering.py:236-237) over the fill -> 0.5*c*0.5 + c*0.5 = 0.75 c
    r_px = 0.2 * W / 4.0
    edge = img[cy, int(round(cx + r_px)) - 1]
    want = np.floor((0.25 * np.array([64, 128, 191]) / 255.0 + 0.5 * np.array([64, 128, 191]) / 255.0) * 255 + 0.5)
    np.testing.assert_array_equal(edge, want.astype(np.uint8))
    # coverage ~ area of the 30-gon plus the outer half of the outline
    covered = (img != 255).any(-1).sum()
    area = 0.5 * 30 * np.sin(2 * np.pi / 30) * r_px ** 2
    assert area < covered < area + 2 * np.pi * r_px * 1.0 + 40


def test_agent_alpha_blend_and_overlap_order():
    # two agents (alpha 0.5) at the same place: second blends over the first, each result rounded to 1/255
    W = H = 100
    rgba = np.array([[1.0, 0.0, 0.0, 0.5], [0.0, 0.0, 1.0, 0.5]], np.float32)
    img = R.rasterise(R.primitives(np.zeros((2, 2)), [0.4, 0.4], rgba), width=W, height=H)
    q = lambda v: np.floor(v * 255 + 0.5) / 255
    first = q(np.array([1.0, 0.0, 0.0]) * 0.5 + 1.0 * 0.5)
    first_edge = first                                                 # centre pixel is far from the outline
    second = q(np.array([0.0, 0.0, 1.0]) * 0.5 + first_edge * 0.5)
    np.testing.assert_array_equal(img[50, 50], np.floor(second * 255 + 0.5).astype(np.uint8))


def test_comm_circles():
    # a speaker with dim_c = 3 saying word 1: three small circles along its diameter, the second one black
    W = H = 300
    size = 0.3
    prims = R.primitives(np.zeros((1, 2)), [size], np.array([[0.25, 0.25, 0.25, 0.5]], np.float32),
                         comm_dim=[3], comm=[1])
    assert len(prims) == 4
    cs = np.float32(size) / 3.0
    np.testing.assert_allclose([p[0] for p in prims[1:]], [-2 * cs, 0.0, 2 * cs], atol=1e-12)
    img = R.rasterise(prims, width=W, height=H)
    px = lambda wx: int((wx + 2.0) * W / 4.0)
    assert tuple(img[H // 2, px(0.0)]) == (0, 0, 0)
    assert tuple(img[H // 2, px(-2 * cs)]) == (255, 255, 255) and tuple(img[H // 2, px(2 * cs)]) == (255, 255, 255)
    assert tuple(img[H // 2 - int(0.8 * size * W / 4), px(0.0)]) == (159, 159, 159)   # agent body: 0.25*.5 + 1*.5
