"""The CPU restatement of the reference's draw list (oracle/render_ref.py, render/render.py:64-98): known-answer
checks of colours, draw order and camera geometry.  (No reference pixels exist to pin against: Taichi's GUI /
GGUI cannot run on this box -- parity unpinned, see the oracle's header.)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import render_ref  # noqa: E402


def test_render2d_colours_and_painter_order():
    pos = np.array([[0.25, 0.25], [0.75, 0.25], [0.25, 0.75]], np.float32)
    el = np.array([[0, 1, 2]], np.int32)
    img = render_ref.render2d(pos, el, np.array([0.5], np.float32), [[0.75, 0.75]], [0.1], 64, 64)
    assert img.shape == (64, 64, 3)
    # triangle interior: int((0.5 + 0.13) * 255), int(0.13 * 255)   (ti.rgb_to_hex truncates, render.py:71)
    assert tuple(img[40, 24]) == (160, 33, 33)
    # y is up: vertex 2 at (0.25, 0.75) is drawn in the TOP half as a 0xAAAA00 dot over the triangle
    assert tuple(img[16, 16]) == (0xAA, 0xAA, 0x00)
    # block disc radius 0.1 * 64 px at (0.75, 0.75) -> (48, 16), colour 0x343434; background black
    assert tuple(img[16, 48]) == (0x34, 0x34, 0x34) and tuple(img[16, 56]) == (0, 0, 0)
    assert tuple(img[60, 60]) == (0, 0, 0)
    # a block over the triangle hides it (drawn last, render.py:73-75)
    img2 = render_ref.render2d(pos, el, np.zeros(1, np.float32), [[0.4, 0.4]], [0.3], 64, 64)
    assert tuple(img2[40, 24]) == (0x34, 0x34, 0x34)
    # phi saturates the red channel, negative phi clips at 0
    assert tuple(render_ref.render2d(pos, el, [5.0], [], [], 64, 64)[40, 24]) == (255, 33, 33)
    assert tuple(render_ref.render2d(pos, el, [-5.0], [], [], 64, 64)[40, 24]) == (0, 33, 33)


def test_render3d_camera_and_depth():
    # camera on the +z axis looking at the origin: x right, y up
    cam = render_ref.Camera((0, 0, 10), (0, 0, 0), (0, 1, 0), 90.0, 200, 100)
    x, y, d = cam.project((0, 0, 0))
    assert (x, y, d) == (100.0, 50.0, 10.0)
    x1, y1, _ = cam.project((1, 1, 0))
    assert x1 > 100 and y1 < 50 and abs((x1 - 100) - 50 / 10) < 1e-4      # f = H/2 / tan(45 deg) = 50 px
    assert cam.project((0, 0, 10.5)) is None                              # behind the eye
    pos = np.array([[-1, -1, 0], [1, -1, 0], [0, 1, 0], [-1, -1, -5], [1, -1, -5], [0, 1, -5]], np.float32) + np.float32(2.5)
    view = dict(eye=(2.5, 2.5, 12.0), lookat=(2.5, 2.5, 0.0), up=(0, 1, 0), fov_deg=45.0, W=128, H=128)
    near = render_ref.render3d(pos, [[0, 1, 2]], **view)
    both = render_ref.render3d(pos, [[3, 4, 5], [0, 1, 2]], **view)
    assert (near == (252, 173, 71)).all(axis=2).any()                     # the box is always drawn
    grey = lambda im: (im[..., 0] == im[..., 1]) & (im[..., 1] == im[..., 2]) & (im[..., 0] > 0)
    assert grey(near).sum() > 20
    # the far triangle projects smaller, inside the near one's outline: both visible, and nothing of the near
    # wireframe is lost to it (nearest sample wins regardless of draw order)
    assert (grey(both) & grey(near)).sum() == grey(near).sum() and grey(both).sum() > grey(near).sum()
    # facing away from the light -> ambient only: 0.5 * 0.8 * 255 = 102
    assert near[grey(near)].min() >= 102
