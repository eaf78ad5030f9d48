"""the CUDA rasteriser (femb_render, csrc/kernels_render.cu) against the CPU restatement of the
reference's draw list (oracle/render_ref.py; render/render.py:64-98).

Scan conversion is float arithmetic at pixel centres: nvcc contracts a*b+c into FMAs, numpy does not, so a
pixel exactly on an edge may fall either way.  Tolerance written here: at most 0.1 % of the pixels differ, and
the pixels that are NOT within one pixel of any primitive's outline agree exactly (colour values are exact)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finite-element-method_b200"))
sys.path.insert(0, ROOT)
pytestmark = pytest.mark.gpu

MAX_DIFF_FRAC = 1e-3


def _tri_grid(n, x0=0.3, y0=0.4, h=0.02):
    xs, ys = np.meshgrid(np.arange(n + 1), np.arange(n + 1), indexing="ij")
    v = np.stack([xs.ravel() * h + x0, ys.ravel() * h + y0], 1).astype(np.float32)
    vid = lambda i, j: i * (n + 1) + j
    e = []
    for i in range(n):
        for j in range(n):
            e.append([vid(i, j), vid(i + 1, j), vid(i, j + 1)])
            e.append([vid(i + 1, j), vid(i + 1, j + 1), vid(i, j + 1)])
    return v, np.asarray(e, np.int32)


def _handle2d(v, e):
    from femb200 import _capi
    return _capi.Handle(v, e, (0.0, 0.0), 1000.0, 1e4, 1e4, 1.0, False)


def _dilate(mask):
    out = mask.copy()
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            out |= np.roll(np.roll(mask, dy, 0), dx, 1)
    return out


def _compare(got, ref):
    assert got.shape == ref.shape and got.dtype == np.uint8
    diff = (got != ref).any(axis=2)
    assert diff.mean() <= MAX_DIFF_FRAC, "{} of {} pixels differ".format(diff.sum(), diff.size)
    # outline = where the oracle's picture changes colour; away from it the images are identical
    edges = np.zeros(diff.shape, bool)
    edges[1:, :] |= (ref[1:] != ref[:-1]).any(axis=2)
    edges[:, 1:] |= (ref[:, 1:] != ref[:, :-1]).any(axis=2)
    assert not (diff & ~_dilate(edges)).any()


def test_render2d_matches_oracle():
    from femb200 import _capi
    from oracle import render_ref
    v, e = _tri_grid(8)
    rng = np.random.default_rng(3)
    h = _handle2d(v, e)
    pos = v + rng.uniform(-0.004, 0.004, v.shape).astype(np.float32)
    h.set(_capi.F_POS, pos)
    centers, radii = np.array([[0.5, 0.2], [0.15, 0.45]], np.float32), np.array([0.1, 0.06], np.float32)
    h.blocks_set(centers, radii)
    h.energy_grad()                                   # writes phi (explicit_auto_diff.py:29)
    phi = h.get(_capi.F_ELEM_PHI)
    assert np.abs(phi).max() > 0
    n0 = _capi.launch_count()
    img = h.render(640, 640)
    assert _capi.launch_count() - n0 == 5             # fill, triangles, dots, blocks, resolve
    ref = render_ref.render2d(pos, e, phi, centers, radii, 640, 640)
    _compare(img, ref)
    # exact colours: block centre, a vertex dot, background, a triangle interior
    assert tuple(img[int((1 - 0.2) * 640), int(0.5 * 640)]) == (0x34, 0x34, 0x34)
    px, py = int(pos[0, 0] * 640), int((1 - pos[0, 1]) * 640)
    assert tuple(img[py, px]) == (0xAA, 0xAA, 0x00)
    assert tuple(img[5, 5]) == (0, 0, 0)
    assert (img[..., 1] == 33).any()                  # int(0.13 * 255)


def test_render2d_other_sizes_and_blocks_cover_mesh():
    from femb200 import _capi
    from oracle import render_ref
    v, e = _tri_grid(3, 0.45, 0.45, 0.03)
    h = _handle2d(v, e)
    centers, radii = np.array([[0.5, 0.5]], np.float32), np.array([0.2], np.float32)
    h.blocks_set(centers, radii)
    img = h.render(200, 120)
    assert img.shape == (120, 200, 3)
    ref = render_ref.render2d(v, e, np.zeros(len(e), np.float32), centers, radii, 200, 120)
    _compare(img, ref)
    # the block is drawn last: it hides the whole mesh (render.py:73-75 after :70-72)
    assert not (img == (0xAA, 0xAA, 0x00)).all(axis=2).any()
    # an image smaller than the last one reuses the buffers; a bigger one grows them
    assert h.render(64, 64).shape == (64, 64, 3)
    assert h.render(700, 300).shape == (300, 700, 3)


def _box_mesh():
    from femb200 import meshgen
    v, el = meshgen.kuhn_box(4, 3, 5, 0.25)
    faces = meshgen.boundary_faces(v, el)
    return v.astype(np.float32), el.astype(np.int32), np.asarray(faces, np.int32).reshape(-1, 3)


def test_render3d_matches_oracle():
    from femb200 import _capi
    from oracle import render_ref
    v, el, faces = _box_mesh()
    center = (2.0, 1.5, 2.0)
    h = _capi.Handle(v, el, center, 1000.0, 1e4, 1e4, 1.0, False)
    pos = h.get(_capi.F_POS)
    img = h.render(640, 640, prims=faces, prims_version=1)
    ref = render_ref.render3d(pos, faces)
    _compare(img, ref)
    assert (img == (252, 173, 71)).all(axis=2).any()        # box colour (0.99, 0.68, 0.28) * 255 rounded
    assert ((img[..., 0] == img[..., 1]) & (img[..., 0] >= 102)).any()   # grey wireframe, >= 0.5 * 0.8
    # other camera, smaller frame; the mesh moves with the positions
    cam = dict(eye=(8.0, 6.0, 9.0), lookat=(2.5, 2.0, 2.5), up=(0.0, 1.0, 0.0), fov_deg=60.0)
    h.set(_capi.F_POS, pos + np.float32(0.3))
    img2 = h.render(320, 240, prims=faces, prims_version=1, **cam)
    ref2 = render_ref.render3d(pos + np.float32(0.3), faces, W=320, H=240, **cam)
    _compare(img2, ref2)


def test_render_errors():
    from femb200 import _capi
    v, e = _tri_grid(2)
    h = _handle2d(v, e)
    with pytest.raises(RuntimeError):
        h.render(0, 10)
    with pytest.raises(RuntimeError):
        h.render(64, 64, prims=e[:3])                 # 2D primitives are the elements
    v3, el, faces = _box_mesh()
    h3 = _capi.Handle(v3, el, (1, 1, 1), 1000.0, 1e4, 1e4, 1.0, False)
    bad = faces.copy()
    bad[0, 0] = len(v3)
    with pytest.raises(RuntimeError):
        h3.render(64, 64, prims=bad, prims_version=2)
    with pytest.raises(RuntimeError):
        h3.render(64, 64, prims=faces, prims_version=3, eye=(1, 1, 1), lookat=(1, 1, 1))
    assert h3.render(64, 64, prims=None).shape == (64, 64, 3)     # box only


def test_headless_render_class_writes_png_and_gif(tmp_path):
    import subprocess
    pkg = os.path.join(ROOT, "finite-element-method_b200")
    out = tmp_path / "out"
    r = subprocess.run([sys.executable, "main.py", "--config", "default.json", "--out", str(out), "--frames", "12"],
                       cwd=pkg, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    from PIL import Image
    last = np.asarray(Image.open(out / "last.png"))
    assert last.shape == (640, 640, 3) and (last == (0xAA, 0xAA, 0x00)).all(axis=2).any()
    gif = Image.open(out / "video.gif")
    # 12 frames x 10 steps x 5e-4 s = 0.06 s of virtual time at 60 fps: frames at t > 0, 1/60, 2/60, 3/60
    assert gif.n_frames in (3, 4)
