"""Host logic of the headless Render (finite-element-method_b200/render/render.py) without a GPU: the output_fps
gate of render.py:77/101 on VIRTUAL time, compositing of several objects, and the files finish() writes.  The
rasteriser itself is covered by tests/test_gpu_render.py."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finite-element-method_b200"))


class _Field:
    def __init__(self, a):
        self.a = a

    def to_numpy(self):
        return self.a


class _FakeHandle:
    def __init__(self, colour, box):
        self.colour, self.box, self.calls = colour, box, []

    def render(self, width, height, **kw):
        self.calls.append(kw)
        img = np.zeros((height, width, 3), np.uint8)
        (y0, y1), (x0, x1) = self.box
        img[y0:y1, x0:x1] = self.colour
        return img


class _FakeObject:
    def __init__(self, colour, box, n=5, dim=2):
        self._h = _FakeHandle(colour, box)
        self.particles = type("P", (), {})()
        self.particles.pos = _Field(np.zeros((n, dim), np.float32))
        self.phi = _Field(np.zeros(2 * n, np.float32))
        self.indices = np.arange(6, dtype=np.int32)
        self.synced = 0

    def _sync_blocks(self, blocks):
        self.synced += 1


def test_gif_gate_compositing_and_outputs(tmp_path):
    import constants
    from PIL import Image
    from render.render import Render

    constants.dim = 2
    r = Render({"is_output_gif": True, "output_fps": 60}, output_dir=str(tmp_path), save_png=True)
    assert (r.width, r.height) == (640, 640) and r.widget.running and not r.widget.is_pressed("x")
    a = _FakeObject((200, 10, 10), ((0, 100), (0, 100)))
    b = _FakeObject((10, 200, 10), ((50, 150), (50, 150)))
    dt_frame = 10 * 5e-4                     # default.json: sim_count 10 x delta_time 5e-4 per rendered frame
    vt = 0.0
    for i in range(12):
        vt += dt_frame
        img = r.render([a, b], object(), vt, [])
        assert img.shape == (640, 640, 3)
        assert tuple(img[10, 10]) == (200, 10, 10) and tuple(img[120, 120]) == (10, 200, 10)
        assert tuple(img[75, 75]) == (10, 200, 10)              # later objects are drawn over earlier ones
        assert tuple(img[300, 300]) == (0, 0, 0)
    assert a.synced == 12 and len(a._h.calls) == 12             # every frame is drawn (the reference shows it) ...
    # ... but only frames past the next 1/60 s mark of virtual time go to the video (render.py:77)
    vt2, cnt = 0.0, 0
    for i in range(12):
        vt2 += dt_frame
        if vt2 / (1.0 / 60) > cnt:
            cnt += 1
    assert len(r.frames) == cnt == r.output_frame_cnt and 3 <= cnt <= 4
    assert sorted(os.listdir(tmp_path / "frames")) == ["%06d.png" % k for k in range(cnt)]
    path = r.finish()
    assert path.endswith("video.gif") and Image.open(path).n_frames >= 1
    last = np.asarray(Image.open(tmp_path / "last.png"))
    assert tuple(last[10, 10]) == (200, 10, 10)
    st = np.load(tmp_path / "frames.npz")
    assert st["pos"].shape == (12, 10, 2) and st["phi"].shape == (12, 20)


def test_no_gif_config_and_3d_camera_arguments(tmp_path):
    import constants
    from render.render import Render

    constants.dim = 3
    r = Render({"is_output_gif": False}, output_dir=str(tmp_path))
    assert r.output_fps == 60 and not r.keep_frames
    o3 = _FakeObject((9, 9, 9), ((0, 4), (0, 4)), dim=3)
    r.camera.position(1.0, 2.0, 3.0)
    r.camera.lookat(0.0, 0.0, 0.0)
    assert r.render([o3]) is not None                           # no blocks, no virtual time: still draws
    kw = o3._h.calls[0]
    assert kw["eye"] == (1.0, 2.0, 3.0) and kw["lookat"] == (0.0, 0.0, 0.0) and kw["up"] == (0.0, 1.0, 0.0)
    assert kw["fov_deg"] == 45.0 and kw["prims"].shape == (2, 3)
    assert o3.synced == 0 and r.frames == []
    r.finish()
    assert sorted(os.listdir(tmp_path)) == ["frames.npz", "last.png"]
    assert Render({"is_output_gif": False}, output_dir=str(tmp_path / "none")).finish() is None


def test_text_overlay_draws_the_messages():
    """render/render.py:108-111: gui.text(msg) per message, top left."""
    from render.render import draw_text
    img = np.zeros((64, 200, 3), np.uint8)
    out = draw_text(img, ["frame_cnt: 7", "time: 0.0350"])
    assert out.shape == img.shape and out.dtype == np.uint8
    assert out[:34].any() and not out[40:].any()          # two text lines near the top, nothing below
    assert not img.any()                                   # the input frame is not modified
