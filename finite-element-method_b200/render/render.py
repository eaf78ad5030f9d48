# coding=utf-8
"""Headless Render (reference: render/render.py:9-106).

The reference opens a Taichi GUI (2D, render.py:26) or a GGUI/Vulkan window (3D, render.py:30); graphics
APIs are unavailable on this box, so there is no window.  The PICTURE is still produced: `render()` has the
library rasterise what render2d / render3d draw (femb_render, csrc/kernels_render.cu) straight from the
positions resident on the GPU into a device frame buffer, and gets back RGB8 pixels:

  2D  triangles coloured (phi + 0.13, 0.13, 0.13), vertex dots radius 2 colour 0xAAAA00, blocks 0x343434
  3D  the (0..5)^3 box lines (0.99, 0.68, 0.28) + the surface mesh wireframe under the reference's camera
      (position (-6.36, 3.49, 2.44), look-at (-5.40, 3.19, 2.43), render.py:37-39), ambient 0.8 and the point
      light at (3.5, 3.5, 3.5)

With `is_output_gif` the video frames are gated at `output_fps` of VIRTUAL time exactly like render.py:77/101 and
`finish()` writes `<output_dir>/video.gif` (the reference's VideoManager.make_video, main.py:131-133);
`save_png=True` also writes each frame as `<output_dir>/frames/NNNNNN.png` (render.py:81); `last.png` and the
per-frame state (`frames.npz`: positions, phi in 2D) are always written.  `widget` and
`camera` exist so that code written against the reference's attributes still runs (no events, never pressed).
"""
import os

import numpy as np

import constants


class _Headless:
    running = True

    def get_event(self, *a):
        return None

    def is_pressed(self, *keys):
        return False


class _Camera(_Headless):
    def __init__(self):
        self.pos = (-6.36, 3.49, 2.44)         # render.py:37
        self.at = (-5.40, 3.19, 2.43)          # render.py:38
        self.upv = (0.0, 1.0, 0.0)             # render.py:39
        self.fov_deg = 45.0                    # GGUI default

    def position(self, x, y, z):
        self.pos = (x, y, z)

    def lookat(self, x, y, z):
        self.at = (x, y, z)

    def up(self, x, y, z):
        self.upv = (x, y, z)

    def fov(self, deg):
        self.fov_deg = float(deg)


class Render:
    def __init__(self, config, output_dir='./output', save_png=False, keep_frames=None):
        self.width, self.height = 640, 640                # render.py:11-12
        self.resolution = (self.width, self.height)
        self.is_output_gif = config.get('is_output_gif')  # render.py:15
        self.output_fps = config.get('output_fps', 60)    # render.py:16
        self.frame_rate = self.output_fps
        self.frame_time = 1.0 / self.output_fps           # render.py:18
        self.output_frame_cnt = 0
        self.output_dir = output_dir
        self.save_png = save_png
        self.widget, self.camera = _Headless(), _Camera()
        self.keep_frames = bool(self.is_output_gif) if keep_frames is None else keep_frames
        self.frames = []            # uint8 [H, W, 3] images kept for the video
        self.last_image = None
        self.rendered = 0
        # positions (+ phi in 2D) per frame -> frames.npz.  Default: on for small scenes only (a 16 M-tet box would
        # copy 32 MB to the host and keep it, every frame); `keep_state` in the config overrides
        self.keep_state = config.get('keep_state', None)
        self.keep_state_max_vertices = 200000
        self.max_frames = int(config.get('max_frames', 2000))   # frames kept in memory for the GIF
        self.states = []

    # -- one frame ------------------------------------------------------------------------------
    def image(self, soft_objects, circle_blocks=None):
        """Rasterise all objects on the GPU; later objects are drawn over earlier ones (render.py:66, 97)."""
        out = None
        for obj in soft_objects:
            if circle_blocks is not None:
                obj._sync_blocks(circle_blocks)
            if constants.dim == 2:
                img = obj._h.render(self.width, self.height)
            else:
                cam = self.camera
                img = obj._h.render(self.width, self.height, prims=obj.indices.reshape(-1, 3), prims_version=1,
                                    eye=cam.pos, lookat=cam.at, up=cam.upv, fov_deg=cam.fov_deg)
            if out is None:
                out = img
            else:
                drawn = img.any(axis=2)
                out[drawn] = img[drawn]
        return out

    def render(self, soft_objects, circle_blocks=None, virtual_time=None, msgs=None):
        """render.py:110-117: every call draws (the reference shows the window each frame); the frame goes
        to the video only when virtual time has passed the next 1/output_fps mark (render.py:77 / :101)."""
        img = self.image(soft_objects, circle_blocks)
        if msgs:
            img = draw_text(img, msgs)       # gui.text(msg) overlay, render.py:108-111 (frame count / time)
        self.last_image = img
        self.rendered += 1
        keep_state = self.keep_state
        if keep_state is None:
            keep_state = sum(getattr(o, "particle_cnt", 0) for o in soft_objects) <= self.keep_state_max_vertices
        if keep_state:
            st = {'pos': np.concatenate([o.particles.pos.to_numpy() for o in soft_objects])}
            if constants.dim == 2:
                st['phi'] = np.concatenate([o.phi.to_numpy() for o in soft_objects])
            self.states.append(st)
        due = virtual_time is None or (virtual_time / self.frame_time) > self.output_frame_cnt
        if due and (self.keep_frames or self.save_png):
            if self.keep_frames and len(self.frames) < self.max_frames:
                self.frames.append(img)
            if self.save_png:
                from PIL import Image
                d = os.path.join(self.output_dir, 'frames')
                os.makedirs(d, exist_ok=True)
                Image.fromarray(img).save(os.path.join(d, '{:06d}.png'.format(self.output_frame_cnt)))
            self.output_frame_cnt += 1
        return img

    # -- end of run -------------------------------------------------------------------------------
    def finish(self):
        """Write what the reference's make_video would (main.py:131-133): video.gif; always last.png."""
        if self.last_image is None:
            return None
        from PIL import Image
        os.makedirs(self.output_dir, exist_ok=True)
        path = os.path.join(self.output_dir, 'last.png')
        Image.fromarray(self.last_image).save(path)
        if self.states:
            out = {k: np.stack([f[k] for f in self.states]) for k in self.states[0]}
            np.savez_compressed(os.path.join(self.output_dir, 'frames.npz'), **out)
        if self.frames and self.is_output_gif:
            path = os.path.join(self.output_dir, 'video.gif')
            ims = [Image.fromarray(f) for f in self.frames]
            ims[0].save(path, save_all=True, append_images=ims[1:], duration=max(1, int(round(1000.0 / self.output_fps))),
                        loop=0)
        return path


def draw_text(img, msgs, origin=(6, 6), color=(255, 255, 255)):
    """The reference's `gui.text(msg, pos)` overlay (render/render.py:108-111: one line per message from the top
    left): PIL's built-in bitmap font blitted on the host into the finished frame -- no graphics stack involved."""
    from PIL import Image, ImageDraw
    im = Image.fromarray(img)
    d = ImageDraw.Draw(im)
    x, y = origin
    for m in msgs:
        d.text((x, y), str(m), fill=tuple(color))
        y += 14
    return np.asarray(im).copy()
