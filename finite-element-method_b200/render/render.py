# coding=utf-8
"""Headless stand-in for the reference's Render (render/render.py:9-106).

The reference opens a Taichi GUI (2D, render.py:26) or a GGUI/Vulkan window (3D, render.py:30).
Graphics APIs are unavailable on this box, so there is no window: `render()` snapshots what the
reference would draw -- vertex positions, per-element energy phi (the triangle colour,
render.py:65-71) and the blocks -- and `finish()` writes them to `<output_dir>/frames.npz`.
`widget` / `camera` exist so that code written against the reference's attributes still imports.
"""
import os

import numpy as np

import constants


class _Headless:
    running = True

    def get_event(self):
        return None

    def is_pressed(self, *keys):
        return False


class Render:
    def __init__(self, config, output_dir='./output'):
        self.resolution = (640, 640)                      # render.py:11-12
        self.is_output_gif = config.get('is_output_gif')  # render.py:15
        self.frame_rate = config.get('output_fps', 60)    # render.py:16
        self.output_dir = output_dir
        self.widget, self.camera = _Headless(), _Headless()
        self.frames = []

    def render(self, soft_objects, circle_blocks):
        for obj in soft_objects:
            frame = {'pos': obj.particles.pos.to_numpy()}
            if constants.dim == 2:
                frame['phi'] = obj.phi.to_numpy()         # colour = (phi + 0.13, 0.13, 0.13), render.py:68-71
            self.frames.append(frame)

    def finish(self):
        if not self.frames:
            return None
        os.makedirs(self.output_dir, exist_ok=True)
        path = os.path.join(self.output_dir, 'frames.npz')
        out = {'pos': np.stack([f['pos'] for f in self.frames])}
        if 'phi' in self.frames[0]:
            out['phi'] = np.stack([f['phi'] for f in self.frames])
        np.savez_compressed(path, **out)
        return path
