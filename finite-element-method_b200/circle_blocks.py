# coding=utf-8
"""Obstacles: same constructor and attributes as the reference's CircleBlocks
(circle_blocks.py:6-25): blocks_count, blocks[i].id / .center / .radius."""
import numpy as np


class Block:
    __slots__ = ("id", "center", "radius")

    def __init__(self, id, center, radius):
        self.id = id
        self.center = np.asarray(center, dtype=np.float32)
        self.radius = float(radius)


class CircleBlocks:

    def __init__(self, configs):
        configs = configs or []
        self.blocks_count = len(configs)
        self.blocks = [Block(c.get('id'), c.get('block_center'), c.get('block_radius')) for c in configs]
        self.version = 0

    def centers(self):
        return np.array([b.center for b in self.blocks], dtype=np.float32)

    def radii(self):
        return np.array([b.radius for b in self.blocks], dtype=np.float32)
