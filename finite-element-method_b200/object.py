# coding=utf-8
"""Scene object: same constructor, attributes and field names as the reference's
object.Object (object.py:41-93, 337-362), backed by libfemb200 device buffers
instead of Taichi fields.  Dm^-1, volumes and lumped masses are computed on the
GPU at construction (femb_object_create).

Mesh sources (config dict of one entry of "objects"):
  dim == 2: "side_length" + "subdivisions"           (reference, object.py:95-126)
  dim == 3: "obj": path to .obj/.stl                 (reference, object.py:128-161; tetrahedralised
                                                      by femb200.meshio -- TetGen itself is a
                                                      third-party C++ package that is not available)
            "cells": [nx, ny, nz] + "h"              (extension: structured Kuhn tet box)
            "side_length" + "subdivisions", no "obj" (extension, SURVEY.md 8 f1: the 2D keys in 3D --
                                                      a Kuhn cube of subdivisions^3 cells)
  any dim : Object.from_arrays(config, vertices, elements)   (extension: bring your own mesh)
"""
import numpy as np

import constants
from femb200 import _capi, fields, meshgen


class Object:

    def __init__(self, config, vertices=None, elements=None, faces=None, tile_elems=0, device=-1):
        dim = constants.dim
        self.rho = config.get("rho")
        E, nu = config.get('E'), config.get('nu')
        self.mu, self.s_lambda = E / 2 / (1 + nu), E * nu / (1 + nu) / (1 - 2 * nu)   # object.py:47-48
        self.E, self.nu = E, nu
        self.damping = config.get('damping')
        if vertices is None:
            vertices, faces, elements = self.load_obj(config)
        else:
            self.center = np.asarray(config.get('center', [0.0] * dim), dtype=np.float32)
            if faces is None:
                faces = elements if dim == 2 else meshgen.boundary_faces(np.asarray(vertices), elements)
        vertices = np.ascontiguousarray(vertices, dtype=np.float32)
        elements = np.ascontiguousarray(elements, dtype=np.int32)
        self.particle_cnt = vertices.shape[0]
        self.mesh_cnt = faces.shape[0]
        self.element_cnt = elements.shape[0]
        # the reference allocates the linear-system vectors only for the implicit path (object.py:78-93)
        self.implicit = not constants.use_explicit_method
        self._h = _capi.Handle(vertices, elements, self.center, self.rho, self.mu, self.s_lambda, self.damping,
                               implicit=self.implicit, tile_elems=tile_elems, device=device)
        self.particles = fields.Particles(self)
        self.elements = fields.Elements(self, elements)
        self.meshs = fields.Meshs(self, np.asarray(faces, dtype=np.int32))
        self.indices = np.asarray(faces, dtype=np.int32).reshape(-1)           # object.py:65, 371-373
        self.phi = fields.ElementField(self, _capi.F_ELEM_PHI)                  # sized by elements (fixes object.py:68)
        self.U = fields.EnergyScalar(self)
        if self.implicit:
            self.vec_b = fields.VertexField(self, _capi.F_VEC_B, writable=False)
            self.vec_x = fields.VertexField(self, _capi.F_VEC_X)
            self.past_vec_x = fields.VertexField(self, _capi.F_PAST_X)
            # dense matrices are never formed; debug densifiers for small meshes only (object.py:82,91,92)
            self.matrix_A = fields.DenseOperator(self, 'matrix_A')
            self.matrix_AT = fields.DenseOperator(self, 'matrix_AT')
            self.matrix_ATA = fields.DenseOperator(self, 'matrix_ATA')
        # extension (synthetic benchmark meshes): initial state x = X + jitter * h * U(-1, 1)
        if config.get('jitter'):
            rng = np.random.default_rng(config.get('seed', 0))
            h = config.get('h') or (config.get('side_length', 1.0) / max(config.get('subdivisions', 1), 1))
            jit = rng.uniform(-1.0, 1.0, vertices.shape).astype(np.float32)
            pos = (vertices + self.center[None, :].astype(np.float32)).astype(np.float32)
            self._h.set(_capi.F_POS, pos + np.float32(config.get('jitter') * h) * jit)
        self._pending_grad = False   # neo_hookean_1_grad requested, not yet materialised
        self._force_clean = True     # particles.force is known to be all zero
        self._blocks_seen = None
        self.last_solve = None
        if config.get('verbose', True):
            print('Vertex count: {}'.format(self.particle_cnt))
            print('Mesh count: {}'.format(self.mesh_cnt))
            print('Element count: {}'.format(self.element_cnt))

    @classmethod
    def from_arrays(cls, config, vertices, elements, faces=None, **kw):
        return cls(config, vertices=vertices, elements=elements, faces=faces, **kw)

    # -- mesh construction -------------------------------------------------
    def construct_2d_mesh(self, config):
        vertices, faces = meshgen.grid_2d(config.get('side_length'), config.get('subdivisions'))
        self.center = np.asarray(config.get('center'), dtype=np.float32)
        return vertices, faces, faces

    def load_obj(self, config):
        if constants.dim == 2:
            return self.construct_2d_mesh(config)
        self.center = np.asarray(config.get('center'), dtype=np.float32)
        if config.get('cells') is not None:
            nx, ny, nz = config.get('cells')
            vertices, elements = meshgen.kuhn_box(nx, ny, nz, config.get('h'))
        elif config.get('obj') is None and config.get('side_length') and config.get('subdivisions'):
            n = int(config.get('subdivisions'))
            vertices, elements = meshgen.kuhn_box(n, n, n, config.get('side_length') / n)
        else:
            from femb200 import meshio
            vertices, elements, self.surface_source = meshio.tetrahedralize_file(config.get('obj'))
        faces = meshgen.boundary_faces(vertices, elements)
        return vertices, faces, elements

    # -- helpers used by the solver wrappers ---------------------------------
    def _sync_blocks(self, circle_blocks):
        key = (id(circle_blocks), getattr(circle_blocks, 'version', 0), circle_blocks.blocks_count)
        if key != self._blocks_seen:
            self._h.blocks_set(circle_blocks.centers(), circle_blocks.radii())
            self._blocks_seen = key

    def _materialize_force(self):
        if self._pending_grad:
            self._h.explicit_grad()
            self._pending_grad = False
            self._force_clean = False

    # -- output (object.py:322-335) -----------------------------------------
    def update_obj(self):
        self._export_points = self.particles.pos.to_numpy()

    def save_obj(self, file_name):
        pts = getattr(self, '_export_points', None)
        if pts is None:
            pts = self.particles.pos.to_numpy()
        faces = self.indices.reshape(-1, 3)
        used = np.unique(faces)
        remap = np.full(self.particle_cnt, -1, dtype=np.int64)
        remap[used] = np.arange(used.size)
        with open(file_name, 'w') as f:
            for p in pts[used]:
                f.write('v {:.8f} {:.8f} {:.8f}\n'.format(*(list(p) + [0.0])[:3]))
            for t in remap[faces] + 1:
                f.write('f {} {} {}\n'.format(*t))

    def count_inverted(self, pos=None):
        """Number of elements with det F <= 0 (the explicit kernel takes log(det F) unclamped,
        explicit.py:24-25, so these produce NaN): sign(det Ds) != sign(det Dm)."""
        pos = self.particles.pos.to_numpy() if pos is None else pos
        ref = self.particles.ref_pos.to_numpy()
        el = self.elements.vertex_indices.to_numpy()

        def dets(p):
            e = (p[el[:, 1:]] - p[el[:, :1]]).astype(np.float64)
            return np.linalg.det(e)
        return int((dets(pos) * dets(ref) <= 0).sum())

    def sync(self):
        self._h.sync()
