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

Multi-GPU (extension; the reference is single-device, main.py:40): when the process runs under
torch.distributed with WORLD_SIZE > 1 (one process per GPU, `torchrun main.py ...`), a STRUCTURED 3D config
("cells"+"h" or "side_length"+"subdivisions") is cut into x-slabs of cell layers by Object itself: this rank
builds and owns only its slab, binds the peer-memory halo exchange (femb200.dist.SlabHalo) and the solver calls
(neo_hookean_1_grad + kinematic, implicit_solver_neo_hookean, advect_implicit) work unchanged -- the halo sums
and the dot-product sums happen inside the library's kernels.  Fields are LOCAL to the rank
(`particles.pos.to_numpy()` returns this slab, shared interface planes included on both owners);
`obj.gather(obj.particles.pos)` assembles the global array on every rank.  Unstructured meshes are far too small
to shard: every rank then simply holds a replica.
"""
import numpy as np

import constants
from femb200 import _capi, fields, meshgen


class Object:

    def __init__(self, config, vertices=None, elements=None, faces=None, tile_elems=0, device=-1, partition=True):
        dim = constants.dim
        self._slab = None            # (rank, world, i0, i1, (nx, ny, nz)) when this object is one x-slab of a box
        self._halo = None
        self._partition = partition
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
        halo_kw = {}
        if self._slab is not None:
            from femb200 import dist as fdist
            rank, world, i0, i1, (nx, ny, nz) = self._slab
            self._halo = fdist.SlabHalo(rank, world, i1 - i0, ny, nz)
            halo_kw = dict(halo_verts=self._halo.vertex_ids, halo_owned=self._halo.owned)
            device = fdist.local_device() if device < 0 else device
        self._h = _capi.Handle(vertices, elements, self.center, self.rho, self.mu, self.s_lambda, self.damping,
                               implicit=self.implicit, tile_elems=tile_elems, device=device, **halo_kw)
        if self._halo is not None:
            # graph capture needs a real stream (the legacy default stream cannot be captured)
            self._stream = fdist.make_stream(self._h)
            self._halo.attach(self._h, mode=config.get('comm'))
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
            if self._slab is None:
                jit = rng.uniform(-1.0, 1.0, vertices.shape).astype(np.float32)
            else:
                # drawn for the GLOBAL vertex array so that both owners of a shared plane see the same values
                _, _, i0, i1, (nx, ny, nz) = self._slab
                plane = (ny + 1) * (nz + 1)
                jit = rng.uniform(-1.0, 1.0, ((nx + 1) * plane, 3)).astype(np.float32)[i0 * plane:(i1 + 1) * plane]
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
        box = None
        if config.get('cells') is not None:
            box = tuple(int(c) for c in config.get('cells')) + (float(config.get('h')),)
        elif config.get('obj') is None and config.get('side_length') and config.get('subdivisions'):
            n = int(config.get('subdivisions'))
            box = (n, n, n, config.get('side_length') / n)
        if box is not None:
            nx, ny, nz, h = box
            from femb200 import dist as fdist
            rank, world = fdist.rank_world() if self._partition else (0, 1)
            if world > 1:
                if nx < world:
                    raise ValueError("cannot cut %d cell layers into %d slabs" % (nx, world))
                i0, i1 = fdist.slab_range(nx, rank, world)
                self._slab = (rank, world, i0, i1, (nx, ny, nz))
                vertices, elements = meshgen.kuhn_box(nx, ny, nz, h, x_cells=(i0, i1))
            else:
                vertices, elements = meshgen.kuhn_box(nx, ny, nz, h)
        else:
            from femb200 import meshio
            vertices, elements, self.surface_source = meshio.tetrahedralize_file(config.get('obj'))
        if config.get('surface', True):
            faces = meshgen.boundary_faces(vertices, elements)
            if self._slab is not None:
                faces = self._drop_interface_faces(faces)
        else:                       # extension: skip the surface of huge benchmark meshes ("surface": false)
            faces = np.zeros((0, 3), np.int32)
        return vertices, faces, elements

    def _drop_interface_faces(self, faces):
        """Faces of a slab that lie in a plane shared with a neighbour rank are interior faces of the box."""
        rank, world, i0, i1, (nx, ny, nz) = self._slab
        plane = (ny + 1) * (nz + 1)
        layer = faces // plane
        keep = np.ones(len(faces), bool)
        if rank > 0:
            keep &= ~(layer == 0).all(axis=1)
        if rank < world - 1:
            keep &= ~(layer == (i1 - i0)).all(axis=1)
        return faces[keep]

    # -- helpers used by the solver wrappers ---------------------------------
    def _sync_blocks(self, circle_blocks):
        key = (id(circle_blocks), getattr(circle_blocks, 'version', 0), circle_blocks.blocks_count)
        if key != self._blocks_seen:
            self._h.blocks_set(circle_blocks.centers(), circle_blocks.radii())
            self._blocks_seen = key

    def _materialize_force(self):
        if self._pending_grad:
            if self._halo is not None:
                raise RuntimeError("on a slab-partitioned object the gradient exists only inside the fused step "
                                   "(neo_hookean_1_grad followed by kinematic): particles.force cannot be read or "
                                   "the state rewritten in between")
            self._h.explicit_grad()
            self._pending_grad = False
            self._force_clean = False

    # -- the calls the solver wrappers make (single GPU: the handle; slab: the handle + its halo exchange)
    def _explicit_step(self, dt, g_dir, n=1):
        if self._halo is not None:
            self._halo.explicit_steps(self._h, dt, g_dir, n)
        else:
            self._h.explicit_step(dt, g_dir, n)

    def _implicit_solve(self, dt, method, preconditioned, **kw):
        if self._halo is not None:
            if method == 2:
                raise RuntimeError("jacobi_order 'reference' (the serial sweep, rows ordered by global vertex id) runs "
                                   "on one GPU; slab-partitioned objects use the simultaneous Jacobi or CG")
            return self._halo.implicit_solve(self._h, dt, method, preconditioned, **kw)
        return self._h.implicit_solve(dt, method, preconditioned, **kw)

    @property
    def global_particle_cnt(self):
        if self._slab is None:
            return self.particle_cnt
        _, _, _, _, (nx, ny, nz) = self._slab
        return (nx + 1) * (ny + 1) * (nz + 1)

    def gather(self, field):
        """Global [Nv, dim] array of a vertex field on EVERY rank (slab objects; a no-op copy otherwise).  Shared
        interface planes are bit-identical on both owners; the left owner's copy is taken."""
        local = field.to_numpy()
        if self._slab is None:
            return local
        from femb200 import dist as fdist
        rank, world, i0, i1, (nx, ny, nz) = self._slab
        plane = (ny + 1) * (nz + 1)
        parts = fdist.all_gather_arrays(local if rank == world - 1 else local[:(i1 - i0) * plane])
        return np.concatenate(parts, axis=0)

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
        explicit.py:24-25, so these produce NaN).  Counted on the device from the current positions; a host
        array `pos` is checked with numpy instead (sign(det Ds) != sign(det Dm))."""
        if pos is None:
            return self._h.diagnostics()["inverted_elements"]
        ref = self.particles.ref_pos.to_numpy()
        el = self.elements.vertex_indices.to_numpy()

        def dets(p):
            e = (p[el[:, 1:]] - p[el[:, :1]]).astype(np.float64)
            return np.linalg.det(e)
        return int((dets(pos) * dets(ref) <= 0).sum())

    def sync(self):
        """Wait for every enqueued kernel and (non-blocking) field transfer of this object."""
        self._h.transfers_wait()
        self._h.sync()
