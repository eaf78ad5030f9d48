"""Mesh generators and mesh utilities used by object.Object.

* grid_2d: the reference's regular 2D triangle grid (object.py:95-126), vectorised.
* kuhn_box: structured Kuhn/Freudenthal tet box -- the synthetic large-mesh input of
  BASELINE configs 3/4; optional x-slab selection for one rank of a
  slab-partitioned multi-GPU run.
* boundary_faces: surface triangles of a tet mesh, outward oriented (object.py:255-308).
"""
import itertools

import numpy as np


def grid_2d(side_length, subdivisions):
    s = int(subdivisions)
    t = np.linspace(0, side_length, s + 1)
    # vertex id i*(s+1)+j -> (x_i, y_j)
    vertices = np.stack(np.meshgrid(t, t, indexing="ij"), axis=-1).reshape(-1, 2)
    i, j = np.meshgrid(np.arange(s), np.arange(s), indexing="ij")
    p1 = (i * (s + 1) + j).ravel()
    p2, p3 = p1 + 1, p1 + s + 1
    p4 = p3 + 1
    faces = np.empty((s * s, 2, 3), dtype=np.int32)
    faces[:, 0] = np.stack([p1, p2, p4], axis=1)
    faces[:, 1] = np.stack([p1, p4, p3], axis=1)
    return vertices, faces.reshape(-1, 3)


def kuhn_box(nx, ny, nz, h, x_cells=None, dtype=np.float32):
    """Vertex id (i*(ny+1)+j)*(nz+1)+k at (i,j,k)*h; cell -> 6 tets
    [v000, v000+e_p1, v000+e_p1+e_p2, v111] over the permutations p of the axes.
    x_cells=(i0, i1) builds only cell layers i0..i1-1 (vertex planes i0..i1) with
    LOCAL vertex ids, coordinates still global."""
    i0, i1 = (0, nx) if x_cells is None else x_cells
    lx = i1 - i0
    sy, sz = (ny + 1) * (nz + 1), nz + 1
    gi, gj, gk = np.meshgrid(np.arange(i0, i1 + 1), np.arange(ny + 1), np.arange(nz + 1), indexing="ij")
    verts = (np.stack([gi, gj, gk], axis=-1).reshape(-1, 3) * float(h)).astype(dtype)
    ci, cj, ck = np.meshgrid(np.arange(lx), np.arange(ny), np.arange(nz), indexing="ij")
    v000 = (ci * sy + cj * sz + ck).ravel().astype(np.int64)
    stride = (sy, sz, 1)
    tets = np.empty((v000.size, 6, 4), dtype=np.int32)
    for t, p in enumerate(itertools.permutations(range(3))):
        a = v000 + stride[p[0]]
        b = a + stride[p[1]]
        tets[:, t, 0], tets[:, t, 1], tets[:, t, 2], tets[:, t, 3] = v000, a, b, b + stride[p[2]]
    return verts, tets.reshape(-1, 4)


def boundary_faces(vertices, tets):
    """Faces that belong to exactly one tet, wound so the normal points away from the
    tet's fourth vertex (same rule as object.py:286-297)."""
    tets = np.asarray(tets)
    combos = np.array([[0, 1, 2], [0, 1, 3], [0, 2, 3], [1, 2, 3]])
    opp = np.array([3, 2, 1, 0])
    faces = tets[:, combos].reshape(-1, 3)
    inner = tets[:, opp].reshape(-1)
    key = np.sort(faces, axis=1)
    _, first, counts = np.unique(key, axis=0, return_index=True, return_counts=True)
    keep = first[counts == 1]
    f, inn = key[keep], inner[keep]
    p0, p1, p2, pi = vertices[f[:, 0]], vertices[f[:, 1]], vertices[f[:, 2]], vertices[inn]
    flip = np.einsum("ij,ij->i", np.cross(p1 - p0, p2 - p0), pi - p0) >= 0
    f[flip] = f[flip][:, [0, 2, 1]]
    return f.astype(np.int32)
