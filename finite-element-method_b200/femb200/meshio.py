"""Surface-mesh ingestion and tetrahedralisation for the 3D `obj` config key (reference object.py:128-161).

The reference reads the surface with trimesh / pyvista and meshes it with TetGen
(`tet.tetrahedralize(order=1, mindihedral=10, minratio=5.0)`, object.py:141).  None of the three
is a dependency here: `.obj` and binary/ascii `.stl` are parsed directly, and the volume mesh is
a constrained-by-rejection Delaunay mesh (scipy): surface vertices first (TetGen also keeps the
input points first -- the reference's export relies on it, object.py:311-327), then interior
lattice points, keeping the tets whose centroid lies inside the surface (generalised winding
number).  Real TetGen is used when `tetgen` + `pyvista` are importable.  The solver only needs
*some* conforming tet mesh: parity tests feed identical (vertices, tets) to both sides.
"""
import os
import struct

import numpy as np


def read_obj(path):
    """Vertices [n,3] float64 and triangles [m,3] int32 of a Wavefront .obj (polygons are fanned)."""
    verts, faces = [], []
    with open(path, "r", errors="replace") as f:
        for line in f:
            if line.startswith("v "):
                p = line.split()
                verts.append((float(p[1]), float(p[2]), float(p[3])))
            elif line.startswith("f "):
                idx = []
                for tok in line.split()[1:]:
                    i = int(tok.split("/")[0])
                    idx.append(i - 1 if i > 0 else len(verts) + i)
                for k in range(1, len(idx) - 1):
                    faces.append((idx[0], idx[k], idx[k + 1]))
    if not verts or not faces:
        raise ValueError("%s: no vertices/faces found" % path)
    return np.asarray(verts, dtype=np.float64), np.asarray(faces, dtype=np.int32)


def read_stl(path):
    """Binary or ascii STL -> welded vertices [n,3] and triangles [m,3]."""
    with open(path, "rb") as f:
        data = f.read()
    tris = None
    if len(data) >= 84:
        n = struct.unpack_from("<I", data, 80)[0]
        if 84 + 50 * n == len(data):
            rec = np.frombuffer(data, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=n,
                                offset=84)
            tris = rec["v"].astype(np.float64)
    if tris is None:
        pts = [tuple(float(x) for x in ln.split()[1:4]) for ln in data.decode("ascii", "replace").splitlines()
               if ln.strip().startswith("vertex")]
        tris = np.asarray(pts, dtype=np.float64).reshape(-1, 3, 3)
    if tris.size == 0:
        raise ValueError("%s: empty STL" % path)
    verts, inv = np.unique(tris.reshape(-1, 3).round(9), axis=0, return_inverse=True)
    return verts, inv.reshape(-1, 3).astype(np.int32)


def read_surface(path):
    ext = os.path.splitext(path)[1].lower()
    if ext == ".stl":
        return read_stl(path)
    return read_obj(path)


def weld(verts, faces, tol=1e-9):
    """Merge duplicate points (the reference does the same for .obj seams, object.py:163-190) and drop
    degenerate triangles."""
    key = np.round(verts / max(tol, 1e-300)).astype(np.int64) if tol > 0 else verts
    _, first, inv = np.unique(key, axis=0, return_index=True, return_inverse=True)
    order = np.argsort(first)                      # keep first-occurrence order
    rank = np.empty_like(order)
    rank[order] = np.arange(order.size)
    new_faces = rank[inv.reshape(-1)][faces]
    ok = (new_faces[:, 0] != new_faces[:, 1]) & (new_faces[:, 1] != new_faces[:, 2]) & (new_faces[:, 0] != new_faces[:, 2])
    return verts[first[order]], new_faces[ok].astype(np.int32)


def winding_number(points, verts, faces, chunk=2048):
    """Generalised winding number of `points` w.r.t. the triangle soup (1 inside a closed, outward-oriented
    surface, 0 outside): sum of signed solid angles / 4 pi (Van Oosterom & Strackee)."""
    a0, b0, c0 = verts[faces[:, 0]], verts[faces[:, 1]], verts[faces[:, 2]]
    out = np.empty(len(points))
    for s in range(0, len(points), chunk):
        p = points[s:s + chunk, None, :]
        a, b, c = a0[None] - p, b0[None] - p, c0[None] - p
        la, lb, lc = np.linalg.norm(a, axis=2), np.linalg.norm(b, axis=2), np.linalg.norm(c, axis=2)
        num = np.einsum("ijk,ijk->ij", a, np.cross(b, c))
        den = la * lb * lc + np.einsum("ijk,ijk->ij", a, b) * lc + np.einsum("ijk,ijk->ij", b, c) * la + \
            np.einsum("ijk,ijk->ij", c, a) * lb
        out[s:s + chunk] = np.arctan2(num, den).sum(axis=1) / (2.0 * np.pi)
    return out


def tetrahedralize(verts, faces, spacing=None, min_quality=1e-3):
    """Tet mesh of the volume enclosed by (verts, faces).  Returns (points [n,3] float32 with the surface
    vertices FIRST and unchanged, tets [m,4] int32).  Every tet has non-zero volume; orientation is not
    normalised (the reference keeps negative-determinant rest tets too: object.py:389-391 uses |det|)."""
    from scipy.spatial import Delaunay, cKDTree

    verts = np.asarray(verts, dtype=np.float64)
    faces = np.asarray(faces, dtype=np.int32)
    if winding_number(_interior_probe(verts, faces), verts, faces)[0] < 0:
        faces = faces[:, [0, 2, 1]]                # inward-oriented input
    if spacing is None:
        e = np.concatenate([verts[faces[:, 0]] - verts[faces[:, 1]], verts[faces[:, 1]] - verts[faces[:, 2]],
                            verts[faces[:, 2]] - verts[faces[:, 0]]])
        spacing = float(np.linalg.norm(e, axis=1).mean())
    lo, hi = verts.min(axis=0), verts.max(axis=0)
    axes = [np.arange(lo[k] + 0.5 * spacing, hi[k], spacing) for k in range(3)]
    grid = np.stack(np.meshgrid(*axes, indexing="ij"), axis=-1).reshape(-1, 3)
    if len(grid):
        # body-centred offset on alternate layers gives better-shaped tets than a plain cubic lattice
        layer = np.rint((grid[:, 2] - axes[2][0]) / spacing).astype(np.int64) if len(axes[2]) else 0
        grid = grid + (layer % 2)[:, None] * np.array([0.5, 0.5, 0.0]) * spacing
        inside = winding_number(grid, verts, faces) > 0.5
        grid = grid[inside]
    if len(grid):
        far = cKDTree(verts).query(grid)[0] > 0.6 * spacing
        grid = grid[far]
    points = np.concatenate([verts, grid]) if len(grid) else verts.copy()
    tets = Delaunay(points).simplices.astype(np.int32)
    cent = points[tets].mean(axis=1)
    keep = winding_number(cent, verts, faces) > 0.5
    p = points[tets]
    vol = np.abs(np.einsum("ij,ij->i", p[:, 1] - p[:, 0], np.cross(p[:, 2] - p[:, 0], p[:, 3] - p[:, 0]))) / 6.0
    edge = np.linalg.norm(p[:, [1, 2, 3, 2, 3, 3]] - p[:, [0, 0, 0, 1, 1, 2]], axis=2).max(axis=1)
    keep &= vol > min_quality * edge ** 3          # drop slivers (TetGen: minratio / mindihedral, object.py:141)
    tets = tets[keep]
    if len(tets) == 0:
        raise ValueError("tetrahedralisation produced no elements (is the surface closed?)")
    # compact unused points but keep the order (surface vertices stay first when all are used)
    used = np.zeros(len(points), dtype=bool)
    used[tets.reshape(-1)] = True
    remap = np.cumsum(used) - 1
    return points[used].astype(np.float32), remap[tets].astype(np.int32)


def _interior_probe(verts, faces):
    """A point slightly behind the first non-degenerate triangle (inside for an outward-oriented surface)."""
    for f in faces:
        a, b, c = verts[f[0]], verts[f[1]], verts[f[2]]
        n = np.cross(b - a, c - a)
        ln = np.linalg.norm(n)
        if ln > 0:
            return ((a + b + c) / 3.0 - 1e-3 * np.sqrt(ln) * n / ln)[None, :]
    raise ValueError("surface has no non-degenerate triangle")


def tetrahedralize_file(path, spacing=None):
    """object.py:133-146: surface file -> (nodes, tets, (surface vertices, surface triangles))."""
    verts, faces = weld(*read_surface(path))
    try:                                           # the reference's own mesher when it happens to be installed
        import pyvista as pv
        import tetgen

        tet = tetgen.TetGen(pv.read(path))
        tet.tetrahedralize(order=1, mindihedral=10, minratio=5.0)
        return np.asarray(tet.node, dtype=np.float32), np.asarray(tet.elem, dtype=np.int32), (verts, faces)
    except ImportError:
        pass
    nodes, tets = tetrahedralize(verts, faces, spacing)
    return nodes, tets, (verts, faces)
