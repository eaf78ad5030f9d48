"""Surface ingestion + fallback tetrahedraliser (reference object.py:128-161, 255-308)."""
import os
import struct

import numpy as np
import pytest

from femb200 import meshgen, meshio

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "finite-element-method_b200", "obj")


def _surface_volume(v, f):
    a, b, c = v[f[:, 0]], v[f[:, 1]], v[f[:, 2]]
    return abs(np.einsum("ij,ij->i", a, np.cross(b, c)).sum() / 6.0)


def _tet_volumes(n, t):
    p = n[t].astype(np.float64)
    return np.einsum("ij,ij->i", p[:, 1] - p[:, 0], np.cross(p[:, 2] - p[:, 0], p[:, 3] - p[:, 0])) / 6.0


@pytest.mark.parametrize("name", ["blob100.obj", "blob.obj"])
def test_tetrahedralize_repo_surfaces(name):
    nodes, tets, (sv, sf) = meshio.tetrahedralize_file(os.path.join(OBJ, name))
    assert nodes.dtype == np.float32 and tets.dtype == np.int32
    assert np.array_equal(nodes[:len(sv)], sv.astype(np.float32))        # surface points first, unchanged
    vol = _tet_volumes(nodes, tets)
    assert (np.abs(vol) > 0).all()
    assert abs(np.abs(vol).sum() - _surface_volume(sv, sf)) / _surface_volume(sv, sf) < 0.02
    # conforming: every face belongs to one or two tets; the boundary is closed (each edge in 2 faces)
    combos = np.array([[0, 1, 2], [0, 1, 3], [0, 2, 3], [1, 2, 3]])
    faces = np.sort(tets[:, combos].reshape(-1, 3), axis=1)
    _, counts = np.unique(faces, axis=0, return_counts=True)
    assert counts.max() <= 2
    bf = meshgen.boundary_faces(nodes, tets)
    assert len(bf) == (counts == 1).sum()
    edges = np.sort(np.concatenate([bf[:, [0, 1]], bf[:, [1, 2]], bf[:, [2, 0]]]), axis=1)
    _, ec = np.unique(edges, axis=0, return_counts=True)
    assert (ec % 2 == 0).all()
    # outward orientation of the extracted surface (object.py:286-297): positive enclosed volume
    a, b, c = nodes[bf[:, 0]].astype(np.float64), nodes[bf[:, 1]].astype(np.float64), nodes[bf[:, 2]].astype(np.float64)
    assert np.einsum("ij,ij->i", a, np.cross(b, c)).sum() > 0


def test_obj_parser_polygons_negative_indices_and_slashes(tmp_path):
    p = tmp_path / "quad.obj"
    p.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvt 0 0\nf 1/1 2/1 3/1 4/1\nf -4//1 -3//1 -2//1\n")
    v, f = meshio.read_obj(str(p))
    assert v.shape == (4, 3) and f.tolist() == [[0, 1, 2], [0, 2, 3], [0, 1, 2]]


def test_binary_stl_cube_roundtrip(tmp_path):
    v = np.array([[x, y, z] for x in (0, 1) for y in (0, 1) for z in (0, 1)], dtype=np.float32)
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    tris = [(q[0], q[1], q[2]) for q in quads] + [(q[0], q[2], q[3]) for q in quads]
    p = tmp_path / "cube.stl"
    with open(p, "wb") as fh:
        fh.write(b"\0" * 80 + struct.pack("<I", len(tris)))
        for t in tris:
            fh.write(struct.pack("<3f", 0, 0, 0) + v[list(t)].tobytes() + b"\0\0")
    nodes, tets, (sv, sf) = meshio.tetrahedralize_file(str(p))
    assert len(sv) == 8 and len(sf) == 12
    assert abs(np.abs(_tet_volumes(nodes, tets)).sum() - 1.0) < 1e-6


def test_weld_merges_duplicates_keeps_order():
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 0, 0], [0, 1, 0]], dtype=np.float64)
    f = np.array([[0, 1, 3], [2, 1, 3], [0, 2, 3]], dtype=np.int32)
    wv, wf = meshio.weld(v, f)
    assert wv.tolist() == [[0, 0, 0], [1, 0, 0], [0, 1, 0]]
    assert wf.tolist() == [[0, 1, 2], [0, 1, 2]]          # the degenerate third triangle is dropped


@pytest.mark.parametrize("name", ["k4_analytic", "spot_analytic", "spot100_cg"])
def test_boundary_faces_equal_the_reference_extract_surface(name):
    """meshgen.boundary_faces against the surface the REFERENCE's own Object.extract_surface (object.py:255-308)
    produced for the same tets (stored in the reference fixtures, tests/golden/make_ref_golden.py): same triangles,
    same winding."""
    import ref_fixtures as rf
    fx = rf.load(name)
    bf = meshgen.boundary_faces(fx["raw_verts"].astype(np.float64), fx["elems"])

    def canon(F):
        return set(tuple(np.roll(f, -int(np.argmin(f)))) for f in np.asarray(F).tolist())
    assert canon(bf) == canon(fx["faces"]) and len(bf) == len(fx["faces"])


@pytest.mark.parametrize("name,nv,nf", [("spot100.obj", 119, 234), ("spot.obj", 587, 1170), ("cube.stl", 8, 12)])
def test_tetrahedralize_reference_assets(name, nv, nf):
    """The assets BASELINE configs 2 and 5 name (reference obj/, vendored as input data) and the binary STL."""
    nodes, tets, (sv, sf) = meshio.tetrahedralize_file(os.path.join(OBJ, name))
    assert len(sv) == nv and len(sf) == nf
    assert np.array_equal(nodes[:len(sv)], sv.astype(np.float32))
    vol = np.abs(_tet_volumes(nodes, tets))
    assert (vol > 0).all() and abs(vol.sum() - _surface_volume(sv, sf)) / _surface_volume(sv, sf) < 0.03
