#!/usr/bin/env python
"""CPU model of the shared-memory wavefronts of the tile kernels, from the tiling tables alone (no GPU).

The tile kernels are shared-memory / latency bound before they are HBM bound (35.3 M LSU
wavefronts per C4 launch, 11.0 M of them bank conflicts -- profiles/r02_ncu_full_explicit.txt).  Which conflicts a
launch has is decided entirely on the host, by csrc/tiling.cpp: the slot of every element inside its tile (bank of
its staged results = slot mod 32, lane of its thread = slot mod 32) and the order of every vertex's incidence list.
This tool counts, for a tiling,

  gather   the JDS gather: per round and warp of 32 local vertices, D loads of stage[plane * S + slot]; wavefronts
           of one load = the largest number of active lanes that fall into one bank (slot mod 32)
  corners  the element phase's vertex reads: 4 (3 in 2D) LDS.128 of sv[local index] per element, 32 consecutive
           slots per warp; a 128-bit load is served a quarter warp (8 lanes) at a time, lanes that read the same
           float4 share a wavefront, distinct float4 with equal (index mod 8) conflict
  other    accesses that are conflict-free by construction (plane-major staging stores, list reads, vertex staging)

and prints totals scaled to a mesh size, so the model can be set against the ncu counters of a real launch and a
change of the host-side ordering can be judged before it is given GPU time.

  python tools/bank_model.py --cells 48 48 48 [--scale-to-elems 15980544] [--env FEMB_SLOT_ORDER=id] [--env FEMB_LIST_ORDER=greedy]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "finite-element-method_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def kuhn_box(nx, ny, nz, h):
    from femb200 import meshgen
    return meshgen.kuhn_box(nx, ny, nz, h)


def max_bank_multiplicity(banks, valid, nbank):
    """banks, valid: [n_groups, lanes] -> per group the largest count of valid lanes in one bank."""
    g = banks.shape[0]
    cnt = np.zeros((g, nbank), np.int32)
    rows = np.repeat(np.arange(g), banks.shape[1])
    np.add.at(cnt, (rows[valid.ravel()], banks.ravel()[valid.ravel()]), 1)
    return cnt.max(axis=1)


def model(t, dim):
    """t: femb200._capi.HostTiling.  Returns a dict of wavefront counts for ONE pass of a G_STAR tile kernel."""
    tiles = t.array(2).reshape(-1, 12)
    jds = t.array(4)
    rounds = t.array(5)
    lidx = t.array(6).reshape(-1, 2)
    S = t.info["stage_stride"]
    nc = dim + 1
    out = dict(gather_loads=0, gather_wavefronts=0, corner_loads=0, corner_wavefronts=0, corner_ideal=0, other=0,
               incidences=0, elements=0, tiles=len(tiles))
    for (slot_off, n_elem, lv_off, n_lv, jds_off, r_off, n_rounds, n_i, int_base, *_pad) in tiles:
        # ---- gather
        off = jds_off
        nw = (n_lv + 31) // 32
        for k in range(n_rounds):
            cnt = int(rounds[r_off + k])
            ent = jds[off:off + cnt].astype(np.int64)
            off += cnt
            bank = np.zeros(nw * 32, np.int64)
            bank[:cnt] = (ent % S) % 32
            valid = np.zeros(nw * 32, bool)
            valid[:cnt] = True
            active = valid.reshape(nw, 32).any(axis=1)
            wf = max_bank_multiplicity(bank.reshape(nw, 32), valid.reshape(nw, 32), 32)
            out["gather_loads"] += dim * int(active.sum())
            out["gather_wavefronts"] += dim * int(wf.sum())
            out["other"] += int(active.sum())                 # the u16 list read: consecutive lanes, one wavefront
        out["incidences"] += n_elem * nc
        out["elements"] += n_elem
        # ---- element phase: corner reads, 32 consecutive slots per warp, 8 lanes per 128-bit wavefront
        ne32 = (n_elem + 31) // 32 * 32
        w = lidx[slot_off:slot_off + ne32]
        loc = np.stack([w[:, 0] & 0xffff, w[:, 0] >> 16, w[:, 1] & 0xffff, w[:, 1] >> 16], axis=1)[:, :nc].astype(np.int64)
        valid = (np.arange(ne32) < n_elem)
        for c in range(nc):
            idx = loc[:, c].reshape(-1, 8)
            v = valid.reshape(-1, 8)
            # distinct addresses per bank group: sort each quarter warp, count distinct idx per (idx mod 8)
            wf = np.zeros(idx.shape[0], np.int64)
            for b in range(8):
                m = v & (idx % 8 == b)
                key = np.where(m, idx, -1)
                key.sort(axis=1)
                distinct = ((key[:, 1:] != key[:, :-1]) & (key[:, 1:] >= 0)).sum(axis=1) + (key[:, 0] >= 0)
                wf = np.maximum(wf, distinct)
            act = v.any(axis=1)
            out["corner_loads"] += int(valid.reshape(-1, 32).any(axis=1).sum())
            out["corner_wavefronts"] += int(wf[act].sum())
            out["corner_ideal"] += int(act.sum())
        # ---- conflict-free by construction: staging stores (D*D + D planes... NSTAGE words per element, one wavefront
        # per warp and word), vertex staging (STS.128: 4 wavefronts per warp) + local-vertex table reads (LDS.64: 2)
        out["other"] += (ne32 // 32) * (dim * dim) + nw * 6
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cells", type=int, nargs=3, default=[48, 48, 48])
    ap.add_argument("--h", type=float, default=0.0055)
    ap.add_argument("--tile-elems", type=int, default=768)
    ap.add_argument("--scale-to-elems", type=int, default=15980544, help="report totals scaled to this many elements (C4)")
    ap.add_argument("--env", action="append", default=[], help="KEY=VALUE set before the tiling is built")
    a = ap.parse_args()
    for kv in a.env:
        k, v = kv.split("=", 1)
        os.environ[k] = v
    from femb200 import _capi
    verts, el = kuhn_box(*a.cells, a.h)
    t = _capi.HostTiling(verts.astype(np.float32), el, a.tile_elems)
    r = model(t, 3)
    s = a.scale_to_elems / r["elements"]
    print("tiles %d, elements %d, scale to %d elements: x%.2f" % (r["tiles"], r["elements"], a.scale_to_elems, s))
    print("gather : %.3f wavefronts per load  (%.2f M wavefronts, %.2f M of them conflicts)" % (
        r["gather_wavefronts"] / r["gather_loads"], s * r["gather_wavefronts"] / 1e6,
        s * (r["gather_wavefronts"] - r["gather_loads"]) / 1e6))
    print("corners: %.3f wavefronts per quarter-warp load  (%.2f M wavefronts, %.2f M conflicts)" % (
        r["corner_wavefronts"] / r["corner_ideal"], s * r["corner_wavefronts"] / 1e6,
        s * (r["corner_wavefronts"] - r["corner_ideal"]) / 1e6))
    tot = r["gather_wavefronts"] + r["corner_wavefronts"] + r["other"]
    conf = (r["gather_wavefronts"] - r["gather_loads"]) + (r["corner_wavefronts"] - r["corner_ideal"])
    print("total  : %.2f M wavefronts, %.2f M conflicts   (ncu, C4 explicit tile kernel: 35.31 M / 10.99 M)" % (
        s * tot / 1e6, s * conf / 1e6))
    return r


if __name__ == "__main__":
    main()
