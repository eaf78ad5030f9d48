#!/usr/bin/env python
"""Kernel timeline of the tile passes on every rank (femb_trace_enable / femb_trace_get): the last explicit steps of a
run and the last solver iteration, times in microseconds relative to the first traced kernel of the rank.

  python tools/trace_steps.py [--x-cells N]                                 one GPU (optionally a slab-sized mesh)
  python -m torch.distributed.run --nproc-per-node 8 ... tools/trace_steps.py   one rank per GPU

Columns: start of CTA 0, last CTA end, last end of a wait for the neighbours' halo flags, release of own flags,
and the first sender CTA's own timeline (released / values stored / fence done)."""
import argparse
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "finite-element-method_b200"))
# the timeline code lives in the diagnostics build of the library (csrc/dev.cuh: FEMB_DIAG_BUILD)
os.environ.setdefault("FEMB200_LIB", os.path.join(ROOT, "finite-element-method_b200", "femb200", "libfemb200_diag.so"))
import bench  # noqa: E402


def table(rows):
    t0 = min(r[1] for r in rows if r[1])
    out = []
    for name, a, b, w, f, s0, s1, s2, _ in rows:
        us = lambda t: "%8.1f" % ((t - t0) / 1e3) if t else "       -"
        out.append("  %-32s start %s  end %s  dur %6.1f  wait_end %s  flags %s  sender: go %s stored %s fenced %s" % (
            name, us(a), us(b), (b - a) / 1e3 if a and b else 0.0, us(w), us(f), us(s0), us(s1), us(s2)))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--x-cells", type=int, default=0)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--out", default="gpurun_out/trace.txt")
    a = ap.parse_args()
    import torch
    from femb200 import dist as fdist
    rank, world = fdist.ensure_process_group()
    torch.cuda.set_device(fdist.local_device())
    args = types.SimpleNamespace(tile=0, comm=None)
    lines = []
    for wname, kind in (("c4_explicit", "explicit"), ("c4_cg", "cg_ata"), ("c4_jacobi", "jacobi")):
        w = dict(bench.WORKLOADS[wname])
        if a.x_cells:
            w["cells"] = (a.x_cells,) + tuple(w["cells"][1:])
        prob = bench.Problem(w, kind != "explicit", args, torch)
        h = prob.h
        h.trace_enable(True)
        if kind == "explicit":
            for _ in range(12):
                prob.run(kind, w["dt"], a.steps)       # one cached graph of a.steps steps, replayed
        else:
            prob.run(kind, w["dt"], 60)
        h.sync()
        rows = h.trace_get()
        h.trace_enable(False)
        lines.append("== %s  rank %d/%d  tiles %d (halo %d)  slots %d" % (wname, rank, world, prob.info.get("n_tiles", 0), prob.info.get("n_halo_tiles", 0), len(rows)))
        lines += table(rows)
        prob.close()
    text = "\n".join(lines)
    if world > 1:
        import torch.distributed as dist
        allt = [None] * world
        dist.all_gather_object(allt, text)
        text = "\n\n".join(allt)
    if rank == 0:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        open(a.out, "w").write(text + "\n")
    if world > 1:
        fdist.shutdown()


if __name__ == "__main__":
    main()
