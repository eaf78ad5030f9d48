#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric: tet-element updates/s (explicit) and solver iterations/s (CG, Jacobi).

  python bench.py [--gpus N --steps K --warmup W] [--workload all|c4_explicit|c3_explicit|c4_cg|c4_jacobi]
  python bench.py --impl reference ...      # the reference algorithm on the host cores (oracle C/OpenMP port)

One "step" = one pass of the hot path over the whole mesh: for the explicit workloads one symplectic-Euler step
(element gradient + vertex update, main.py:103-110); for the implicit workloads one solver iteration.  The JSON
line's headline (`metric`, `value`) is the explicit step on C4 (the 16 M-tet box the metric is quoted on); with the
default `--workload all` the same line carries a `solvers` block (AtA-CG at dt 5e-4 and Jacobi at dt 1e-4 on the
same mesh: it/s, roofline, cpu_baseline, e2e) and `extras.c3_explicit` (BASELINE config 3, 1 M tets).
Everything goes through the drop-in modules (object.Object, solver.*): `value` is device-timed with the state
resident in HBM; `e2e` repeats the measurement with HOST buffers crossing PCIe every step.  At N > 1 the mesh is
cut into x-slabs by object.Object itself (strong scaling) and `config.parity` records a check of the N-GPU result
against a 1-GPU replay on rank 0.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "finite-element-method_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

MAT = dict(E=4e4, nu=0.4, rho=1000.0, damping=10.0)  # config/demo_3d.json:16-21
G_DIR = [0.0, -1.0, 0.0]

WORKLOADS = {
    # name: cells, h, offset, dt, kind                          
    "c4_explicit": dict(cells=(144, 136, 136), h=0.0055, offset=(0.1, 0.1, 0.1), dt=1e-4, kind="explicit"),
    "c3_explicit": dict(cells=(55, 55, 55), h=0.4 / 55, offset=(0.3, 0.5, 0.3), dt=1e-4, kind="explicit"),
    "c4_cg": dict(cells=(144, 136, 136), h=0.0055, offset=(0.1, 0.1, 0.1), dt=5e-4, kind="cg_ata"),
    "c4_jacobi": dict(cells=(144, 136, 136), h=0.0055, offset=(0.1, 0.1, 0.1), dt=1e-4, kind="jacobi"),
}
RESET_EVERY = 1000  # explicit: restore the initial state every 1000 steps so the body never reaches the floor
UNITS = {"explicit": ("tet_element_updates_per_sec", "tet-updates/s"), "cg_ata": ("solver_iterations_per_sec", "iterations/s"),
         "cg": ("solver_iterations_per_sec", "iterations/s"), "jacobi": ("solver_iterations_per_sec", "iterations/s")}


def lame(E, nu):
    return E / 2 / (1 + nu), E * nu / (1 + nu) / (1 - 2 * nu)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons WHILE the GPU is under the bench's load (B200_PROFILING.md recipe).  The
    headline timed region alone lasts milliseconds, so the sampler spans it AND the roofline / solver / e2e legs
    that follow back to back (seconds of the same kernels)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index, self.t0 = [], None, index, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            self.t0 = time.perf_counter()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        time.sleep(0.05)
        window = time.perf_counter() - self.t0
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "window_s": round(window, 2),
                "window": "headline timed region + roofline, solver and e2e legs, back to back"}


def algorithmic_bytes(kind, ne, nv):
    """BASELINE.md section 3 / SURVEY.md 8d byte model (fp32, 3D)."""
    if kind == "explicit":
        return 56 * ne + 88 * nv
    if kind == "cg_ata":
        return 104 * ne + 192 * nv
    if kind == "cg":
        return 52 * ne + 160 * nv
    return 52 * ne + 88 * nv  # jacobi


def top_config(w, kind):
    """The top-level JSON config (README.md:73-110 schema) of a workload."""
    return {"dim": 3, "auto_diff": False, "use_explicit_method": kind == "explicit", "delta_time": w["dt"],
            "g_dir": G_DIR, "implicit_method": 0 if kind == "jacobi" else 1, "preconditioned": 1 if kind == "cg_ata" else 0,
            "blocks": [{"id": 0, "block_center": [0.5, 0.5, 0.5], "block_radius": 0.0}]}   # demo_3d.json:24-28


def object_config(w):
    return {"id": 0, "rho": MAT["rho"], "E": MAT["E"], "nu": MAT["nu"], "damping": MAT["damping"],
            "center": list(w["offset"]), "cells": list(w["cells"]), "h": w["h"], "jitter": 0.1, "seed": 0,
            "surface": False, "verbose": False}


class Problem:
    """One mesh on the GPU(s) behind object.Object, plus what the timed legs need."""

    def __init__(self, w, implicit, args, torch, partition=True):
        import utils
        from femb200 import _capi as c
        from femb200 import dist as fdist
        utils.sys_init(top_config(w, "cg_ata" if implicit else "explicit"))
        from object import Object
        self.c, self.torch, self.w = c, torch, w
        ocfg = object_config(w)
        if args.comm:
            ocfg["comm"] = args.comm
        self.obj = Object(ocfg, tile_elems=args.tile, partition=partition)
        self.h = self.obj._h
        if self.obj._halo is None:
            self.stream = fdist.make_stream(self.h)
        else:
            self.stream = self.obj._stream
        self.pos0 = self.h.get(c.F_POS)
        self.vel0 = np.zeros_like(self.pos0)
        self.info = self.h.info()
        nx, ny, nz = w["cells"]
        self.ne_global, self.nv_global = nx * ny * nz * 6, (nx + 1) * (ny + 1) * (nz + 1)
        self.ne_local, self.nv_local = self.obj.element_cnt, self.obj.particle_cnt

    def reset(self):
        self.h.set(self.c.F_POS, self.pos0)
        self.h.set(self.c.F_VEL, self.vel0)

    def run(self, kind, dt, n):
        if kind == "explicit":
            done = 0
            while done < n:
                m = min(RESET_EVERY, n - done)
                self.obj._explicit_step(dt, G_DIR, m)
                done += m
                if done < n:
                    self.reset()
        else:
            # one solve of exactly n iterations (tol = 0 => fixed count, no host polling)
            method, pre = (0 if kind == "jacobi" else 1), (1 if kind == "cg_ata" else 0)
            self.obj._implicit_solve(dt, method, pre, max_iter=n, tol=0.0, fetch=False)

    def close(self):
        if self.obj._halo is not None:
            self.obj._halo.close()
        self.h.close()


def timed(prob, kind, dt, steps, warmup, barrier, world, dist):
    """W warm-up steps, then exactly `steps` steps between barriers, CUDA events on the launch stream, max over ranks."""
    torch, c = prob.torch, prob.c
    prob.reset()
    prob.run(kind, dt, warmup)
    # prime the CUDA graph of the timed chunk size (capture + instantiate must not be timed), then put the state back
    prob.run(kind, dt, min(RESET_EVERY, steps) if kind == "explicit" else steps)
    prob.reset()
    barrier()
    halo = prob.obj._halo
    replayed0 = getattr(halo, "replayed_launches", 0) if halo is not None else 0
    launches0 = c.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(prob.stream)
    prob.run(kind, dt, steps)
    e1.record(prob.stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = c.launch_count() - launches0
    if halo is not None:   # kernels inside replayed torch CUDA graphs are not seen by the library's counter
        launches += getattr(halo, "replayed_launches", 0) - replayed0
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    finite = bool(np.isfinite(prob.h.get(c.F_POS)).all())
    iters = None
    if kind != "explicit":
        finite = finite and bool(np.isfinite(prob.h.get(c.F_VEC_X)).all())
        iters = int(prob.h.last_solve_iters())
        if iters != steps:   # the solve must really have run `steps` iterations (no early exit on convergence / NaN)
            raise SystemExit("solver ran %d iterations, expected %d: timing invalid" % (iters, steps))
    return ms, int(launches), finite, iters


def roofline(prob, kind, dt, ms_per_step, workload, world):
    """Dominant kernel = the tile kernel of the pass: algorithmic bytes per launch / live CUDA-event time."""
    peak, peak_src = peaks()
    h, info = prob.h, prob.info
    ne_l, nv_l = prob.ne_local, prob.nv_local
    bts = algorithmic_bytes(kind, ne_l, nv_l)
    if kind == "explicit":
        prob.reset()
        h.explicit_step_timed(dt, G_DIR, 5)
        t_ms, b_ms = h.explicit_step_timed(dt, G_DIR, 50)
        itc = os.environ.get("FEMB_ITC", "0") == "1"
        # in-tile completion: the tile kernel also finishes the boundary vertices, i.e. it IS the whole step
        # (a halo boundary kernel remains on a slab); otherwise it reads 12 B and leaves a partial per boundary vertex
        tile_bytes = bts if itc else 56 * ne_l + 88 * info["n_interior"] + 12 * info["n_boundary"]
        ach = tile_bytes / (t_ms / 50 * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "k_tile<3,OpExplicit<3,EX_STEP>>" + (" (boundary vertices finished in-tile)" if itc else ""),
                "achieved": round(ach, 1), "peak": peak,
                "unit": "GB/s", "frac": round(ach / peak, 4), "traffic": None, "peak_source": peak_src,
                "bytes_per_launch": tile_bytes, "tile_ms": round(t_ms / 50, 5), "boundary_ms": round(b_ms / 50, 5),
                "step_bytes": bts, "step_frac": round(bts / (ms_per_step * 1e-3) / 1e9 / peak, 4),
                "note": "tile_ms + boundary_ms are timed launch by launch outside the CUDA graph and can exceed ms_per_step by ~2 %"}
    else:
        it_ach = bts / (ms_per_step * 1e-3) / 1e9
        roof = {"bound": "hbm", "peak": peak, "unit": "GB/s", "traffic": None, "peak_source": peak_src,
                "iteration_bytes": bts, "iteration_achieved": round(it_ach, 1), "iteration_frac": round(it_ach / peak, 4)}
        if prob.obj._halo is None:
            h.implicit_apply_timed(dt, 0, 5)
            t_ms, b_ms = h.implicit_apply_timed(dt, 0, 50)
            tile_bytes = 52 * ne_l + 28 * info["n_interior"] + 12 * info["n_boundary"]
            ach = tile_bytes / (t_ms / 50 * 1e-3) / 1e9
            roof.update({"kernel": "operator pass A d: tile kernel of OpApply<3,0>", "achieved": round(ach, 1),
                         "frac": round(ach / peak, 4), "bytes_per_launch": tile_bytes, "tile_ms": round(t_ms / 50, 5),
                         "boundary_ms": round(b_ms / 50, 5)})
        else:
            roof.update({"kernel": "solver iteration (all kernels)", "achieved": round(it_ach, 1),
                         "frac": round(it_ach / peak, 4), "bytes_per_launch": bts})
    # measured DRAM traffic of the dominant kernel (one ncu --set full capture, committed under profiles/)
    for name in ("r02_traffic.json", "r01_traffic.json"):
        tpath = os.path.join(ROOT, "profiles", name)
        if os.path.exists(tpath) and world == 1:
            tr = json.load(open(tpath)).get(workload)
            if tr:
                roof["traffic"] = tr["traffic_bytes"]
                roof["traffic_source"] = tr["source"]
                break
    return roof


def e2e_explicit(prob, w, steps, barrier, world, dist):
    """The same step through the public, reference-shaped API with HOST buffers: every step copies pos and vel in
    from pinned host memory, runs neo_hookean_1_grad + kinematic and copies pos out.  The copies are enqueued with
    blocking=False (extension of from_numpy/to_numpy), so H2D of step k+1, the kernels of step k and D2H of step
    k-1 overlap; the blocking (reference-literal) sequence is timed next to it."""
    import utils
    torch, obj = prob.torch, prob.obj
    utils.sys_init(top_config(w, "explicit"))
    from circle_blocks import CircleBlocks
    from solver.explicit import neo_hookean_1_grad
    from solver.kinematic import kinematic
    blocks = CircleBlocks(top_config(w, "explicit")["blocks"])
    P = obj.particles
    pin = lambda a: torch.from_numpy(a.copy()).pin_memory().numpy()  # noqa: E731
    np_pos, np_vel = pin(prob.pos0), pin(prob.vel0)
    outs = [pin(prob.pos0), pin(prob.pos0)]

    def loop(n, blocking):
        for k in range(n):
            P.pos.from_numpy(np_pos, blocking=blocking)          # H2D
            P.vel.from_numpy(np_vel, blocking=blocking)          # H2D
            neo_hookean_1_grad(obj)
            kinematic(obj, blocks)                               # one fused pass (+ halo exchange on a slab)
            P.pos.to_numpy(out=outs[k & 1], blocking=blocking)   # D2H
        obj.sync()

    res = {}
    for blocking in (True, False):
        loop(3, blocking)
        barrier()
        t0 = time.perf_counter()
        loop(steps, blocking)
        barrier()
        dt_s = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt_s], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt_s = float(t.item())
        res[blocking] = dt_s
    ok = bool(np.isfinite(outs[0]).all() and np.isfinite(outs[1]).all() and not np.array_equal(outs[0], prob.pos0))
    return {"value": prob.ne_global * steps / res[False], "unit": "tet-updates/s",
            "h2d_bytes_per_step": int(np_pos.nbytes + np_vel.nbytes), "d2h_bytes_per_step": int(outs[0].nbytes),
            "steps": steps, "ms_per_step": res[False] / steps * 1e3, "blocking_ms_per_step": res[True] / steps * 1e3,
            "result_checked": ok,
            "api": "object.Object: particles.pos/vel.from_numpy(blocking=False) -> neo_hookean_1_grad -> kinematic -> "
                   "particles.pos.to_numpy(out=pinned, blocking=False); obj.sync() at the end"}


def e2e_solver(prob, w, kind, solves, barrier, world, dist):
    """implicit_solver_neo_hookean(obj) to the reference's own stopping rule, host pos / vel in, vel out, per solve."""
    import utils
    torch, obj = prob.torch, prob.obj
    utils.sys_init(top_config(w, kind))
    from solver.implicit import implicit_solver_neo_hookean
    P = obj.particles
    pin = lambda a: torch.from_numpy(a.copy()).pin_memory().numpy()  # noqa: E731
    np_pos, np_vel, np_out = pin(prob.pos0), pin(prob.vel0), pin(prob.vel0)

    def one():
        P.pos.from_numpy(np_pos)                 # H2D
        P.vel.from_numpy(np_vel)                 # H2D
        implicit_solver_neo_hookean(obj)         # assemble + solve, vel = x (implicit.py:201-223)
        P.vel.to_numpy(out=np_out)               # D2H
        return obj.last_solve["iters"], obj.last_solve["resid"]

    one()
    barrier()
    t0 = time.perf_counter()
    its = 0
    for _ in range(solves):
        it, resid = one()
        its += it
    barrier()
    dt_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt_s = float(t.item())
    return {"value": its / dt_s, "unit": "iterations/s", "h2d_bytes_per_step": int((np_pos.nbytes + np_vel.nbytes) * solves / max(its, 1)),
            "d2h_bytes_per_step": int(np_out.nbytes * solves / max(its, 1)), "solves": solves, "iterations_per_solve": its / solves,
            "final_residual": float(resid), "ms_per_solve": dt_s / solves * 1e3,
            "h2d_bytes_per_solve": int(np_pos.nbytes + np_vel.nbytes), "d2h_bytes_per_solve": int(np_out.nbytes),
            "api": "object.Object: particles.pos/vel.from_numpy -> implicit_solver_neo_hookean (reference stop rule) -> "
                   "particles.vel.to_numpy; a step of the metric = one solver iteration"}


def parity_multi(prob, args, torch, dist, rank, world):
    """N > 1: the slab run against a 1-GPU replay of the same calls on rank 0 (positions after 100 explicit steps,
    x after 20 fixed iterations of each solver) + bit-identity of every shared plane on its two owners."""
    from femb200 import dist as fdist
    obj, c, w = prob.obj, prob.c, prob.w
    _, _, i0, i1, (nx, ny, nz) = obj._slab
    plane = (ny + 1) * (nz + 1)
    res = {"explicit_steps": 100, "solver_iterations": 20}
    jobs = [("explicit", WORKLOADS["c4_explicit"]["dt"], 100, c.F_POS)]
    if obj.implicit:
        jobs += [("cg_ata", WORKLOADS["c4_cg"]["dt"], 20, c.F_VEC_X), ("jacobi", WORKLOADS["c4_jacobi"]["dt"], 20, c.F_VEC_X)]
    got = {}
    planes_ok = True
    for kind, dt, n, field in jobs:
        prob.reset()
        prob.run(kind, dt, n)
        local = prob.h.get(field)
        parts = fdist.all_gather_arrays(local)
        planes_ok = planes_ok and all(np.array_equal(parts[r][-plane:], parts[r + 1][:plane]) for r in range(world - 1))
        got[kind] = np.concatenate([p if r == world - 1 else p[:-plane] for r, p in enumerate(parts)], axis=0)
    res["shared_planes_bit_identical"] = bool(planes_ok)
    if rank == 0:
        single = Problem(w, obj.implicit, args, torch, partition=False)
        worst = 0.0
        for kind, dt, n, field in jobs:
            single.reset()
            single.run(kind, dt, n)
            want = single.h.get(field).astype(np.float64)
            err = float(np.linalg.norm(got[kind] - want) / max(np.linalg.norm(want), 1e-300))
            res["rel_l2_" + kind] = err
            tol = 1e-6 if kind == "explicit" else 1e-4
            worst = max(worst, err / tol)
        single.h.close()
        res["tolerance"] = {"explicit_pos": 1e-6, "solver_x": 1e-4}
        res["status"] = "pass" if (worst <= 1.0 and planes_ok) else "FAIL"
        res["against"] = "1-GPU replay of the same public calls on rank 0 (itself pinned to the reference fixtures by tests/)"
    if world > 1:
        dist.barrier()
    return res


# ---------------------------------------------------------------------- CPU side (oracle C/OpenMP port)
class CpuProblem:
    """The workload's mesh and state for the C/OpenMP restatement (oracle/fem_ref.c) on all host cores."""

    def __init__(self, w, cells=None):
        from oracle import cref
        from oracle import fem_oracle as o
        self.cref, self.o = cref, o
        self.threads = cref.set_threads(os.cpu_count())
        self.cells = tuple(cells or w["cells"])
        self.w = w
        verts, self.tets = o.kuhn_box(*self.cells, w["h"], dtype=np.float32)
        self.pos0 = o.particles_init(verts, w["offset"])
        del verts
        self.ref, self.vol, self.mass = cref.elements_init(self.pos0, self.tets, MAT["rho"])
        rng = np.random.default_rng(0)
        self.pos = (self.pos0 + np.float32(0.1 * w["h"]) * rng.uniform(-1, 1, self.pos0.shape).astype(np.float32)).astype(np.float32)
        self.mu, self.lam = o.lame(MAT["E"], MAT["nu"])
        self.ne = len(self.tets)
        self.full = self.cells == tuple(w["cells"])
        self.H = None

    def describe(self):
        return "%s %dx%dx%d-cell Kuhn box (%d tets)" % ("the full" if self.full else "a", *self.cells, self.ne)

    def explicit(self, dt, steps):
        pos, vel, force = self.pos.copy(), np.zeros_like(self.pos), np.zeros_like(self.pos)
        t0 = time.perf_counter()
        self.cref.explicit_steps(pos, vel, force, self.mass, self.tets, self.ref, self.vol, self.mu, self.lam, dt, G_DIR,
                                 MAT["damping"], [], steps)
        return time.perf_counter() - t0

    def solver(self, kind, dt, iters):
        if self.H is None:
            self.H = self.cref.hessian_blocks(self.pos, self.tets, self.ref, self.vol, self.mu, self.lam)
            self.b = {}
        if dt not in self.b:
            f = np.zeros_like(self.pos)     # b = vel + dt/m * f with vel = 0; f = -dE/dx (log-J convention differs by rounding only here)
            self.cref.explicit_grad(self.pos, self.tets, self.ref, self.vol, self.mu, self.lam, f)
            self.b[dt] = (np.float32(-dt) * f / self.mass[:, None]).astype(np.float32)
        b = self.b[dt]
        t0 = time.perf_counter()
        if kind == "jacobi":
            _, it, _ = self.cref.jacobi(self.H, self.tets, self.mass, dt, b, np.zeros_like(b), iter_max=iters, tol=0.0)
        else:
            _, it, _ = self.cref.cg(self.H, self.tets, self.mass, dt, b, 1 if kind == "cg_ata" else 0, iter_max=iters, tol=0.0)
        assert it == iters, (it, iters)
        return time.perf_counter() - t0

    def run(self, kind, dt, n):
        return self.explicit(dt, n) if kind == "explicit" else self.solver(kind, dt, n)


def cpu_baseline(cpu, kind, dt, budget_s=10.0):
    """~budget_s seconds of the reference algorithm on the host cores, after a 2-step calibration."""
    t_cal = cpu.run(kind, dt, 2)
    n = int(max(3, min(2000, budget_s / max(t_cal / 2, 1e-6))))
    t = cpu.run(kind, dt, n)
    units = cpu.ne if kind == "explicit" else 1
    value = units * n / t
    if kind != "explicit" and not cpu.full:      # iterations/s of a sub-box, scaled to the full mesh
        value *= cpu.ne / (cpu.w["cells"][0] * cpu.w["cells"][1] * cpu.w["cells"][2] * 6)
    return {"value": value, "unit": UNITS[kind][1], "cores": cpu.threads, "kind": "port",
            "sample": "%d %s of %s, %.1f s of CPU work; C/OpenMP restatement of the reference algorithm, %d threads "
                      "(the reference's own ti.cpu path needs taichi, which is not installable here)"
                      % (n, "steps" if kind == "explicit" else "iterations", cpu.describe(), t, cpu.threads)}


def run_ours(args):
    import torch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torch.distributed.run" % (args.gpus, world))
    torch.cuda.set_device(local_rank)
    from femb200 import _capi as c
    from femb200 import dist as fdist
    dist = None
    if world > 1:
        import torch.distributed as dist
        fdist.ensure_process_group("nccl")
    c.set_host_threads(max(1, (os.cpu_count() or 1) // world))   # torchrun exports OMP_NUM_THREADS=1

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    everything = args.workload == "all"
    head_name = "c4_explicit" if everything else args.workload
    w = dict(WORKLOADS[head_name])
    if args.x_cells:      # diagnostic only: a thinner box (e.g. the per-rank slab of a larger N); not a bench line
        w["cells"] = (args.x_cells,) + tuple(w["cells"][1:])
    kind = w["kind"]
    prob = Problem(w, everything or kind != "explicit", args, torch)

    sampler = ClockSampler(local_rank)
    sampler.start()
    ms, launches, finite, iters = timed(prob, kind, w["dt"], args.steps, args.warmup, barrier, world, dist)
    roof = roofline(prob, kind, w["dt"], ms / args.steps, head_name, world)
    solvers = {}
    if everything:
        for name in ("c4_cg", "c4_jacobi"):
            sw = dict(WORKLOADS[name], cells=w["cells"])
            s_ms, s_launch, s_fin, s_it = timed(prob, sw["kind"], sw["dt"], args.solver_iters, 10, barrier, world, dist)
            solvers[name] = {"metric": UNITS[sw["kind"]][0], "value": args.solver_iters / (s_ms * 1e-3), "unit": "iterations/s",
                             "ms_per_iter": s_ms / args.solver_iters, "iterations": args.solver_iters, "warmup": 10,
                             "dt": sw["dt"], "solver": "A^T A-CG (preconditioned: 1)" if sw["kind"] == "cg_ata" else "weighted Jacobi",
                             "gpu_launches": s_launch, "state_finite": s_fin,
                             "roofline": roofline(prob, sw["kind"], sw["dt"], s_ms / args.solver_iters, name, world),
                             "e2e": e2e_solver(prob, sw, sw["kind"], 2 if sw["kind"] == "cg_ata" else 5, barrier, world, dist)}
    if kind == "explicit":
        e2e = e2e_explicit(prob, w, max(3, min(args.e2e_steps, args.steps)), barrier, world, dist)
    else:
        e2e = e2e_solver(prob, w, kind, 2 if kind == "cg_ata" else 5, barrier, world, dist)
    clocks = sampler.stop()

    parity = None
    if world > 1 and not args.no_parity:
        parity = parity_multi(prob, args, torch, dist, rank, world)

    extras = {}
    if everything and world == 1:
        w3 = dict(WORKLOADS["c3_explicit"])
        p3 = Problem(w3, False, args, torch)
        ms3, l3, f3, _ = timed(p3, "explicit", w3["dt"], 2000, 50, barrier, world, dist)
        r3 = roofline(p3, "explicit", w3["dt"], ms3 / 2000, "c3_explicit", world)
        extras["c3_explicit"] = {"value": p3.ne_global * 2000 / (ms3 * 1e-3), "unit": "tet-updates/s", "ms_per_step": ms3 / 2000,
                                 "steps": 2000, "n_tets": p3.ne_global, "gpu_launches": l3, "state_finite": f3, "roofline": r3,
                                 "cache": "71 MB working set: L2-resident (126 MB L2); the fraction is against the HBM figure"}
        p3.close()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cp = CpuProblem(w, cells=args.cpu_cells)
        cpu = cpu_baseline(cp, kind, w["dt"])
        for name, s in solvers.items():
            sw = WORKLOADS[name]
            s["cpu_baseline"] = cpu_baseline(cp, sw["kind"], sw["dt"], budget_s=8.0)
    halo = prob.obj._halo
    if rank == 0:
        units = prob.ne_global if kind == "explicit" else 1
        out = {
            "metric": UNITS[kind][0], "value": units * args.steps / (ms * 1e-3), "unit": UNITS[kind][1],
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic Kuhn tet box, seeded 10% h vertex jitter, material from config/demo_3d.json",
            "config": {"workload": head_name, "cells": list(w["cells"]), "n_tets": prob.ne_global, "n_verts": prob.nv_global,
                       "dt": w["dt"], "partition": "x-slabs, %d rank(s), cut by object.Object" % world,
                       "comm": None if halo is None else ("peer memory (in-kernel NVLink stores + flags)" if halo.peer
                                                          else "NCCL P2P + all_reduce"), "tile_elems": prob.info["tile_elems"],
                       "n_tiles_rank0": prob.info["n_tiles"],
                       "tile_tables": {"list_order": os.environ.get("FEMB_LIST_ORDER", "matching"),
                                       "slot_order": os.environ.get("FEMB_SLOT_ORDER", "bank"),
                                       "vertex_order": os.environ.get("FEMB_VERTEX_ORDER", "bank")},
                       "cache": "inputs %.0f MB per rank > 126 MB L2" %
                       (algorithmic_bytes(kind, prob.ne_local, prob.nv_local) / 1e6),
                       "state_reset_every": RESET_EVERY if kind == "explicit" else None, "state_finite": finite,
                       "solver_iterations_run": iters, "parity": parity},
            "clocks": clocks, "gpu_launches": int(launches), "roofline": roof, "e2e": e2e, "cpu_baseline": cpu,
        }
        if solvers:
            out["solvers"] = solvers
        if extras:
            out["extras"] = extras
        print(json.dumps(out))
    prob.close()
    if world > 1:
        fdist.shutdown()


def run_reference(args):
    """--impl reference: the reference algorithm on the host cores (oracle C/OpenMP port -- the reference is Taichi-JIT
    Python and taichi is not installable, so kind = "port"), on the FULL mesh of the workload with every host core
    (the thread count is set here: torchrun exports OMP_NUM_THREADS=1)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = "c4_explicit" if args.workload == "all" else args.workload
    w = dict(WORKLOADS[name])
    if args.x_cells:
        w["cells"] = (args.x_cells,) + tuple(w["cells"][1:])
    kind = w["kind"]
    full = w["cells"][0] * w["cells"][1] * w["cells"][2] * 6
    warm, steps = max(1, args.warmup), max(1, args.steps)
    cells = tuple(args.cpu_cells or w["cells"])
    while True:     # the full mesh unless K + W steps of it would exceed ~150 s of CPU work (then a sub-box)
        cp = CpuProblem(w, cells=cells)
        t_cal = cp.run(kind, w["dt"], 2)
        if t_cal / 2 * (steps + warm) <= 150.0 or max(cells) <= 16:
            break
        cells = tuple(max(16, c * 3 // 4) for c in cells)
    cp.run(kind, w["dt"], warm)
    t = cp.run(kind, w["dt"], steps)
    units = cp.ne if kind == "explicit" else 1
    value = units * steps / t
    if kind != "explicit" and not cp.full:
        value *= cp.ne / full
    sample = "%d %s of %s on %d threads" % (steps, "steps" if kind == "explicit" else "iterations", cp.describe(), cp.threads)
    out = {
        "impl": "reference", "metric": UNITS[kind][0], "value": value, "unit": UNITS[kind][1],
        "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": t / steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic Kuhn tet box, seeded 10% h vertex jitter, material from config/demo_3d.json",
        "config": {"workload": name, "cells": list(w["cells"]), "n_tets": full, "n_verts": int(np.prod([c + 1 for c in w["cells"]])),
                   "dt": w["dt"], "sample": sample, "same_config": bool(cp.full)},
        "cpu_baseline": {"value": value, "unit": UNITS[kind][1], "cores": cp.threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNITS[kind][1], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if args.workload == "all":     # the other half of the metric, on the same mesh, bounded
        out["solvers"] = {}
        for sname in ("c4_cg", "c4_jacobi"):
            sw = WORKLOADS[sname]
            b = cpu_baseline(cp, sw["kind"], sw["dt"], budget_s=8.0)
            out["solvers"][sname] = {"metric": UNITS[sw["kind"]][0], "value": b["value"], "unit": "iterations/s",
                                     "e2e": {"value": b["value"], "unit": "iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                                     "cpu_baseline": b}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="all", choices=["all"] + sorted(WORKLOADS))
    ap.add_argument("--tile", type=int, default=0, help="elements per tile (0 = library default)")
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--solver-iters", type=int, default=200, help="timed iterations of each solver in the `solvers` block")
    ap.add_argument("--cpu-cells", type=int, nargs=3, default=None, help="CPU legs: a sub-box instead of the full mesh")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the check against the 1-GPU replay")
    ap.add_argument("--x-cells", type=int, default=0, help="diagnostic: override the box's cell count along x")
    ap.add_argument("--comm", default=None, choices=["peer", "nccl"], help="multi-GPU transport (default: peer memory)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
