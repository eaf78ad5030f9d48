#!/usr/bin/env python
"""bench.py -- tet-element updates/s of the explicit Neo-Hookean step (BASELINE.json metric).

  python bench.py [--gpus N --steps K --warmup W] [--workload c4_explicit|c3_explicit|c4_cg|c4_jacobi]
  python bench.py --impl reference ...      # the reference algorithm on the host cores (oracle C port)

One "step" = one pass of the hot path over the whole mesh: for the explicit
workloads one symplectic-Euler step (element gradient + vertex update,
main.py:103-110); for the implicit workloads one solver iteration.  Inputs
(mesh, Dm^-1, state) are resident in HBM when the timed region starts; `e2e`
repeats the measurement through the reference-shaped Python API with host
buffers (pinned) crossing PCIe every step.  Prints ONE JSON line on rank 0.
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


def lame(E, nu):
    return E / 2 / (1 + nu), E * nu / (1 + nu) / (1 - 2 * nu)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
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
                "samples": len(sm)}


def slab(cells, rank, world):
    """x-slab [i0, i1) of cell layers owned by `rank`."""
    nx = cells[0]
    base, rem = divmod(nx, world)
    i0 = rank * base + min(rank, rem)
    return i0, i0 + base + (1 if rank < rem else 0)


def make_state(w, rank, world, seed=0):
    """Local slab mesh + jittered initial positions (x = X + 0.1 h U(-1,1)).  The jitter is
    drawn for the GLOBAL vertex array so that every rank sees the same values on shared interface planes."""
    from femb200 import meshgen

    nx, ny, nz = w["cells"]
    i0, i1 = slab(w["cells"], rank, world)
    verts, tets = meshgen.kuhn_box(nx, ny, nz, w["h"], x_cells=(i0, i1))
    plane = (ny + 1) * (nz + 1)
    rng = np.random.default_rng(seed)
    jit = rng.uniform(-1.0, 1.0, ((nx + 1) * plane, 3)).astype(np.float32)[i0 * plane:(i1 + 1) * plane]
    center = np.asarray(w["offset"], dtype=np.float32)
    pos = (verts + center[None, :]).astype(np.float32) + np.float32(0.1 * w["h"]) * jit
    return verts, tets, center, pos.astype(np.float32), (i0, i1)


def algorithmic_bytes(kind, ne, nv):
    """BASELINE.md section 3 / SURVEY.md 8d byte model (fp32, 3D)."""
    if kind == "explicit":
        return 56 * ne + 88 * nv
    if kind == "cg_ata":
        return 104 * ne + 192 * nv
    if kind == "cg":
        return 52 * ne + 160 * nv
    return 52 * ne + 88 * nv  # jacobi


def run_ours(args):
    import torch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torch.distributed.run" % (args.gpus, world))
    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from femb200 import _capi as c

    w = dict(WORKLOADS[args.workload])
    if args.x_cells:      # diagnostic only: a thinner box (e.g. the per-rank slab of a larger N); not a bench line
        w["cells"] = (args.x_cells,) + tuple(w["cells"][1:])
    kind = w["kind"]
    mu, lam = lame(MAT["E"], MAT["nu"])
    verts, tets, center, pos0, (i0, i1) = make_state(w, rank, world)
    nx, ny, nz = w["cells"]
    ne_global = nx * ny * nz * 6
    nv_global = (nx + 1) * (ny + 1) * (nz + 1)
    halo = None
    if world > 1:
        from femb200 import dist as fdist
        halo = fdist.SlabHalo(rank, world, i1 - i0, ny, nz)
    h = c.Handle(verts, tets, center, MAT["rho"], mu, lam, MAT["damping"], implicit=(kind != "explicit"),
                 tile_elems=args.tile, device=local_rank, halo_verts=None if halo is None else halo.vertex_ids,
                 halo_owned=None if halo is None else halo.owned)
    stream = torch.cuda.Stream()          # a real stream: the legacy default stream cannot be graph-captured
    torch.cuda.set_stream(stream)
    h.set_stream(stream.cuda_stream)
    if halo is not None:
        halo.attach(h, mode=args.comm)
    info = h.info()
    vel0 = np.zeros_like(pos0)
    h.set(c.F_POS, pos0)
    dt = w["dt"]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def reset_state():
        h.set(c.F_POS, pos0)
        h.set(c.F_VEL, vel0)

    def run_steps(n):
        if kind == "explicit":
            done = 0
            while done < n:
                m = min(RESET_EVERY, n - done)
                if halo is None:
                    h.explicit_step(dt, G_DIR, m)
                else:
                    halo.explicit_steps(h, dt, G_DIR, m)
                done += m
                if done < n:
                    reset_state()
        else:
            # one solve of exactly n iterations (tol = 0 => fixed count, no host polling)
            method, pre = (0 if kind == "jacobi" else 1), (1 if kind == "cg_ata" else 0)
            if halo is None:
                h.implicit_solve(dt, method, pre, max_iter=n, tol=0.0, fetch=False)
            else:
                halo.implicit_solve(h, dt, method, pre, max_iter=n, tol=0.0, fetch=False)

    launches0 = None
    run_steps(args.warmup)
    # prime the CUDA graph of the timed chunk size (capture + instantiate must not be timed), then
    # put the state back so the timed region starts from the same configuration
    run_steps(min(RESET_EVERY, args.steps) if kind == "explicit" else args.steps)
    reset_state()
    barrier()
    replayed0 = getattr(halo, "replayed_launches", 0) if halo is not None else 0
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = c.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    run_steps(args.steps)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = c.launch_count() - launches0
    if halo is not None:   # kernels inside replayed torch CUDA graphs are not seen by the library's counter
        launches += getattr(halo, "replayed_launches", 0) - replayed0
    clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    finite = bool(np.isfinite(h.get(c.F_POS)).all())
    iters_done = None
    if kind != "explicit":
        # the solve must really have run `steps` iterations (no early exit on convergence / NaN)
        finite = finite and bool(np.isfinite(h.get(c.F_VEC_X)).all())
        iters_done = int(h.last_solve_iters())
        if iters_done != args.steps:
            raise SystemExit("solver ran %d iterations, expected %d: timing invalid" % (iters_done, args.steps))

    # ---- per-kernel roofline of the dominant kernel (tile kernel), timed live with CUDA events
    roof = None
    peak, peak_src = peaks()
    if kind == "explicit":
        reset_state()
        h.explicit_step_timed(dt, G_DIR, 5)
        t_ms, b_ms = h.explicit_step_timed(dt, G_DIR, 50)
        ne_l, nv_l = len(tets), len(verts)
        tile_bytes = 56 * ne_l + 88 * info["n_interior"] + 12 * info["n_boundary"]
        ach = tile_bytes / (t_ms / 50 * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "k_tile<3,OpExplicit<3,EX_STEP>>", "achieved": round(ach, 1),
                "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4), "traffic": None, "peak_source": peak_src,
                "bytes_per_launch": tile_bytes, "tile_ms": round(t_ms / 50, 5), "boundary_ms": round(b_ms / 50, 5),
                "step_bytes": algorithmic_bytes(kind, ne_l, nv_l),
                "step_frac": round(algorithmic_bytes(kind, ne_l, nv_l) / ((t_ms + b_ms) / 50 * 1e-3) / 1e9 / peak, 4)}
    else:
        # dominant kernel = the operator pass's tile kernel: 52 B/tet (indices + H_e), per interior vertex d in,
        # q out, 1/m (28 B), per boundary vertex d in (12 B); timed live, one launch at a time
        ne_l, nv_l = len(tets), len(verts)
        bts = algorithmic_bytes(kind, ne_l, nv_l)
        it_ach = bts / (ms / args.steps * 1e-3) / 1e9
        roof = {"bound": "hbm", "peak": peak, "unit": "GB/s", "traffic": None, "peak_source": peak_src,
                "iteration_bytes": bts, "iteration_achieved": round(it_ach, 1), "iteration_frac": round(it_ach / peak, 4)}
        if halo is None:
            h.implicit_apply_timed(dt, 0, 5)
            t_ms, b_ms = h.implicit_apply_timed(dt, 0, 50)
            tile_bytes = 52 * ne_l + 28 * info["n_interior"] + 12 * info["n_boundary"]
            ach = tile_bytes / (t_ms / 50 * 1e-3) / 1e9
            roof.update({"kernel": "k_tile_persistent<3,OpApply<3,0>> (operator pass A d)", "achieved": round(ach, 1),
                         "frac": round(ach / peak, 4), "bytes_per_launch": tile_bytes, "tile_ms": round(t_ms / 50, 5),
                         "boundary_ms": round(b_ms / 50, 5)})
        else:
            roof.update({"kernel": "solver iteration (all kernels)", "achieved": round(it_ach, 1),
                         "frac": round(it_ach / peak, 4), "bytes_per_launch": bts})

    # measured DRAM traffic of the dominant kernel (one ncu --set full capture, committed under profiles/)
    tpath = os.path.join(ROOT, "profiles", "r01_traffic.json")
    if os.path.exists(tpath) and world == 1:
        tr = json.load(open(tpath)).get(args.workload)
        if tr:
            roof["traffic"] = tr["traffic_bytes"]
            roof["traffic_source"] = tr["source"]

    # ---- e2e: the same step through the reference-shaped Python API with pinned host buffers
    e2e = measure_e2e(args, h, c, kind, w, pos0, vel0, ne_global, world, torch, barrier, halo)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(kind, w, sample_cells=args.cpu_cells)
    if rank == 0:
        units = ne_global if kind == "explicit" else 1
        value = units * args.steps / (ms * 1e-3)
        out = {
            "metric": "tet_element_updates_per_sec" if kind == "explicit" else "solver_iterations_per_sec",
            "value": value, "unit": "tet-updates/s" if kind == "explicit" else "iterations/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic Kuhn tet box, seeded 10% h vertex jitter, random-free material from config/demo_3d.json",
            "config": {"workload": args.workload, "cells": list(w["cells"]), "n_tets": ne_global, "n_verts": nv_global,
                       "dt": dt, "partition": "x-slabs, %d rank(s)" % world,
                       "comm": None if halo is None else ("peer memory (in-kernel NVLink stores + flags)" if halo.peer
                                                          else "NCCL P2P + all_reduce"), "tile_elems": info["tile_elems"],
                       "n_tiles_rank0": info["n_tiles"], "cache": "inputs %.0f MB per rank > 126 MB L2" %
                       (algorithmic_bytes(kind, len(tets), len(verts)) / 1e6),
                       "state_reset_every": RESET_EVERY if kind == "explicit" else None, "state_finite": finite,
                       "solver_iterations_run": iters_done},
            "clocks": clocks, "gpu_launches": int(launches), "roofline": roof, "e2e": e2e, "cpu_baseline": cpu,
        }
        print(json.dumps(out))
    if halo is not None:
        halo.close()
    h.close()
    if world > 1:
        fdist.shutdown()


def measure_e2e(args, h, c, kind, w, pos0, vel0, ne_global, world, torch, barrier, halo=None):
    """Same metric through the public, reference-shaped Python API (solver.explicit.neo_hookean_1_grad +
    solver.kinematic.kinematic / solver.implicit.implicit_solver_neo_hookean) with HOST buffers: every step
    copies the step's state in from pinned host memory and reads the result back (to_numpy/from_numpy)."""
    import constants
    import utils
    from circle_blocks import CircleBlocks
    from femb200 import fields
    from solver.explicit import neo_hookean_1_grad
    from solver.kinematic import kinematic

    if kind != "explicit":
        return {"value": None, "unit": "iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": "e2e is reported for the explicit workloads only"}
    utils.sys_init({"dim": 3, "auto_diff": False, "use_explicit_method": True, "delta_time": w["dt"], "g_dir": G_DIR,
                    "implicit_method": 1, "preconditioned": 1})

    class _Obj:  # the object.Object surface the solver wrappers touch, around the already-built handle
        pass

    obj = _Obj()
    obj._h, obj.implicit, obj.particle_cnt = h, False, h.n_vert
    obj._pending_grad, obj._force_clean, obj._blocks_seen = False, True, None
    from object import Object
    obj._sync_blocks = Object._sync_blocks.__get__(obj)
    obj._materialize_force = Object._materialize_force.__get__(obj)
    obj.particles = fields.Particles(obj)
    blocks = CircleBlocks([{"id": 0, "block_center": [0.5, 0.5, 0.5], "block_radius": 0.0}])  # demo_3d.json:24-28
    pin_pos = torch.from_numpy(pos0.copy()).pin_memory()
    pin_vel = torch.from_numpy(vel0.copy()).pin_memory()
    pin_out = torch.empty_like(pin_pos).pin_memory()
    np_pos, np_vel, np_out = pin_pos.numpy(), pin_vel.numpy(), pin_out.numpy()
    steps = max(3, min(args.e2e_steps, args.steps))

    def one():
        obj.particles.pos.from_numpy(np_pos)      # H2D
        obj.particles.vel.from_numpy(np_vel)      # H2D
        if halo is None:
            neo_hookean_1_grad(obj)
            kinematic(obj, blocks)
        else:                                     # same step, split around the slab halo exchange
            halo.explicit_steps(h, w["dt"], G_DIR, 1)
        h.get(c.F_POS, out=np_out)                # D2H (== particles.pos.to_numpy() into a pinned buffer)

    for _ in range(3):
        one()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    barrier()
    dt_s = time.perf_counter() - t0
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([dt_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt_s = float(t.item())
    return {"value": ne_global * steps / dt_s, "unit": "tet-updates/s", "h2d_bytes_per_step": int(np_pos.nbytes + np_vel.nbytes),
            "d2h_bytes_per_step": int(np_out.nbytes), "steps": steps, "ms_per_step": dt_s / steps * 1e3,
            "api": "particles.pos/vel.from_numpy -> neo_hookean_1_grad -> kinematic -> particles.pos.to_numpy"}


def cpu_problem(w, cells):
    from oracle import fem_oracle as o

    verts, tets = o.kuhn_box(*cells, w["h"])
    pos0 = o.particles_init(verts.astype(np.float32), w["offset"])
    ref, vol, mass = o.elements_init(pos0, tets, MAT["rho"])
    rng = np.random.default_rng(0)
    pos = (pos0 + np.float32(0.1 * w["h"]) * rng.uniform(-1, 1, pos0.shape).astype(np.float32)).astype(np.float32)
    return tets, pos, ref, vol, mass


def cpu_run(kind, w, cells, steps):
    """Times `steps` steps of the oracle's C/OpenMP port on a `cells` sample of the workload; returns (units/s, s)."""
    from oracle import cref
    from oracle import fem_oracle as o

    mu, lam = o.lame(MAT["E"], MAT["nu"])
    tets, pos, ref, vol, mass = cpu_problem(w, cells)
    if kind == "explicit":
        vel, force = np.zeros_like(pos), np.zeros_like(pos)
        cref.explicit_steps(pos, vel, force, mass, tets, ref, vol, mu, lam, w["dt"], G_DIR, MAT["damping"], [], 1)
        t0 = time.perf_counter()
        cref.explicit_steps(pos, vel, force, mass, tets, ref, vol, mu, lam, w["dt"], G_DIR, MAT["damping"], [], steps)
        t = time.perf_counter() - t0
        return len(tets) * steps / t, t
    H = cref.hessian_blocks(pos, tets, ref, vol, mu, lam)
    b = o.compute_linear_system_vector_b(pos, np.zeros_like(pos), tets, ref, vol, mass, mu, lam, w["dt"])
    t0 = time.perf_counter()
    _, it, _ = cref.cg(H, tets, mass, w["dt"], b, 1 if kind == "cg_ata" else 0, iter_max=steps, tol=0.0)
    t = time.perf_counter() - t0
    return it / t, t


def cpu_baseline(kind, w, sample_cells, budget_s=12.0):
    """The reference algorithm (oracle C/OpenMP port) on all host cores, on a bounded sample of the workload:
    a short calibration run sizes the step count to about `budget_s` seconds of CPU work."""
    from oracle import cref

    cells = tuple(min(a, b) for a, b in zip(w["cells"], sample_cells))
    ne = cells[0] * cells[1] * cells[2] * 6
    _, t_cal = cpu_run(kind, w, cells, 3)
    n = int(max(5, min(2000, budget_s / max(t_cal / 3, 1e-6))))
    value, t = cpu_run(kind, w, cells, n)
    scale = 1.0 if kind == "explicit" else ne / (w["cells"][0] * w["cells"][1] * w["cells"][2] * 6)
    return {"value": value * scale, "unit": "tet-updates/s" if kind == "explicit" else "iterations/s (scaled to the full mesh)",
            "cores": cref.threads(), "kind": "port",
            "sample": "%dx%dx%d-cell Kuhn box (%d tets) of the same workload, %.1f s of CPU work, C/OpenMP restatement "
                      "(taichi unavailable: the reference's own ti.cpu path cannot run here)" % (*cells, ne, t)}


def run_reference(args):
    """--impl reference: the reference algorithm on the host cores (oracle C/OpenMP port -- the reference is
    Taichi-JIT Python and taichi is not installable, so kind = "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import cref

    w = dict(WORKLOADS[args.workload])
    if args.x_cells:      # diagnostic only: a thinner box (e.g. the per-rank slab of a larger N); not a bench line
        w["cells"] = (args.x_cells,) + tuple(w["cells"][1:])
    kind = w["kind"]
    full = w["cells"][0] * w["cells"][1] * w["cells"][2] * 6
    # exactly --steps steps; the sample (a sub-box of the workload) is shrunk until they fit ~90 s of CPU work
    cells = tuple(min(a, b) for a, b in zip(w["cells"], args.cpu_cells))
    while True:
        _, t_cal = cpu_run(kind, w, cells, 2)
        if t_cal / 2 * (args.steps + args.warmup) <= 90.0 or max(cells) <= 16:
            break
        cells = tuple(max(16, c * 3 // 4) for c in cells)
    ne = cells[0] * cells[1] * cells[2] * 6
    cpu_run(kind, w, cells, max(1, args.warmup))
    steps = max(1, args.steps)
    value, t = cpu_run(kind, w, cells, steps)
    if kind != "explicit":
        value *= ne / full
    sample = "%dx%dx%d-cell sample (%d tets) of %s, %d step(s)" % (*cells, ne, args.workload, steps)
    print(json.dumps({
        "impl": "reference", "metric": "tet_element_updates_per_sec" if kind == "explicit" else "solver_iterations_per_sec",
        "value": value, "unit": "tet-updates/s" if kind == "explicit" else "iterations/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": max(1, args.warmup), "ms_per_step": t / steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic Kuhn tet box (bounded sample of the workload)",
        "config": {"workload": args.workload, "cells": list(w["cells"]), "n_tets": full, "sample": sample},
        "cpu_baseline": {"value": value, "unit": "tet-updates/s" if kind == "explicit" else "iterations/s",
                         "cores": cref.threads(), "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "tet-updates/s" if kind == "explicit" else "iterations/s",
                "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4_explicit", choices=sorted(WORKLOADS))
    ap.add_argument("--tile", type=int, default=0, help="elements per tile (0 = library default)")
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--cpu-cells", type=int, nargs=3, default=[96, 96, 96], help="CPU sample size in cells")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--x-cells", type=int, default=0, help="diagnostic: override the box's cell count along x")
    ap.add_argument("--comm", default=None, choices=["peer", "nccl"], help="multi-GPU transport (default: peer memory)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
