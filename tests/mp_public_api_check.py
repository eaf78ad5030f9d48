"""Run under torchrun (world_size >= 2, one GPU per rank): the reference-shaped PUBLIC calls on a box that
object.Object cuts into x-slabs by itself, against the same calls on one GPU (partition=False on rank 0).
main.py:101-113's loop, unchanged: fem(obj) / kinematic / advect_implicit.  Prints 'PUBLIC_SLAB_OK' on rank 0."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "finite-element-method_b200"), ROOT]


def config(world, explicit, method=1, pre=1, dt=1e-4):
    return {"dim": 3, "delta_time": dt, "sim_count": 1, "auto_diff": False, "use_explicit_method": explicit,
            "implicit_method": method, "preconditioned": pre, "g_dir": [0, -1, 0],
            "objects": [{"id": 0, "rho": 1000, "center": [0.3, 0.5, 0.3], "cells": [4 * world + 1, 6, 5], "h": 0.4 / 55,
                         "E": 4e4, "nu": 0.4, "damping": 10, "jitter": 0.1, "seed": 0, "verbose": False}],
            "blocks": [{"id": 0, "block_center": [0.32, 0.48, 0.32], "block_radius": 0.03}]}


def run(cfg, steps, partition, vel=None):
    """main.py:36-61 + 101-113 on the product modules."""
    import utils
    utils.sys_init(cfg)
    import constants
    import solver.kinematic as ki
    from circle_blocks import CircleBlocks
    from object import Object
    from solver.explicit import neo_hookean_1_grad
    from solver.implicit import advect_implicit, implicit_solver_neo_hookean
    obj = Object(cfg["objects"][0], tile_elems=96, partition=partition)
    blocks = CircleBlocks(cfg["blocks"])
    if vel is not None:
        if obj._slab is not None:
            _, _, i0, i1, (nx, ny, nz) = obj._slab
            plane = (ny + 1) * (nz + 1)
            vel = vel[i0 * plane:(i1 + 1) * plane]
        obj.particles.vel.from_numpy(vel)
    its = []
    for _ in range(steps):
        if constants.use_explicit_method:
            neo_hookean_1_grad(obj)
            ki.kinematic(obj, blocks)
        else:
            implicit_solver_neo_hookean(obj)
            its.append(obj.last_solve["iters"])
            advect_implicit(obj, blocks)
    out = dict(pos=obj.gather(obj.particles.pos), vel=obj.gather(obj.particles.vel), mass=obj.gather(obj.particles.mass),
               local_pos=obj.particles.pos.to_numpy(), its=its, slab=obj._slab, peer=None if obj._halo is None else obj._halo.peer)
    if obj._halo is not None:
        obj._halo.close()
    obj._h.close()
    return out


def rel(a, b):
    return float(np.linalg.norm(a.astype(np.float64) - b) / max(np.linalg.norm(b), 1e-300))


def main():
    import torch
    from femb200 import dist as fdist
    rank, world = fdist.ensure_process_group()
    torch.cuda.set_device(fdist.local_device())
    ok = True
    nv = (4 * world + 2) * 7 * 6
    gvel = (0.05 * np.random.default_rng(1).uniform(-1, 1, (nv, 3))).astype(np.float32)
    cases = [("explicit", config(world, True), 24, None, 1e-6),
             ("cg", config(world, False, 1, 0, 1e-4), 3, gvel, 1e-6),
             ("cg_ata", config(world, False, 1, 1, 5e-4), 3, gvel, 1e-6),
             ("jacobi", config(world, False, 0, 0, 1e-4), 3, gvel, 1e-6)]
    for name, cfg, steps, vel, tol in cases:
        got = run(cfg, steps, True, vel)
        assert got["slab"] is not None and got["pos"].shape == (nv, 3), (got["slab"], got["pos"].shape)
        # every rank gathered the same global arrays
        chk = fdist.all_gather_arrays(got["pos"][None])
        same = all(np.array_equal(chk[0], c) for c in chk)
        if rank == 0:
            want = run(cfg, steps, False, vel)
            e_pos, e_vel, e_m = rel(got["pos"], want["pos"]), rel(got["vel"], want["vel"]), rel(got["mass"], want["mass"])
            good = same and e_pos <= tol and e_m <= 1e-6 and e_vel <= (1e-4 if name == "explicit" else 2e-3)
            good = good and (not got["its"] or max(abs(a - b) for a, b in zip(got["its"], want["its"])) <= 2)
            print(name, "pos %.2e vel %.2e mass %.2e iters %s/%s peer=%s %s" % (e_pos, e_vel, e_m, got["its"], want["its"],
                                                                            got["peer"], "ok" if good else "FAIL"), flush=True)
            ok = ok and good
        import torch.distributed as dist
        dist.barrier()
    if rank == 0:
        print("PUBLIC_SLAB_OK" if ok else "PUBLIC_SLAB_FAIL", flush=True)
    fdist.shutdown()


if __name__ == "__main__":
    main()
