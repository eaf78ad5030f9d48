"""Run under torchrun (world_size >= 2, one GPU per rank): slab-partitioned explicit steps vs the
same global mesh on one GPU.  Prints 'SLAB_OK' on rank 0 when every check passes."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "finite-element-method_b200"), ROOT]
from femb200 import _capi as c, dist as fdist, meshgen  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    nx, ny, nz, hh = 4 * world + 1, 6, 5, 0.4 / 55
    center = np.array([0.3, 0.5, 0.3], np.float32)
    E, nu = 4e4, 0.4
    mu, lam = E / 2 / (1 + nu), E * nu / (1 + nu) / (1 - 2 * nu)
    plane = (ny + 1) * (nz + 1)
    jit = np.random.default_rng(0).uniform(-1, 1, ((nx + 1) * plane, 3)).astype(np.float32)
    gverts, gtets = meshgen.kuhn_box(nx, ny, nz, hh)
    gpos = (gverts + center).astype(np.float32) + np.float32(0.1 * hh) * jit
    i0, i1 = fdist.slab_range(nx, rank, world)
    verts, tets = meshgen.kuhn_box(nx, ny, nz, hh, x_cells=(i0, i1))
    halo = fdist.SlabHalo(rank, world, i1 - i0, ny, nz)
    h = c.Handle(verts, tets, center, 1000.0, mu, lam, 10.0, implicit=True, tile_elems=96, halo_verts=halo.vertex_ids,
                 halo_owned=halo.owned)
    h.set_stream(stream.cuda_stream)
    halo.attach(h)
    h.set(c.F_POS, gpos[i0 * plane:(i1 + 1) * plane])
    dt, g, n = 1e-4, [0, -1, 0], 24
    halo.explicit_steps(h, dt, g, 12, use_graph=False)
    halo.explicit_steps(h, dt, g, 12, use_graph=True)     # CUDA-graph path (kernels + NCCL P2P)
    torch.cuda.synchronize()
    mine = dict(pos=h.get(c.F_POS), vel=h.get(c.F_VEL), mass=h.get(c.F_MASS), i0=i0, i1=i1)
    # ---- implicit path on the slabs: every solver, eager and graph-replayed, then advect
    sl_me = slice(i0 * plane, (i1 + 1) * plane)
    gvel = (0.05 * np.random.default_rng(1).uniform(-1, 1, gpos.shape)).astype(np.float32)
    SOLVES = [("cg", 1, 0, 1e-4, False), ("cg_ata", 1, 1, 5e-4, True), ("jacobi", 0, 0, 1e-4, True)]
    for name, method, pre, sdt, use_graph in SOLVES:
        h.set(c.F_POS, gpos[sl_me])
        h.set(c.F_VEL, gvel[sl_me])
        h.set(c.F_VEL_G, np.zeros_like(gvel[sl_me]))
        for _ in range(2):      # second solve replays the cached graph
            it, res = halo.implicit_solve(h, sdt, method, pre, use_graph=use_graph)
            h.advect_implicit(sdt, g)
        torch.cuda.synchronize()
        mine[name] = dict(vel=h.get(c.F_VEL), pos=h.get(c.F_POS), b=h.get(c.F_VEC_B), it=it, res=res)
    allr = [None] * world
    dist.all_gather_object(allr, mine)
    ok = True
    if rank == 0:
        hs = c.Handle(gverts, gtets, center, 1000.0, mu, lam, 10.0, implicit=True, tile_elems=96)
        hs.set(c.F_POS, gpos)
        hs.explicit_step(dt, g, n)
        want_p, want_v, want_m = hs.get(c.F_POS), hs.get(c.F_VEL), hs.get(c.F_MASS)
        for r, d in enumerate(allr):
            sl = slice(d["i0"] * plane, (d["i1"] + 1) * plane)
            for name, got, want in (("pos", d["pos"], want_p[sl]), ("vel", d["vel"], want_v[sl]), ("mass", d["mass"], want_m[sl])):
                err = np.linalg.norm(got - want) / np.linalg.norm(want)
                if not err <= (1e-6 if name != "vel" else 1e-4):
                    ok = False
                    print("rank", r, name, "rel err", err)
            if r + 1 < world:   # the shared plane is bit-identical on both owners
                nxt = allr[r + 1]
                for name in ("pos", "vel", "mass"):
                    if not np.array_equal(d[name][-plane:], nxt[name][:plane]):
                        ok = False
                        print("shared plane differs between ranks", r, r + 1, name)
        for name, method, pre, sdt, _ in SOLVES:
            hs.set(c.F_POS, gpos)
            hs.set(c.F_VEL, gvel)
            hs.set(c.F_VEL_G, np.zeros_like(gvel))
            for _ in range(2):
                it1, res1 = hs.implicit_solve(sdt, method, pre)
                hs.advect_implicit(sdt, g)
            want = dict(vel=hs.get(c.F_VEL), pos=hs.get(c.F_POS), b=hs.get(c.F_VEC_B))
            for r, d in enumerate(allr):
                sl = slice(d["i0"] * plane, (d["i1"] + 1) * plane)
                for f, tol in (("b", 1e-4), ("vel", 2e-3), ("pos", 1e-6)):   # second step: b inherits the solver tolerance
                    err = np.linalg.norm(d[name][f] - want[f][sl]) / np.linalg.norm(want[f][sl])
                    if not err <= tol:
                        ok = False
                        print("solve", name, "rank", r, f, "rel err", err)
                if d[name]["it"] != allr[0][name]["it"]:
                    ok = False
                    print("solve", name, "iteration counts differ between ranks")
                if r + 1 < world and not np.array_equal(d[name]["vel"][-plane:], allr[r + 1][name]["vel"][:plane]):
                    ok = False
                    print("solve", name, "shared plane velocity differs between ranks", r, r + 1)
            print("solve", name, "iters slab/single", allr[0][name]["it"], it1, "resid", allr[0][name]["res"], res1)
        print("COMM", "peer" if halo.peer else "nccl")
        print("SLAB_OK" if ok else "SLAB_FAIL")
    halo.close()
    h.close()
    fdist.shutdown()


if __name__ == "__main__":
    main()
