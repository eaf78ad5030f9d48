#!/usr/bin/env python
"""Generates tests/golden/ref_*.npz by running the UNMODIFIED reference sources on the CPU.

The reference (/root/reference: solver/explicit.py:8, solver/explicit_auto_diff.py:8 under ti.ad.Tape,
solver/kinematic.py:9, solver/implicit.py:201 and :407, object.py:337-362, circle_blocks.py:15, main.py:8-12 and
the step order of main.py:101-113) is imported as it lies and executed by oracle/ti_shim (a plain-Python, numpy
scalar stand-in for the `taichi` names those files use; the Taichi semantics it assumes are listed in its module
docstring and in DESIGN.md section 0).  The vectors written here are therefore REFERENCE outputs, not outputs of
oracle/fem_oracle.py.  /root/reference exists only in the build container: the GPU box reads the committed .npz.

    python tests/golden/make_ref_golden.py [case ...]        # all cases: about 10 minutes on 8 cores

Each case stores its inputs (config JSON, mesh, initial pos / vel) and, for the recorded sub-steps, what the
reference's kernels produced (names below).  One file per case: tests/golden/ref_<case>.npz.
"""
import copy
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref_harness as rh  # noqa: E402

DEFAULT = json.load(open(os.path.join(rh.REFERENCE, "default.json"))) if rh.available() else None
DEMO3D = json.load(open(os.path.join(rh.REFERENCE, "config", "demo_3d.json"))) if rh.available() else None

_KUHN = ((0, 1, 2), (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0))


def kuhn_cube(n, side):
    """SURVEY.md A.6 structured cube: vertex (i*(n+1)+j)*(n+1)+k at (i,j,k)*h, six tets per cell."""
    h = side / n
    g = np.arange(n + 1)
    i, j, k = np.meshgrid(g, g, g, indexing="ij")
    verts = np.stack([i, j, k], -1).reshape(-1, 3) * h
    s = ((n + 1) * (n + 1), n + 1, 1)
    tets = []
    for ci in range(n):
        for cj in range(n):
            for ck in range(n):
                v0 = ci * s[0] + cj * s[1] + ck * s[2]
                for p in _KUHN:
                    a = v0 + s[p[0]]
                    b = a + s[p[1]]
                    tets.append((v0, a, b, b + s[p[2]]))
    return verts.astype(np.float64), np.asarray(tets, np.int32)


def asset_mesh(name):
    sys.path.insert(0, os.path.join(ROOT, "finite-element-method_b200"))
    from femb200 import meshio
    nodes, tets, _ = meshio.tetrahedralize_file(os.path.join(ROOT, "finite-element-method_b200", "obj", name))
    return nodes.astype(np.float64), tets.astype(np.int32)


def cfg2d(**kw):
    c = copy.deepcopy(DEFAULT)
    obj = kw.pop("obj", {})
    c.update(kw)
    c["objects"][0].update(obj)
    return c


def cfg3d(**kw):
    c = copy.deepcopy(DEMO3D)
    obj = kw.pop("obj", {})
    c.update(kw)
    c["objects"][0].update(obj)
    c["objects"][0].pop("obj", None)
    return c


def stretch(pos, amount, jitter, seed):
    """10 % stretch about the centroid in x + seeded jitter (forces above rounding noise)."""
    rng = np.random.default_rng(seed)
    c = pos.mean(0)
    p = pos.astype(np.float64).copy()
    p[:, 0] = c[0] + (p[:, 0] - c[0]) * (1.0 + amount)
    p += jitter * rng.uniform(-1, 1, p.shape)
    return p.astype(np.float32)


# name -> (config, mesh (verts, tets) or None, initial-state recipe, steps, recorded steps, keep dense A)
def cases():
    sq = lambda s: {"subdivisions": s}  # noqa: E731
    k4 = kuhn_cube(4, 0.4)
    k3 = kuhn_cube(3, 0.3)
    cube = {"center": [0.3, 0.5, 0.3]}
    blk3 = [{"id": 0, "block_center": [0.5, 0.296, 0.5], "block_radius": 0.2}, {"id": 1, "block_center": [0.1, 0.1, 0.1], "block_radius": 0.0}]
    walls = {"center": [0.75, -0.15, 0.75]}
    rec100 = (1, 2, 10, 50, 100)
    C = {}
    # C1 = default.json (BASELINE config 1), as shipped (autodiff) and analytic
    C["c1_autodiff"] = (cfg2d(), None, None, 100, rec100)
    C["c1_analytic"] = (cfg2d(auto_diff=False), None, None, 100, rec100)
    C["c1_stretch_analytic"] = (cfg2d(auto_diff=False), None, dict(stretch=0.10, jitter=0.002, seed=1), 100, rec100)
    C["c1_stretch_autodiff"] = (cfg2d(), None, dict(stretch=0.10, jitter=0.002, seed=1), 100, rec100)
    C["c1_blocks"] = (cfg2d(auto_diff=False), None, None, 1500, (500, 1000, 1250, 1500))
    # 3D explicit on the 4^3 Kuhn cube (demo_3d.json material, dt), sphere block + disabled block
    C["k4_analytic"] = (cfg3d(obj=cube, blocks=blk3), k4, dict(stretch=0.10, jitter=0.01, seed=2), 100, rec100)
    C["k4_autodiff"] = (cfg3d(obj=cube, blocks=blk3, auto_diff=True), k4, dict(stretch=0.10, jitter=0.01, seed=2), 20, (1, 2, 10, 20))
    C["k4_block"] = (cfg3d(obj=cube, blocks=blk3), k4, dict(vel=(0.0, -1.0, 0.0)), 300, (100, 200, 300))
    C["k4_walls"] = (cfg3d(obj=walls), k4, dict(stretch=0.05, jitter=0.01, seed=3, rand_vel=0.5), 60, (1, 10, 30, 60))
    # implicit (BASELINE config 2's solver on meshes the dense reference can hold)
    imp = dict(auto_diff=False, use_explicit_method=False)
    st2 = dict(stretch=0.10, jitter=0.002, seed=1)
    st3 = dict(stretch=0.10, jitter=0.008, seed=2)
    C["c1_jacobi"] = (cfg2d(implicit_method=0, **imp), None, st2, 10, (1, 2, 5, 10))
    C["c1_cg"] = (cfg2d(implicit_method=1, preconditioned=0, **imp), None, st2, 10, (1, 2, 5, 10))
    C["c1_cg_ata"] = (cfg2d(implicit_method=1, preconditioned=1, **imp), None, st2, 2, (1, 2))
    C["s5_jacobi"] = (cfg2d(implicit_method=0, obj=sq(5), **imp), None, st2, 100, rec100, True)
    C["s5_cg"] = (cfg2d(implicit_method=1, preconditioned=0, obj=sq(5), **imp), None, st2, 100, rec100, True)
    C["s5_cg_ata"] = (cfg2d(implicit_method=1, preconditioned=1, obj=sq(5), **imp), None, st2, 100, rec100, True)
    C["k3_jacobi"] = (cfg3d(implicit_method=0, obj=cube, blocks=blk3, **imp), k3, st3, 20, (1, 2, 10, 20), True)
    C["k3_cg"] = (cfg3d(implicit_method=1, preconditioned=0, obj=cube, blocks=blk3, **imp), k3, st3, 20, (1, 2, 10, 20), True)
    C["k3_cg_ata"] = (cfg3d(implicit_method=1, preconditioned=1, obj=cube, blocks=blk3, **imp), k3, st3, 6, (1, 2, 6), True)
    C["k3_walls_cg"] = (cfg3d(implicit_method=1, preconditioned=0, obj=walls, **imp), k3, dict(stretch=0.05, jitter=0.008, seed=3, rand_vel=0.5), 40, (1, 10, 20, 40))
    C["k3_cg_stiff"] = (cfg3d(implicit_method=1, preconditioned=0, delta_time=4e-3, obj=cube, blocks=blk3, **imp), k3, st3, 10, (1, 2, 5, 10), True)
    C["k3_cg_ata_stiff"] = (cfg3d(implicit_method=1, preconditioned=1, delta_time=4e-3, obj=cube, blocks=blk3, **imp), k3, st3, 4, (1, 2, 4), True)
    # BASELINE configs 2 and 5 on the assets they name (config/demo_3d.json:2-28 with the edits SURVEY.md 8 lists).
    # TetGen is not installable: the tet mesh comes from the product's fallback mesher (femb200.meshio) and the SAME
    # arrays are stored in the fixture and fed to every side.
    spot100, spot = asset_mesh("spot100.obj"), asset_mesh("spot.obj")
    sts = dict(stretch=0.05, jitter=0.002, seed=4)
    C["spot100_cg"] = (cfg3d(implicit_method=1, preconditioned=0, **imp), spot100, sts, 100, rec100)
    C["spot100_cg_ata"] = (cfg3d(implicit_method=1, preconditioned=1, **imp), spot100, sts, 3, (1, 2, 3))
    C["spot100_jacobi"] = (cfg3d(implicit_method=0, **imp), spot100, sts, 30, (1, 2, 10, 30))
    C["spot_analytic"] = (cfg3d(), spot, sts, 100, rec100)
    C["spot_autodiff"] = (cfg3d(auto_diff=True), spot, sts, 20, (1, 2, 10, 20))
    # stiff Jacobi (dt 4x): the serial in-place sweep vs a simultaneous Jacobi part ways here
    C["s5_jacobi_stiff"] = (cfg2d(implicit_method=0, delta_time=2e-3, obj=sq(5), **imp), None, st2, 6, (1, 2, 3, 6), True)
    return C


def run_case(name):
    np.seterr(all="ignore")
    spec = cases()[name]
    config, mesh, init, steps, rec = spec[:5]
    keep_A = len(spec) > 5 and spec[5]
    t0 = time.time()
    ref = rh.load_reference(config)
    c = ref.constants
    blocks = ref.circle_blocks.CircleBlocks(config["blocks"])
    ocfg = config["objects"][0]
    _stdout, sys.stdout = sys.stdout, open(os.devnull, "w")
    obj = rh.make_object(ref, ocfg, *(mesh or ()))
    sys.stdout = _stdout
    out = {"config": np.array(json.dumps(config))}
    P = obj.particles
    out["raw_verts"] = obj.ti_vertices.to_numpy()
    out["elems"] = obj.elements.vertex_indices.to_numpy()
    out["faces"] = obj.ti_faces.to_numpy()
    out["pos0"] = P.pos.to_numpy()
    out["ref"] = obj.elements.ref.to_numpy()
    out["volume"] = obj.elements.volume.to_numpy()
    out["mass"] = P.mass.to_numpy()
    init = init or {}
    pos = out["pos0"]
    if "stretch" in init:
        pos = stretch(pos, init["stretch"], init["jitter"], init["seed"])
        P.pos.from_numpy(pos)
    vel = np.zeros_like(pos)
    if "vel" in init:
        vel[:] = np.asarray(init["vel"], np.float32)
    if "rand_vel" in init:
        vel = (init["rand_vel"] * np.random.default_rng(init["seed"] + 100).uniform(-1, 1, pos.shape)).astype(np.float32)
    P.vel.from_numpy(vel)
    out["pos_init"], out["vel_init"] = pos, vel
    implicit = not (c.use_explicit_method or c.auto_diff == 1)
    series = {}

    def put(key, arr):
        series.setdefault(key, []).append(np.array(arr))

    touched = 0
    for step in range(1, steps + 1):
        r = step in rec
        rh.fem_only(ref, obj)
        if r:
            if c.auto_diff == 1:
                put("grad", P.pos.grad.to_numpy())
                put("U", obj.U.to_numpy())
                put("phi", obj.phi.to_numpy())
            elif c.use_explicit_method:
                put("force", P.force.to_numpy())
            else:
                put("b", obj.vec_b.to_numpy())
                put("x", obj.vec_x.to_numpy())
                put("past_x", obj.past_vec_x.to_numpy())
                if keep_A and step <= 2:
                    put("A", obj.matrix_A.to_numpy())
        v_before = P.vel.to_numpy() if not implicit else None
        rh.advance_only(ref, obj, blocks)
        if r:
            put("pos", P.pos.to_numpy())
            put("vel", P.vel.to_numpy())
            if implicit:
                put("vel_g", P.vel_g.to_numpy())
        if not implicit and config["blocks"]:
            # how many vertices sit inside an enabled block (tells the tests the contact branch ran)
            p = P.pos.to_numpy()
            for b in config["blocks"]:
                if b["block_radius"] > 0:
                    touched = max(touched, int((np.linalg.norm(p - np.asarray(b["block_center"]), axis=1) < b["block_radius"]).sum()))
    for k, v in series.items():
        out["s_" + k] = np.stack(v)
    out["rec"] = np.asarray(rec, np.int32)
    out["steps"] = np.int32(steps)
    out["touched_max"] = np.int32(touched)
    path = os.path.join(HERE, "ref_%s.npz" % name)
    np.savez_compressed(path, **out)
    return name, time.time() - t0, os.path.getsize(path)


if __name__ == "__main__":
    assert rh.available(), "needs the reference tree at " + rh.REFERENCE
    names = sys.argv[1:] or list(cases())
    with mp.Pool(min(len(names), os.cpu_count())) as pool:
        for name, dt, size in pool.imap_unordered(run_case, names):
            print("%-22s %7.1f s %8d bytes" % (name, dt, size), flush=True)
