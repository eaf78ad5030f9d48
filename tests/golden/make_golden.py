#!/usr/bin/env python
"""Generates tests/golden/fem_golden.npz from the numpy oracle (oracle/fem_oracle.py).

NOT reference outputs: the reference (Taichi 1.6) cannot run here, so parity stays "unpinned"
(DESIGN.md section 0).  These vectors pin the ORACLE (and through it the CUDA path) against silent
drift between rounds.  Regenerate with:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
from oracle import fem_oracle as o  # noqa: E402
from helpers import MAT2, MAT3, cube_case, square_case  # noqa: E402


def case(prefix, verts, el, center, pos0, pos, mat, dt, g_dir, blocks, out):
    mu, lam = o.lame(mat["E"], mat["nu"])
    ref, vol, mass = o.elements_init(pos0, el, mat["rho"])
    force = np.zeros_like(pos)
    o.neo_hookean_1_grad(pos, el, ref, vol, mu, lam, force)
    phi, U = o.compute_energy(pos, el, ref, vol, mu, lam)
    p, v = pos.copy(), np.zeros_like(pos)
    o.kinematic(p, v, force.copy(), mass, dt, g_dir, mat["damping"], blocks, auto_diff=0)
    vel = (0.05 * np.random.default_rng(7).uniform(-1, 1, pos.shape)).astype(np.float32)
    H = o.element_hessian_block(pos, el, ref, vol, mu, lam)
    b = o.compute_linear_system_vector_b(pos, vel, el, ref, vol, mass, mu, lam, dt)
    d = np.random.default_rng(8).uniform(-1, 1, pos.shape).astype(np.float32)
    Ad = o.apply_A(H, el, mass, dt, d)
    ATd = o.apply_AT(H, el, mass, dt, d)
    x_cg, it_cg, _ = o.cg_matrix_free(H, el, mass, dt, b, 0)
    x_ata, it_ata, _ = o.cg_matrix_free(H, el, mass, dt, b, 1)
    out.update({prefix + k: val for k, val in dict(
        verts=verts, elems=el.astype(np.int32), center=center, pos=pos, ref=ref, vol=vol, mass=mass, force=force,
        phi=phi, U=np.float32(U), kin_pos=p, kin_vel=v, vel=vel, H=H, b=b, d=d, Ad=Ad, ATd=ATd, x_cg=x_cg,
        x_ata=x_ata, iters=np.array([it_cg, it_ata]), dt=np.float32(dt), g_dir=np.asarray(g_dir, np.float32),
        blocks=np.asarray([[*c, r] for c, r in blocks], np.float32).reshape(len(blocks), -1)).items()})


if __name__ == "__main__":
    out = {}
    v, el, c, p0, p = cube_case(n=3)
    case("cube_", v, el, c, p0, p, MAT3, 1e-4, [0, -1, 0], [((0.32, 0.5, 0.32), 0.05)], out)
    v, el, c, p0, p = square_case(s=4)
    case("square_", v, el, c, p0, p, MAT2, 5e-4, [0, -1], [((0.5, 0.8), 0.08)], out)
    path = os.path.join(HERE, "fem_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")
