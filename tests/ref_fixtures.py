"""Helpers for the tests that use the REFERENCE fixtures tests/golden/ref_*.npz (generated from the unmodified
reference sources by tests/golden/make_ref_golden.py): load a case, replay it with the numpy oracle, replay it
on the GPU through the C ABI."""
import glob
import json
import os

import numpy as np

from oracle import fem_oracle as o

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def names():
    return sorted(os.path.basename(p)[4:-4] for p in glob.glob(os.path.join(GOLDEN, "ref_*.npz")))


def load(name):
    z = np.load(os.path.join(GOLDEN, "ref_%s.npz" % name))
    fx = {k: z[k] for k in z.files}
    fx["config"] = json.loads(str(fx["config"]))
    fx["name"] = name
    return fx


def mode(fx):
    c = fx["config"]
    if c["auto_diff"]:
        return "autodiff"
    return "analytic" if c["use_explicit_method"] else "implicit"


def blocks_of(fx):
    return [(b["block_center"], b["block_radius"]) for b in fx["config"]["blocks"]]


def material(fx):
    oc = fx["config"]["objects"][0]
    mu, lam = o.lame(oc["E"], oc["nu"])
    return oc, mu, lam


def oracle_replay(fx, setup=None):
    """Runs oracle/fem_oracle.py over the case's whole trajectory; returns the same series the fixture holds.
    `setup` = (ref, volume, mass) to use instead of the oracle's own (to separate set-up from step errors)."""
    cfg = fx["config"]
    oc, mu, lam = material(fx)
    dt, g = cfg["delta_time"], cfg["g_dir"]
    el = fx["elems"]
    pos0 = o.particles_init(fx["raw_verts"], oc["center"])
    ref, vol, mass = setup or o.elements_init(pos0, el, oc["rho"])
    pos, vel = fx["pos_init"].copy(), fx["vel_init"].copy()
    vel_g = np.zeros_like(pos)
    past_x = np.zeros_like(pos)
    force = np.zeros_like(pos)
    blocks = blocks_of(fx)
    m = mode(fx)
    rec = set(int(s) for s in fx["rec"])
    out = {}

    def put(k, v):
        out.setdefault(k, []).append(np.array(v))

    for step in range(1, int(fx["steps"]) + 1):
        r = step in rec
        if m == "analytic":
            o.neo_hookean_1_grad(pos, el, ref, vol, mu, lam, force)
            if r:
                put("force", force)
            o.kinematic(pos, vel, force, mass, dt, g, oc["damping"], blocks, auto_diff=0)
        elif m == "autodiff":
            phi, U = o.compute_energy(pos, el, ref, vol, mu, lam)
            grad = np.zeros_like(pos)
            o.neo_hookean_1_grad(pos, el, ref, vol, mu, lam, grad)   # analytic gradient == the tape's adjoint (to rounding)
            if r:
                put("grad", grad)
                put("U", U)
                put("phi", phi)
            o.kinematic(pos, vel, grad, mass, dt, g, oc["damping"], blocks, auto_diff=1)
        else:
            A = o.compute_linear_system_matrix_a(pos, el, ref, vol, mass, mu, lam, dt)
            b = o.compute_linear_system_vector_b(pos, vel, el, ref, vol, mass, mu, lam, dt)
            if cfg["implicit_method"] == o.JACOBIN_METHOD:
                x, past_x, it = o.jacobi_iter_field(A, b, past_x)
            else:
                x, it, _ = o.cg(A, b, cfg["preconditioned"])
            vel = x.copy()
            if r:
                put("b", b)
                put("x", x)
                put("past_x", past_x)
                put("iters", it)
                if "s_A" in fx and step <= 2:
                    put("A", A)
            o.advect_implicit(pos, vel, vel_g, dt, g, oc["damping"], blocks)
        if r:
            put("pos", pos)
            put("vel", vel)
            if m == "implicit":
                put("vel_g", vel_g)
    return {k: np.stack(v) for k, v in out.items()}, (ref, vol, mass)
