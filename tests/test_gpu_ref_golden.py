"""GPU parity against the REFERENCE fixtures (tests/golden/ref_*.npz = outputs of the unmodified reference sources
run under oracle/ti_shim, see tests/golden/make_ref_golden.py), driven through the drop-in modules in the reference's
own call order (main.py:38-49 imports, main.py:101-113 step loop) and therefore through the C ABI.

Tolerances are north_star's: per-step forces rel-L2 <= 1e-5 (on deformed states), positions after
100 steps <= 1e-4, fp32.  Solver outputs are compared at convergence level (the reference stops at r.r <= 1e-5
absolute, implicit.py:341, and its "Jacobi" is a serial sweep the GPU does not follow iterate by iterate, H2/H3)."""
import numpy as np
import pytest

import ref_fixtures as rf
from helpers import rel

pytestmark = pytest.mark.gpu
np.seterr(all="ignore")

POS_TOL = 1e-4      # north_star: positions after 100 steps
FORCE_TOL = 1e-5    # north_star: per-step forces


def _run(fx):
    """main.py:36-49 + 101-113 on the product modules; returns the series at the fixture's recorded steps."""
    import utils
    config = fx["config"]
    utils.sys_init(config)
    import constants
    import solver.kinematic as ki
    from circle_blocks import CircleBlocks
    from object import Object
    from solver.explicit import neo_hookean_1_grad
    from solver.explicit_auto_diff import compute_energy
    from solver.implicit import advect_implicit, implicit_solver_neo_hookean

    ocfg = dict(config["objects"][0], verbose=False)
    if constants.dim == 2:
        obj = Object(ocfg)                                      # object.py:95-126 mesh built by the product
    else:
        obj = Object.from_arrays(ocfg, fx["raw_verts"], fx["elems"])
    blocks = CircleBlocks(config.get("blocks"))
    setup = dict(elems=obj.elements.vertex_indices.to_numpy(), pos0=obj.particles.pos.to_numpy(),
                 ref=obj.elements.ref.to_numpy(), volume=obj.elements.volume.to_numpy(), mass=obj.particles.mass.to_numpy())
    obj.particles.pos.from_numpy(fx["pos_init"])
    obj.particles.vel.from_numpy(fx["vel_init"])
    rec = set(int(s) for s in fx["rec"])
    out, iters = {}, []

    def put(k, v):
        out.setdefault(k, []).append(np.array(v))

    for step in range(1, int(fx["steps"]) + 1):
        r = step in rec
        if constants.auto_diff == 0:
            if constants.use_explicit_method:
                neo_hookean_1_grad(obj)
                if r:
                    put("force", obj.particles.force.to_numpy())
            else:
                implicit_solver_neo_hookean(obj)
                iters.append(obj.last_solve["iters"])
                if r:
                    put("b", obj.vec_b.to_numpy())
                    put("x", obj.vec_x.to_numpy())
        else:
            compute_energy(obj)
            if r:
                put("grad", obj.particles.pos.grad.to_numpy())
                put("U", obj.U[None])
                put("phi", obj.phi.to_numpy())
        if constants.use_explicit_method or constants.auto_diff == 1:
            ki.kinematic(obj, blocks)
        else:
            advect_implicit(obj, blocks)
        if r:
            put("pos", obj.particles.pos.to_numpy())
            put("vel", obj.particles.vel.to_numpy())
            if not (constants.use_explicit_method or constants.auto_diff == 1):
                put("vel_g", obj.particles.vel_g.to_numpy())
    return setup, {k: np.stack(v) for k, v in out.items()}, iters


@pytest.mark.parametrize("name", rf.names())
def test_cuda_path_matches_reference_fixture(name):
    fx = rf.load(name)
    setup, out, iters = _run(fx)
    mode = rf.mode(fx)
    rec = [int(s) for s in fx["rec"]]
    # set-up (object.py:337-362): indices identical, Dm^-1 and volume bit-exact, lumped mass to rounding
    assert np.array_equal(setup["elems"], fx["elems"]) and np.array_equal(setup["pos0"], fx["pos0"])
    assert np.array_equal(setup["ref"], fx["ref"]) and np.array_equal(setup["volume"], fx["volume"])
    assert rel(setup["mass"], fx["mass"]) <= 1e-6
    deformed = not (name in ("c1_autodiff", "c1_analytic", "c1_blocks", "k4_block"))
    report = {}
    # first recorded step = the same input state on both sides
    if mode == "analytic" and deformed:
        report["force"] = rel(out["force"][0], fx["s_force"][0])
        assert report["force"] <= FORCE_TOL
    if mode == "autodiff":
        if deformed:
            report["grad"] = rel(out["grad"][0], fx["s_grad"][0])
            assert report["grad"] <= FORCE_TOL
            n_phi = fx["s_phi"].shape[1]
            assert rel(out["phi"][0][:n_phi], fx["s_phi"][0][:n_phi]) <= 1e-5
        assert abs(out["U"][0] - fx["s_U"][0]) <= 1e-5 * max(abs(fx["s_U"][0]), 1e-6) + 1e-9
    if mode == "implicit":
        report["b"] = rel(out["b"][0], fx["s_b"][0])
        assert report["b"] <= FORCE_TOL
        # the solve: both sides stop at an ABSOLUTE residual (r.r <= 1e-5 / ||b - Ax|| <= 1e-5), so x agrees to that level
        report["x"] = rel(out["x"][0], fx["s_x"][0])
        assert report["x"] <= 5e-3
        assert all(0 <= it <= 20000 for it in iters)
    # trajectories: every recorded step up to the end (100 steps where the case has them)
    for k, step in enumerate(rec):
        tol = POS_TOL
        if name == "c1_blocks" and step > 1250:
            # after the body has wrapped round the blocks the contact branch makes the trajectory chaotic: two fp32
            # evaluations that agree to 1e-7 per step part by 6.6e-4 at step 1 500 (round-1 oracle vs reference);
            # the packed-pair build of the element kernel (other rounding, 1e-5 at step 1 250) reaches 7.2e-3
            tol = 2e-2
        report["pos@%d" % step] = rel(out["pos"][k], fx["s_pos"][k])
        assert report["pos@%d" % step] <= tol, (name, step, report)
    if mode == "implicit":
        assert rel(out["vel_g"][-1], fx["s_vel_g"][-1]) <= 1e-5
    print("\n%s: %s" % (name, " ".join("%s=%.1e" % kv for kv in report.items())))


@pytest.mark.parametrize("name", ["c1_stretch_analytic", "k4_analytic", "k4_walls", "k4_block"])
def test_fused_step_loop_matches_reference_trajectory(name):
    """femb_explicit_step(n) -- the graph-captured multi-step call the bench times -- against the reference's
    neo_hookean_1_grad + kinematic trajectory."""
    from femb200 import _capi as c
    fx = rf.load(name)
    oc, mu, lam = rf.material(fx)
    h = c.Handle(fx["raw_verts"], fx["elems"], oc["center"], oc["rho"], mu, lam, oc["damping"], tile_elems=64)
    bl = rf.blocks_of(fx)
    h.blocks_set([b[0] for b in bl], [b[1] for b in bl])
    h.set(c.F_POS, fx["pos_init"])
    h.set(c.F_VEL, fx["vel_init"])
    done = 0
    for k, step in enumerate(int(s) for s in fx["rec"]):
        h.explicit_step(fx["config"]["delta_time"], fx["config"]["g_dir"], step - done)
        done = step
        assert rel(h.get(c.F_POS), fx["s_pos"][k]) <= POS_TOL
        assert rel(h.get(c.F_VEL), fx["s_vel"][k]) <= 1e-3
    h.close()
