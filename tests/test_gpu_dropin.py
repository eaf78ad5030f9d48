"""The drop-in boundary exercised the way the reference's main.py uses it: same module paths,
names and call order (main.py:38-49 import order, :101-113 step loop), results against the oracle.

  C1  default.json as shipped (2D, auto_diff explicit): compute_energy leaves dU/dx in particles.pos.grad, U and
      phi readable; kinematic consumes pos.grad; a run long enough to reach the circle blocks (main.py:103-110)
  C1' analytic explicit (neo_hookean_1_grad -> kinematic) from a 10 % stretched state, 100 steps
  C2' implicit (CG, both `preconditioned` values; Jacobi) on a small 3D mesh: vel = solution; the debug densifier
      obj.matrix_A / matrix_AT / matrix_ATA against the literal dense restatement (implicit.py:114-194, 289-299)
"""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "finite-element-method_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

from helpers import rel  # noqa: E402
from oracle import fem_oracle as o  # noqa: E402

pytestmark = pytest.mark.gpu


def _setup(config_path, **edits):
    """main.py:38-49: read_config, sys_init, THEN import the modules that use constants."""
    import utils
    config = utils.read_config(os.path.join(PKG, config_path))
    for k, v in edits.items():
        if k.startswith("obj_"):
            config["objects"][0][k[4:]] = v
        else:
            config[k] = v
    config["objects"][0]["verbose"] = False
    utils.sys_init(config)
    from circle_blocks import CircleBlocks
    from object import Object
    obj = Object(config["objects"][0])
    blocks = CircleBlocks(config.get("blocks"))
    return config, obj, blocks


def _oracle_blocks(blocks):
    return [(b.center, b.radius) for b in blocks.blocks]


def test_c1_default_json_autodiff_loop_reaches_the_blocks():
    import constants
    import solver.kinematic as ki
    from solver.explicit_auto_diff import compute_energy

    config, obj, blocks = _setup("default.json")
    assert constants.dim == 2 and constants.auto_diff == 1 and obj.particle_cnt == 121 and obj.element_cnt == 200
    assert blocks.blocks_count == len(config["blocks"])
    el = obj.elements.vertex_indices.to_numpy()
    ref, vol = obj.elements.ref.to_numpy(), obj.elements.volume.to_numpy()
    mass = obj.particles.mass.to_numpy()
    pos0 = obj.particles.pos.to_numpy()
    mat = config["objects"][0]
    assert abs(obj.mu - mat["E"] / 2 / (1 + mat["nu"])) < 1e-9

    # one stretched state: energy, phi and dU/dx where kinematic expects it (main.py:107-108)
    stretched = pos0.copy()
    stretched[:, 0] = (stretched[:, 0] - 0.5) * np.float32(1.1) + 0.5
    obj.particles.pos.from_numpy(stretched)
    compute_energy(obj)
    phi, U = o.compute_energy(stretched, el, ref, vol, obj.mu, obj.s_lambda)
    assert abs(obj.U[None] - U) <= 1e-5 * abs(U)
    assert rel(obj.phi.to_numpy(), phi) <= 1e-5
    _, g64 = o.energy_autograd(stretched, el, ref, vol, obj.mu, obj.s_lambda, dtype=np.float64)
    assert rel(obj.particles.pos.grad.to_numpy(), g64) <= 1e-5

    # the shipped loop (main.py:103-110) from the shipped state until the body has hit a block
    obj.particles.pos.from_numpy(pos0)
    obj.particles.vel.from_numpy(np.zeros_like(pos0))
    p, v = pos0.copy(), np.zeros_like(pos0)
    dt, g_dir, damping = config["delta_time"], config["g_dir"], mat["damping"]
    ob = _oracle_blocks(blocks)
    steps, hit = 0, False
    while steps < 1500:
        for _ in range(50):
            compute_energy(obj)
            ki.kinematic(obj, blocks)
            _, grad = o.energy_autograd(p, el, ref, vol, obj.mu, obj.s_lambda)
            o.kinematic(p, v, grad, mass, dt, g_dir, damping, ob, auto_diff=1)
        steps += 50
        got = obj.particles.pos.to_numpy()
        assert np.isfinite(got).all()
        if steps <= 100:
            assert rel(got, p) <= 1e-4 and rel(got - pos0, p - pos0) <= 1e-3      # north_star: 100 steps <= 1e-4
        inside = [np.linalg.norm(got - c[None, :], axis=1).min() < r * 1.02 for c, r in ob if r > 0]
        if any(inside):
            hit = True
            break
    assert hit, "the body never reached a block in 1500 steps"
    # contact is a discontinuity: trajectories stay close but are no longer compared at 1e-4
    assert np.abs(obj.particles.pos.to_numpy() - p).max() <= 5e-3
    # nothing has entered a block deeper than one step of travel
    for c, r in ob:
        if r > 0:
            assert np.linalg.norm(obj.particles.pos.to_numpy() - c[None, :], axis=1).min() >= r - 5e-3


def test_c1_analytic_explicit_from_stretched_state():
    import solver.kinematic as ki
    from solver.explicit import neo_hookean_1_grad

    config, obj, blocks = _setup("config/default_analytic.json")
    el = obj.elements.vertex_indices.to_numpy()
    ref, vol, mass = obj.elements.ref.to_numpy(), obj.elements.volume.to_numpy(), obj.particles.mass.to_numpy()
    pos0 = obj.particles.pos.to_numpy()
    stretched = pos0.copy()
    stretched[:, 1] = (stretched[:, 1] - 0.8) * np.float32(1.1) + 0.8
    obj.particles.pos.from_numpy(stretched)
    mat = config["objects"][0]

    neo_hookean_1_grad(obj)                       # deferred ...
    want = np.zeros_like(stretched)
    o.neo_hookean_1_grad(stretched, el, ref, vol, obj.mu, obj.s_lambda, want)
    assert rel(obj.particles.force.to_numpy(), want) <= 1e-5      # ... and materialised by the read
    obj.particles.force.from_numpy(np.zeros_like(want))

    p, v, f = stretched.copy(), np.zeros_like(stretched), np.zeros_like(stretched)
    ob = _oracle_blocks(blocks)
    for _ in range(100):
        neo_hookean_1_grad(obj)
        ki.kinematic(obj, blocks)
        o.neo_hookean_1_grad(p, el, ref, vol, obj.mu, obj.s_lambda, f)
        o.kinematic(p, v, f, mass, config["delta_time"], config["g_dir"], mat["damping"], ob, auto_diff=0)
    got = obj.particles.pos.to_numpy()
    assert rel(got, p) <= 1e-4 and rel(got - stretched, p - stretched) <= 1e-3
    assert rel(obj.particles.vel.to_numpy(), v) <= 1e-3
    assert np.all(obj.particles.force.to_numpy() == 0)            # kinematic.py:44-45


@pytest.mark.parametrize("method,pre", [(1, 1), (1, 0), (0, 0)])
def test_c2_implicit_solver_and_dense_debug_matrices(method, pre):
    import constants
    from solver.implicit import advect_implicit, implicit_solver_neo_hookean

    dt = 5e-4
    config, obj, blocks = _setup("config/demo_3d_cube.json", use_explicit_method=False, implicit_method=method,
                                 preconditioned=pre, delta_time=dt, obj_subdivisions=3, obj_side_length=0.15)
    assert constants.implicit_method == method and obj.implicit
    el = obj.elements.vertex_indices.to_numpy()
    ref, vol, mass = obj.elements.ref.to_numpy(), obj.elements.volume.to_numpy(), obj.particles.mass.to_numpy()
    rng = np.random.default_rng(2)
    pos = obj.particles.pos.to_numpy()
    pos = (pos + 0.004 * rng.uniform(-1, 1, pos.shape)).astype(np.float32)
    vel = (0.2 * rng.standard_normal(pos.shape)).astype(np.float32)
    obj.particles.pos.from_numpy(pos)
    obj.particles.vel.from_numpy(vel)
    nv, dim = pos.shape

    A = o.compute_linear_system_matrix_a(pos, el, ref, vol, mass, obj.mu, obj.s_lambda, dt)   # [Nv, Nv, dim, dim]
    dense = lambda B: np.asarray(B, np.float64).transpose(0, 2, 1, 3).reshape(nv * dim, nv * dim)
    M = dense(A)
    assert rel(dense(obj.matrix_A.to_numpy()), M) <= 1e-5
    assert rel(dense(obj.matrix_AT.to_numpy()), M.T) <= 1e-5                      # implicit.py:292
    assert rel(dense(obj.matrix_ATA.to_numpy()), M.T @ M if pre == 1 else M) <= 1e-5   # implicit.py:299 / 306
    assert obj.matrix_A.shape == (nv, nv)

    b = o.compute_linear_system_vector_b(pos, vel, el, ref, vol, mass, obj.mu, obj.s_lambda, dt)
    implicit_solver_neo_hookean(obj)
    x = obj.particles.vel.to_numpy()                                                # vel = x (implicit.py:222-223)
    assert rel(obj.vec_b.to_numpy(), b) <= 1e-5 and np.array_equal(obj.vec_x.to_numpy(), x)
    x_direct = np.linalg.solve(M, b.reshape(-1).astype(np.float64)).reshape(nv, dim)
    assert rel(x, x_direct) <= 5e-3 and obj.last_solve["iters"] > 0
    p, v, vg = pos.copy(), x.copy(), np.zeros_like(pos)
    advect_implicit(obj, blocks)
    o.advect_implicit(p, v, vg, dt, config["g_dir"], config["objects"][0]["damping"], _oracle_blocks(blocks))
    assert rel(obj.particles.pos.to_numpy(), p) <= 1e-6
