"""GPU side of the reference-order sweep (femb_implicit_solve method 2; csrc/kernels_sweep.cu): the launches, the
device-side loop control and the CUDA-graph replay of api_solve.cu, through the C ABI and through the drop-in modules
(config key `jacobi_order: "reference"`), against the REFERENCE's own Jacobi runs (tests/golden/ref_*jacobi*.npz),
oracle.jacobi_iter_field iterate by iterate, and -- on a mesh too large for the dense oracle -- the host build of the
same row code (tests/native/sweep_host.cpp, itself checked against the reference in tests/test_sweep_host.py).

These tests were written after the round's GPU budget had been spent: the arithmetic, tables and schedule they
exercise are verified on the CPU (tests/test_sweep_host.py), the launch plumbing has NOT yet run on hardware.  They
are therefore marked xfail(strict=False): XPASS on the driver's box is the evidence that the plumbing works, and a
failure here cannot mask the rest of the GPU suite (the file also sorts last).  FEMB_SWEEP_STRICT=1 makes them plain
tests."""
import os

import numpy as np
import pytest

import ref_fixtures as rf
from helpers import MAT3, cube_case, make_handle, rel
from oracle import fem_oracle as o
from test_sweep_host import lib, solve as host_solve   # noqa: F401  (lib is a fixture)

pytestmark = [pytest.mark.gpu]
if os.environ.get("FEMB_SWEEP_STRICT") != "1":
    pytestmark.append(pytest.mark.xfail(strict=False, reason="method 2 launch plumbing not yet run on hardware "
                                                              "(GPU budget of the round was spent); CPU-verified core"))

SWEEP = 2


def capi():
    from femb200 import _capi
    return _capi


def _handle(fx):
    c = capi()
    oc, mu, lam = rf.material(fx)
    h = c.Handle(fx["raw_verts"].astype(np.float32), fx["elems"], np.asarray(oc["center"], np.float32), oc["rho"], mu, lam,
                 oc["damping"], implicit=True, tile_elems=64)
    h.set(c.F_POS, fx["pos_init"])
    h.set(c.F_VEL, fx["vel_init"])
    return c, h


def _close(h):
    try:
        h.close()
    except Exception:
        pass


def _dense(fx):
    oc, mu, lam = rf.material(fx)
    dt = fx["config"]["delta_time"]
    A = o.compute_linear_system_matrix_a(fx["pos_init"], fx["elems"], fx["ref"], fx["volume"], fx["mass"], mu, lam, dt)
    b = o.compute_linear_system_vector_b(fx["pos_init"], fx["vel_init"], fx["elems"], fx["ref"], fx["volume"], fx["mass"],
                                         mu, lam, dt)
    return A, b, dt


@pytest.mark.parametrize("name", ["c1_jacobi", "s5_jacobi", "s5_jacobi_stiff", "k3_jacobi", "spot100_jacobi"])
def test_first_solve_matches_the_reference_run(name):
    fx = rf.load(name)
    c, h = _handle(fx)
    try:
        A, b, dt = _dense(fx)
        tr = []
        xo, po, ito = o.jacobi_iter_field(A, b, np.zeros_like(b), trace=tr)
        assert np.array_equal(xo, fx["s_x"][0])                    # the oracle IS the reference here (test_ref_golden)
        it, resid = h.implicit_solve(dt, SWEEP, 0)
        # the device takes ||b - A x|| from the tile pass (another summation order): where an error sits within a few
        # rounding floors of 1e-5 the stop test may fall one sweep earlier / later -- compare like with like
        assert abs(it - ito) <= 1, (it, ito)
        if it > len(tr):                                            # one sweep more than the oracle: extend its trace
            xn = o.jacobi_iter_field_once(A, b, tr[-1][1].copy(), tr[-1][1].copy())
            tr.append((float(o.compute_error(A, b, xn)), xn))
        want = tr[it - 1][1]
        assert rel(h.get(c.F_VEC_X), want) <= 1e-5
        assert rel(h.get(c.F_PAST_X), want) <= 1e-5
        assert rel(h.get(c.F_VEL), want) <= 1e-5                    # vel = x (implicit.py:222-223)
        if it == ito:
            assert rel(h.get(c.F_VEC_X), fx["s_x"][0]) <= 1e-5
            assert rel(h.get(c.F_PAST_X), fx["s_past_x"][0]) <= 1e-5
        assert abs(resid - tr[it - 1][0]) <= 1e-3 * tr[it - 1][0] + 2e-7 * np.linalg.norm(b)
    finally:
        _close(h)


@pytest.mark.parametrize("name", ["s5_jacobi_stiff", "k3_jacobi"])
def test_iterate_by_iterate(name):
    """tol == 0 runs exactly k sweeps: the k-th iterate of the device equals the k-th iterate of the serial sweep."""
    fx = rf.load(name)
    A, b, dt = _dense(fx)
    tr = []
    o.jacobi_iter_field(A, b, np.zeros_like(b), trace=tr)
    for k in (1, 2, 3, 5, len(tr)):
        c, h = _handle(fx)
        try:
            it, _ = h.implicit_solve(dt, SWEEP, 0, max_iter=k, tol=0.0)
            assert it == k
            assert rel(h.get(c.F_VEC_X), tr[k - 1][1]) <= 1e-5, k
        finally:
            _close(h)


def test_revert_paths():
    """implicit.py:257-259 on the device: a worse sweep is undone; a worse FIRST sweep brings back the stale vector."""
    verts, el, center, pos0, pos = cube_case(n=3, jitter=0.1, seed=3)
    h, mu, lam = make_handle(verts, el, center, MAT3, implicit=True, tile_elems=64)
    c = capi()
    try:
        dt = 4e-3
        vel = (0.1 * np.random.default_rng(5).standard_normal(pos.shape)).astype(np.float32)
        stale = (0.01 * np.random.default_rng(6).standard_normal(pos.shape)).astype(np.float32)
        ref, vol, mass = h.get(c.F_ELEM_REF), h.get(c.F_ELEM_VOLUME), h.get(c.F_MASS)
        A = o.compute_linear_system_matrix_a(pos, el, ref, vol, mass, mu, lam, dt)
        b = o.compute_linear_system_vector_b(pos, vel, el, ref, vol, mass, mu, lam, dt)
        for past, want_it in ((stale, 7), ((1e7 * stale).astype(np.float32), 1)):
            xo, po, ito = o.jacobi_iter_field(A, b, past)
            assert ito == want_it
            h.set(c.F_POS, pos)
            h.set(c.F_VEL, vel)
            h.set(c.F_PAST_X, past)
            it, _ = h.implicit_solve(dt, SWEEP, 0)
            assert it == ito
            x, px = h.get(c.F_VEC_X), h.get(c.F_PAST_X)
            assert np.array_equal(x, px)                            # the answer is what past_vec_x holds
            assert rel(x, xo) <= 1e-5
            if want_it == 1:
                assert np.array_equal(x, past)                      # the stale vector itself, bit for bit
    finally:
        _close(h)


def test_large_mesh_matches_the_host_build_of_the_same_rows(lib):   # noqa: F811
    """30^3 cells: 162 000 tets in ~220 tiles, levels of up to ~700 rows -> grid launches, chained small levels at
    both ends, multi-tile H planes.  No dense oracle at this size: the host build of the same row code is the checker."""
    verts, el, center, pos0, pos = cube_case(dims=(30, 30, 30), jitter=0.1, seed=4)
    h, mu, lam = make_handle(verts, el, center, MAT3, implicit=True)
    c = capi()
    try:
        dt = 1e-4
        vel = (0.1 * np.random.default_rng(8).standard_normal(pos.shape)).astype(np.float32)
        h.set(c.F_POS, pos)
        h.set(c.F_VEL, vel)
        h.implicit_assemble(dt)
        H, mass, b = h.get(c.F_ELEM_H), h.get(c.F_MASS), h.get(c.F_VEC_B)
        past = np.zeros_like(b)
        xh, ph, ith, errh, trace, _ = host_solve(lib, verts, el, H, mass, b, past, dt, max_trace=4)
        assert 2 <= ith <= 200
        for k in (1, 3):                                             # fixed sweep counts first: isolates the row kernels
            h.set(c.F_VEL, vel)
            h.set(c.F_PAST_X, past)
            it, _ = h.implicit_solve(dt, SWEEP, 0, max_iter=k, tol=0.0)
            assert it == k and rel(h.get(c.F_VEC_X), trace[k - 1]) <= 1e-5, k
        h.set(c.F_VEL, vel)
        h.set(c.F_PAST_X, past)
        it, resid = h.implicit_solve(dt, SWEEP, 0)
        assert abs(it - ith) <= 1                                    # the stop test sits at the rounding floor
        assert rel(h.get(c.F_VEC_X), xh) <= 1e-5
        assert resid <= 1e-5 or it == 20000
    finally:
        _close(h)


def test_drop_in_trajectory_with_reference_order(monkeypatch):
    """main.py's loop (implicit_solver_neo_hookean -> advect_implicit) with jacobi_order = reference over the whole
    recorded trajectory of the reference: vec_x of every recorded step, positions at the end (north_star: <= 1e-4)."""
    import constants
    import utils
    fx = rf.load("s5_jacobi")
    cfg = dict(fx["config"], jacobi_order="reference")
    monkeypatch.delenv("FEMB_JACOBI_ORDER", raising=False)
    utils.sys_init(cfg)
    import circle_blocks
    import object as fobject
    from solver import implicit
    obj = fobject.Object(dict(cfg["objects"][0], verbose=False))          # object.py:95-126 mesh built by the product
    blocks = circle_blocks.CircleBlocks(cfg["blocks"])
    try:
        obj.particles.pos.from_numpy(fx["pos_init"])
        obj.particles.vel.from_numpy(fx["vel_init"])
        rec = [int(s) for s in fx["rec"]]
        k = 0
        for step in range(1, int(fx["steps"]) + 1):
            implicit.implicit_solver_neo_hookean(obj)
            if step in rec:
                assert rel(obj.vec_x.to_numpy(), fx["s_x"][k]) <= (1e-5 if k == 0 else 1e-3), step
                k += 1
            implicit.advect_implicit(obj, blocks)
        assert rel(obj.particles.pos.to_numpy(), fx["s_pos"][-1]) <= 1e-5
        assert constants.jacobi_order == "reference"
    finally:
        utils.sys_init(fx["config"])
