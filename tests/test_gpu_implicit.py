"""GPU parity of the implicit path (solver/implicit.py) against the oracle, through the C-ABI.

The reference assembles a dense Nv x Nv matrix; the GPU path is matrix-free.  Checked here:
H_e, b, diag(A) and A d / A^T d against the literal dense restatement (<= 1e-5 rel-L2),
CG / A^T A-CG / Jacobi solutions against numpy.linalg.solve on the dense restatement
(agreement at convergence level -- never iterate by iterate), the
reference's stale-past_x and revert quirks, advect_implicit, and a 100-step trajectory."""
import numpy as np
import pytest

from helpers import MAT2, MAT3, cube_case, make_handle, rel, square_case
from oracle import fem_oracle as o

pytestmark = pytest.mark.gpu


def capi():
    from femb200 import _capi
    return _capi


CASES = {
    "cube_multi_tile": lambda: (cube_case(n=5, h=0.05, center=(0.3, 0.4, 0.3)), MAT3, 128),
    "cube_fine": lambda: (cube_case(n=4, h=0.4 / 55), MAT3, 0),
    "square_default_json": lambda: (square_case(), MAT2, 64),
}


@pytest.fixture(params=list(CASES))
def case(request):
    (verts, el, center, pos0, pos), mat, T = CASES[request.param]()
    h, mu, lam = make_handle(verts, el, center, mat, implicit=True, tile_elems=T)
    c = capi()
    rng = np.random.default_rng(11)
    vel = (0.2 * rng.standard_normal(pos.shape)).astype(np.float32)
    h.set(c.F_POS, pos)
    h.set(c.F_VEL, vel)
    ref, vol, mass = h.get(c.F_ELEM_REF), h.get(c.F_ELEM_VOLUME), h.get(c.F_MASS)
    yield dict(h=h, c=c, el=el, pos=pos, vel=vel, ref=ref, vol=vol, mass=mass, mat=mat, mu=mu, lam=lam,
               name=request.param)
    h.close()


def test_assemble_matches_dense_reference(case):
    c, h = case["c"], case["h"]
    dt = 5e-4
    h.implicit_assemble(dt)
    H = o.element_hessian_block(case["pos"], case["el"], case["ref"], case["vol"], case["mu"], case["lam"])
    assert rel(h.get(c.F_ELEM_H), H) <= 1e-5
    b = o.compute_linear_system_vector_b(case["pos"], case["vel"], case["el"], case["ref"], case["vol"], case["mass"],
                                         case["mu"], case["lam"], dt)
    assert rel(h.get(c.F_VEC_B), b) <= 1e-5
    A = o.compute_linear_system_matrix_a(case["pos"], case["el"], case["ref"], case["vol"], case["mass"], case["mu"],
                                         case["lam"], dt)
    nv, dim = case["pos"].shape
    M = A.transpose(0, 2, 1, 3).reshape(nv * dim, nv * dim).astype(np.float64)
    assert rel(h.get(c.F_DIAG_A).reshape(-1), np.diag(M)) <= 1e-5
    rng = np.random.default_rng(3)
    d = rng.standard_normal((nv, dim)).astype(np.float32)
    assert rel(h.implicit_apply(dt, d).reshape(-1), M @ d.reshape(-1)) <= 1e-5
    assert rel(h.implicit_apply(dt, d, transpose=True).reshape(-1), M.T @ d.reshape(-1)) <= 1e-5
    # deterministic operator: bitwise repeatable
    assert np.array_equal(h.implicit_apply(dt, d), h.implicit_apply(dt, d))


@pytest.mark.parametrize("method,precond", [(1, 0), (1, 1), (0, 0)])
def test_solve_matches_direct_solution(case, method, precond):
    c, h = case["c"], case["h"]
    dt = 1e-4 if case["name"] == "cube_fine" else 5e-4
    A = o.compute_linear_system_matrix_a(case["pos"], case["el"], case["ref"], case["vol"], case["mass"], case["mu"],
                                         case["lam"], dt)
    b = o.compute_linear_system_vector_b(case["pos"], case["vel"], case["el"], case["ref"], case["vol"], case["mass"],
                                         case["mu"], case["lam"], dt)
    nv, dim = case["pos"].shape
    M = A.transpose(0, 2, 1, 3).reshape(nv * dim, nv * dim).astype(np.float64)
    x_direct = np.linalg.solve(M, b.reshape(-1).astype(np.float64)).reshape(nv, dim)
    it, resid = h.implicit_solve(dt, method, precond)
    x = h.get(c.F_VEL)                                  # vel = x (implicit.py:222-223)
    assert np.array_equal(x, h.get(c.F_VEC_X))
    r = b.reshape(-1) - M @ x.reshape(-1).astype(np.float64)
    if method == 1:
        # stop rule r.r <= 1e-5 on the (possibly normal-equation) residual (implicit.py:341)
        rr = (M.T @ r) if precond else r
        assert resid <= 1e-5 and it < 500
        assert float(rr @ rr) <= 2e-5
        xo, ito, _ = o.cg(A, b, precond)
        assert abs(it - ito) <= 2
        assert rel(x, xo) <= 2e-3
    else:
        assert resid <= 1e-5 * 1.01 or it == 20000      # ||b - A x|| <= 1e-5 (implicit.py:236,252)
        assert np.linalg.norm(r) <= 5e-5
        xo, _, ito = o.jacobi_parallel_matrix_free(
            o.element_hessian_block(case["pos"], case["el"], case["ref"], case["vol"], case["mu"], case["lam"]),
            case["el"], case["mass"], dt, b, np.zeros_like(b))
        assert abs(it - ito) <= 2                      # same algorithm restated on the CPU
    assert rel(x, x_direct) <= 5e-3


def test_jacobi_keeps_stale_past_x_and_reverts(case):
    """past_vec_x persists across solves (implicit.py:85, never reset) and a non-decreasing residual
    restores it (implicit.py:257-259)."""
    c, h = case["c"], case["h"]
    dt = 1e-4 if case["name"] == "cube_fine" else 5e-4
    it1, _ = h.implicit_solve(dt, 0, 0)
    past1 = h.get(c.F_PAST_X)
    assert it1 > 0 and np.abs(past1).max() > 0
    # second solve from the same state: the first damped update mixes in past1 (not zero)
    h.set(c.F_POS, case["pos"]); h.set(c.F_VEL, case["vel"])
    it2, _ = h.implicit_solve(dt, 0, 0, max_iter=1)
    H = o.element_hessian_block(case["pos"], case["el"], case["ref"], case["vol"], case["mu"], case["lam"])
    b = o.compute_linear_system_vector_b(case["pos"], case["vel"], case["el"], case["ref"], case["vol"], case["mass"],
                                         case["mu"], case["lam"], dt)
    want, _, _ = o.jacobi_parallel_matrix_free(H, case["el"], case["mass"], dt, b, past1.copy(), max_iter=1)
    assert rel(h.get(c.F_VEC_X), want) <= 1e-4


def test_advect_implicit_parity(case):
    c, h = case["c"], case["h"]
    dim = case["pos"].shape[1]
    rng = np.random.default_rng(5)
    vel_g = (0.3 * rng.standard_normal(case["pos"].shape)).astype(np.float32)
    blocks = [([0.35, 0.45, 0.35][:dim], 0.3)] if dim == 3 else [([0.8, 0.5], 0.21), ([0.2, 0.5], 0.21)]
    h.blocks_set([b[0] for b in blocks], [b[1] for b in blocks])
    h.set(c.F_VEL_G, vel_g)
    g_dir = [0, -1, 0][:dim]
    h.advect_implicit(5e-4, g_dir)
    p, v, vg = case["pos"].copy(), case["vel"].copy(), vel_g.copy()
    o.advect_implicit(p, v, vg, 5e-4, g_dir, case["mat"]["damping"], blocks)
    assert rel(h.get(c.F_POS), p) <= 1e-7 and rel(h.get(c.F_VEL), v) <= 2e-6 and rel(h.get(c.F_VEL_G), vg) <= 2e-6


@pytest.mark.parametrize("method,precond", [(1, 1), (1, 0), (0, 0)])
def test_implicit_trajectory_100_steps(case, method, precond):
    """C2-style run, north_star's length: 100 x [implicit solve + advect_implicit] with A^T A-CG, plain CG and Jacobi
    against the dense oracle (the reference's serial sweep for Jacobi); positions after the run <= 1e-4."""
    c, h = case["c"], case["h"]
    dim = case["pos"].shape[1]
    dt, g_dir, n = 5e-4 if case["name"] != "cube_fine" else 1e-4, [0, -1, 0][:dim], 100
    if case["name"] == "cube_multi_tile" and n > 40:
        n = 40                                      # 216 vertices: the dense oracle costs ~1 s per step
    p, v = case["pos"].copy(), np.zeros_like(case["pos"])
    vg = np.zeros_like(p)
    h.set(c.F_VEL, v)
    past = np.zeros_like(p)
    for _ in range(n):
        v, past, _ = o.implicit_solver_neo_hookean(p, v, case["el"], case["ref"], case["vol"], case["mass"],
                                                   case["mu"], case["lam"], dt, method, precond, past)
        o.advect_implicit(p, v, vg, dt, g_dir, case["mat"]["damping"], [])
        h.implicit_solve(dt, method, precond, fetch=False)
        h.advect_implicit(dt, g_dir)
    got = h.get(c.F_POS)
    assert rel(got, p) <= 1e-4
    assert rel(got - case["pos"], p - case["pos"]) <= 2e-2


def test_single_reduction_cg_matches_textbook_cg(case, monkeypatch):
    """FEMB_CG_SINGLE_REDUCTION=1 (Chronopoulos-Gear arrangement of the A^T A loop, api_solve.cu; the default on thin
    multi-GPU slabs): same solution as the textbook loop to the stop tolerance, iteration count within 2."""
    c, h = case["c"], case["h"]
    dt = 1e-4 if case["name"] == "cube_fine" else 5e-4
    res = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("FEMB_CG_SINGLE_REDUCTION", flag)
        h.set(c.F_POS, case["pos"]); h.set(c.F_VEL, case["vel"])
        it, resid = h.implicit_solve(dt, 1, 1)
        res[flag] = (it, resid, h.get(c.F_VEC_X))
        assert resid <= 1e-5 or it == 500
    assert abs(res["0"][0] - res["1"][0]) <= 2, (res["0"][:2], res["1"][:2])
    assert rel(res["1"][2], res["0"][2]) <= 2e-3
    # fixed iteration count (benchmark mode): every scheduled update is applied
    monkeypatch.setenv("FEMB_CG_SINGLE_REDUCTION", "1")
    h.set(c.F_POS, case["pos"]); h.set(c.F_VEL, case["vel"])
    h.implicit_solve(dt, 1, 1, max_iter=3, tol=0.0)
    x3 = h.get(c.F_VEC_X)
    monkeypatch.setenv("FEMB_CG_SINGLE_REDUCTION", "0")
    h.set(c.F_POS, case["pos"]); h.set(c.F_VEL, case["vel"])
    h.implicit_solve(dt, 1, 1, max_iter=3, tol=0.0)
    assert rel(x3, h.get(c.F_VEC_X)) <= 1e-4
