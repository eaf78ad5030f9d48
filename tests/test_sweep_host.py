"""The reference-order sweep of the library (implicit method 2, csrc/sweep_core.h + sweep.cpp) checked WITHOUT a GPU:
tests/native/sweep_host.cpp compiles the very functions the CUDA kernels call (sweep_row, sweep_decide) together with
the real tiling and table builders and runs the solve level by level on the host.  Checked here against

* the REFERENCE's own runs (tests/golden/ref_*jacobi*.npz, produced by the unmodified solver/implicit.py): vec_x,
  past_vec_x after every recorded solve of the trajectory -- the simultaneous Jacobi of method 0 cannot pass these
  iterate for iterate, the sweep must;
* oracle.jacobi_iter_field iterate by iterate (error after every sweep, iteration count, revert path);
* the schedule itself: no star edge inside a level, lower-numbered neighbours strictly earlier.
What this cannot cover -- the kernel launches and the CUDA-graph loop -- is tests/test_zz_gpu_sweep.py."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import ref_fixtures as rf
from helpers import MAT3, cube_case, rel
from oracle import fem_oracle as o

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "finite-element-method_b200", "csrc")


def _cuda_include():
    for d in ("/usr/local/cuda/include", os.path.join(os.environ.get("CUDA_HOME", "/nonexistent"), "include")):
        if os.path.exists(os.path.join(d, "vector_types.h")):
            return d
    pytest.skip("CUDA headers (vector_types.h) not found")


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("sweep_host") / "libsweep_host.so")
    src = [os.path.join(ROOT, "tests", "native", "sweep_host.cpp"), os.path.join(CSRC, "tiling.cpp"),
           os.path.join(CSRC, "sweep.cpp")]
    subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-fopenmp", "-shared", "-I" + CSRC, "-I" + _cuda_include()] + src +
                   ["-o", out], check=True)
    L = C.CDLL(out)
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    L.sweep_host_levels.restype = C.c_int
    L.sweep_host_levels.argtypes = [C.c_int, C.c_int, C.c_int, fp, ip, C.c_int, ip, ip, ip]
    L.sweep_host_solve.restype = C.c_int
    L.sweep_host_solve.argtypes = [C.c_int, C.c_int, C.c_int, fp, ip, C.c_int, fp, fp, fp, fp, C.c_float, C.c_float,
                                   C.c_int, C.c_int, C.c_int, fp, fp, ip, fp, fp, fp, C.c_int]
    return L


def _f(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _i(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def levels(L, verts, el, tile_elems=0):
    verts = np.ascontiguousarray(verts, np.float32)
    el = np.ascontiguousarray(el, np.int32)
    lv = np.full(verts.shape[0], -1, np.int32)
    ng, mc = C.c_int32(), C.c_int32()
    n = L.sweep_host_levels(verts.shape[1], verts.shape[0], el.shape[0], _f(verts), _i(el), tile_elems, _i(lv),
                            C.byref(ng), C.byref(mc))
    assert n > 0
    return lv, n, ng.value, mc.value


def solve(L, verts, el, H, mass, b, past, dt, tol=-1.0, max_iter=0, fixed=0, order=0, tile_elems=0, max_trace=64):
    verts = np.ascontiguousarray(verts, np.float32)
    el = np.ascontiguousarray(el, np.int32)
    nv, dim = b.shape
    H, mass, b, past = (np.ascontiguousarray(a, np.float32) for a in (H, mass, b, past))
    x, po = np.zeros_like(b), np.zeros_like(b)
    tr, te = np.zeros((max_trace, nv, dim), np.float32), np.full(max_trace, np.nan, np.float32)
    it, err = C.c_int32(), C.c_float()
    rc = L.sweep_host_solve(dim, nv, el.shape[0], _f(verts), _i(el), tile_elems, _f(H), _f(mass), _f(b), _f(past), dt,
                            tol, max_iter, fixed, order, _f(x), _f(po), C.byref(it), C.byref(err), _f(tr), _f(te),
                            max_trace)
    assert rc > 0
    return x, po, it.value, err.value, tr, te


def _check_schedule(lv, el):
    el = np.asarray(el)
    for j in range(1, el.shape[1]):                      # star edges: corner 0 <-> corner j
        a, c = el[:, 0], el[:, j]
        assert np.all(lv[a] != lv[c])
        lo, hi = np.minimum(a, c), np.maximum(a, c)      # the reference's row order is the vertex id
        assert np.all(lv[lo] < lv[hi])
    # minimal: every row of level l > 0 has a lower-numbered star neighbour in level l - 1
    best = np.zeros(lv.shape[0], np.int64) - 1
    for j in range(1, el.shape[1]):
        for a, c in ((el[:, 0], el[:, j]), (el[:, j], el[:, 0])):
            m = c < a
            np.maximum.at(best, a[m], lv[c[m]])
    assert np.array_equal(best + 1, lv)


@pytest.mark.parametrize("tile_elems", [0, 64])
def test_levels_follow_the_reference_row_order(lib, tile_elems):
    verts, el = o.construct_2d_mesh(0.2, 10)             # C1's mesh
    lv, n, ng, mc = levels(lib, verts, el, tile_elems)
    _check_schedule(lv, el)
    assert ng == 1 and mc <= 256                         # 121 rows: one CTA chains every level
    verts, el, *_ = cube_case(n=6)
    lv, n, ng, mc = levels(lib, verts, el, tile_elems)
    _check_schedule(lv, el)
    i, j, k = np.meshgrid(*(np.arange(7),) * 3, indexing="ij")
    assert np.all(lv <= (i + j + k).reshape(-1)) and 7 <= n <= 19    # Kuhn box: never later than the hyperplane i + j + k
    fx = rf.load("spot100_jacobi")                       # unstructured (the reference's asset)
    lv, n, ng, mc = levels(lib, fx["raw_verts"], fx["elems"], tile_elems)
    _check_schedule(lv, fx["elems"])


def test_large_levels_get_their_own_launch(lib):
    verts, el, *_ = cube_case(dims=(30, 30, 30))
    lv, n, ng, mc = levels(lib, verts, el)
    assert 31 <= n <= 91 and mc <= 256
    sizes = np.bincount(lv)
    big = int((sizes > 256).sum())
    assert big > 0 and ng == big + 2                     # one chain of small levels at either end + one launch per big level


def _system(fx, pos, vel):
    oc, mu, lam = rf.material(fx)
    dt = fx["config"]["delta_time"]
    H = o.element_hessian_block(pos, fx["elems"], fx["ref"], fx["volume"], mu, lam)
    b = o.compute_linear_system_vector_b(pos, vel, fx["elems"], fx["ref"], fx["volume"], fx["mass"], mu, lam, dt)
    return H, b, dt


@pytest.mark.parametrize("name", ["c1_jacobi", "s5_jacobi", "s5_jacobi_stiff", "k3_jacobi", "spot100_jacobi"])
def test_host_sweep_reproduces_the_reference_trajectory(lib, name):
    """The whole recorded trajectory of the reference (assemble -> its serial sweep -> advect), the solve done by the
    library's sweep code on the host: vec_x and past_vec_x of every recorded step."""
    fx = rf.load(name)
    cfg = fx["config"]
    oc = cfg["objects"][0]
    pos, vel = fx["pos_init"].copy(), fx["vel_init"].copy()
    vel_g, past = np.zeros_like(pos), np.zeros_like(pos)
    rec = [int(s) for s in fx["rec"]]
    k = 0
    for step in range(1, int(fx["steps"]) + 1):
        H, b, dt = _system(fx, pos, vel)
        x, past, it, err, _, _ = solve(lib, fx["raw_verts"], fx["elems"], H, fx["mass"], b, past, dt)
        if step in rec:
            # the first recorded solve starts from the reference's exact state; later ones carry the rounding
            # differences of all earlier steps (the bar for those is north_star's trajectory tolerance)
            tol = 2e-6 if k == 0 else 1e-3      # (velocities: the same 1e-3 as tests/test_gpu_ref_golden.py)
            assert rel(b, fx["s_b"][k]) <= tol / 2, (name, step)
            assert rel(x, fx["s_x"][k]) <= tol, (name, step, it)
            assert rel(past, fx["s_past_x"][k]) <= tol, (name, step, it)
            k += 1
        vel = x.copy()
        o.advect_implicit(pos, vel, vel_g, dt, cfg["g_dir"], oc["damping"], rf.blocks_of(fx))
    assert k == len(rec)
    assert rel(pos, fx["s_pos"][-1]) <= 1e-5             # north_star: positions <= 1e-4


@pytest.mark.parametrize("name", ["s5_jacobi_stiff", "k3_jacobi"])
def test_host_sweep_iterate_by_iterate(lib, name):
    fx = rf.load(name)
    A, b = fx["s_A"][0], fx["s_b"][0]
    H, b2, dt = _system(fx, fx["pos_init"], fx["vel_init"])
    assert np.array_equal(b, b2)
    tr = []
    xo, po, ito = o.jacobi_iter_field(A, b, np.zeros_like(b), trace=tr)
    for order in (0, 1):                                  # rows of a level forwards / backwards: same bits
        x, past, it, err, trace, terr = solve(lib, fx["raw_verts"], fx["elems"], H, fx["mass"], b, np.zeros_like(b), dt,
                                              order=order)
        assert it == ito == len(tr)
        for i, (e, xi) in enumerate(tr):
            assert rel(trace[i], xi) <= 2e-6, i
            assert abs(terr[i] - e) <= 1e-4 * e + 1e-7 * np.linalg.norm(b), i   # rounding floor of b - A x
        assert rel(x, xo) <= 2e-6 and rel(past, po) <= 2e-6
        if order == 0:
            first = (x, trace.copy())
        else:
            assert np.array_equal(first[0], x) and np.array_equal(first[1], trace)


def test_host_sweep_revert_and_fixed_mode(lib):
    """The reference's break rule (257-259): a sweep that makes the residual worse is undone (vec_x = past_vec_x); when
    that is the very first sweep, what comes back is the STALE past_vec_x of the previous time step."""
    verts, el, center, pos0, pos = cube_case(n=3, jitter=0.1, seed=3)
    mu, lam = o.lame(MAT3["E"], MAT3["nu"])
    ref, vol, mass = o.elements_init(pos0, el, MAT3["rho"])
    dt = 4e-3                                            # outside the sweep's convergence domain
    vel = (0.1 * np.random.default_rng(5).standard_normal(pos.shape)).astype(np.float32)
    A = o.compute_linear_system_matrix_a(pos, el, ref, vol, mass, mu, lam, dt)
    b = o.compute_linear_system_vector_b(pos, vel, el, ref, vol, mass, mu, lam, dt)
    H = o.element_hessian_block(pos, el, ref, vol, mu, lam)
    stale = (0.01 * np.random.default_rng(6).standard_normal(pos.shape)).astype(np.float32)
    tr = []
    xo, po, ito = o.jacobi_iter_field(A, b, stale, trace=tr)
    errs = [e for e, _ in tr]
    assert ito == 7 and errs[-1] > errs[-2] and np.array_equal(xo, tr[-2][1])   # sweep 7 is undone: the answer is iterate 6
    x, past, it, err, trace, terr = solve(lib, verts, el, H, mass, b, stale, dt)
    assert it == 7 and np.array_equal(x, trace[5]) and np.array_equal(past, x)
    assert rel(x, xo) <= 1e-5 and abs(terr[6] - errs[6]) <= 1e-4 * errs[6]
    # first sweep worse than the start: back to the stale vector, which also stays in past_vec_x
    huge = (1e7 * stale).astype(np.float32)
    xo, po, ito = o.jacobi_iter_field(A, b, huge)
    assert ito == 1 and np.array_equal(xo, huge)
    x, past, it, err, trace, _ = solve(lib, verts, el, H, mass, b, huge, dt)
    assert it == 1 and np.array_equal(x, huge) and np.array_equal(past, huge)
    # fixed mode: exactly n sweeps, no revert, every iterate accepted
    x3, p3, it3, _, trace3, _ = solve(lib, verts, el, H, mass, b, stale, dt, tol=0.0, max_iter=9, fixed=1)
    assert it3 == 9 and np.array_equal(x3, trace3[8]) and np.array_equal(p3, x3)
    assert rel(trace3[6], tr[6][1]) <= 1e-5
