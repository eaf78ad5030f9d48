"""GPU parity of the explicit path against the oracle, through the C-ABI.

Tolerances (BASELINE.json north_star): per-step forces rel-L2 <= 1e-5 on
bit-identical fp32 inputs at >= 5 % jitter; positions after 100 steps <= 1e-4.
Dm^-1 and volumes are required to be BIT-EXACT (contraction-free setup kernel)."""
import numpy as np
import pytest

from helpers import MAT2, MAT3, cube_case, make_handle, rel, square_case
from oracle import fem_oracle as o

pytestmark = pytest.mark.gpu
C = None


def capi():
    global C
    if C is None:
        from femb200 import _capi
        C = _capi
    return C


CASES = {
    "cube_multi_tile": lambda: (cube_case(n=8), MAT3, 256),
    "cube_one_tile": lambda: (cube_case(n=4), MAT3, 0),      # 384 tets < default tile capacity
    "box_ragged": lambda: (cube_case(dims=(9, 5, 7), seed=3), MAT3, 320),
    "square_default_json": lambda: (square_case(), MAT2, 0),
    "square_multi_tile": lambda: (square_case(s=24, seed=4), MAT2, 64),
}


@pytest.fixture(params=list(CASES))
def case(request):
    (verts, el, center, pos0, pos), mat, T = CASES[request.param]()
    h, mu, lam = make_handle(verts, el, center, mat, tile_elems=T)
    yield dict(h=h, verts=verts, el=el, center=center, pos0=pos0, pos=pos, mat=mat, mu=mu, lam=lam)
    h.close()


def test_setup_bit_exact_ref_volume_and_mass(case):
    c, h = capi(), case["h"]
    ref, vol, mass = o.elements_init(case["pos0"], case["el"], case["mat"]["rho"])
    assert np.array_equal(h.get(c.F_POS), case["pos0"])
    assert np.array_equal(h.get(c.F_REF_POS), case["pos0"])
    assert np.array_equal(h.get(c.F_ELEM_VOLUME), vol)
    assert np.array_equal(h.get(c.F_ELEM_REF), ref)
    got = h.get(c.F_MASS)
    assert np.abs(got / mass - 1).max() < 1e-6
    assert abs(got.sum(dtype=np.float64) / (case["mat"]["rho"] * vol.sum(dtype=np.float64)) - 1) < 1e-6
    assert np.all(h.get(c.F_FORCE) == 0) and np.all(h.get(c.F_VEL) == 0)


def test_force_parity_and_accumulation(case):
    c, h = capi(), case["h"]
    h.set(c.F_POS, case["pos"])
    ref, vol = h.get(c.F_ELEM_REF), h.get(c.F_ELEM_VOLUME)
    want = np.zeros_like(case["pos"])
    o.neo_hookean_1_grad(case["pos"], case["el"], ref, vol, case["mu"], case["lam"], want)
    h.explicit_grad()
    got = h.get(c.F_FORCE)
    assert rel(got, want) <= 1e-5, rel(got, want)
    # fp64 adjudicator on the same fp32 inputs
    w64 = np.zeros(case["pos"].shape)
    o.neo_hookean_1_grad(case["pos"].astype(np.float64), case["el"], ref.astype(np.float64), vol.astype(np.float64),
                         case["mu"], case["lam"], w64)
    assert rel(got, w64) <= 1e-5
    h.explicit_grad()                      # force += ... (explicit.py:42-49)
    assert rel(h.get(c.F_FORCE), 2 * got.astype(np.float64)) <= 1e-6
    # run-to-run determinism (no float atomics)
    h.set(c.F_FORCE, np.zeros_like(got))
    h.explicit_grad()
    assert np.array_equal(h.get(c.F_FORCE), got)


def test_rest_state_force_is_rounding_noise(case):
    c, h = capi(), case["h"]
    h.explicit_grad()
    f = h.get(c.F_FORCE)
    scale = case["mu"] * h.get(c.F_ELEM_VOLUME).max() ** (1 - 1 / case["pos"].shape[1])
    assert np.abs(f).max() < 1e-4 * scale


def test_energy_and_autodiff_gradient(case):
    """C5's oracle: analytic CUDA gradient vs reverse-mode AD of the restated energy."""
    c, h = capi(), case["h"]
    h.set(c.F_POS, case["pos"])
    ref, vol = h.get(c.F_ELEM_REF), h.get(c.F_ELEM_VOLUME)
    U = h.energy_grad(fetch=True)
    phi_want, U_want = o.compute_energy(case["pos"], case["el"], ref, vol, case["mu"], case["lam"])
    assert rel(h.get(c.F_ELEM_PHI), phi_want) <= 1e-5
    assert abs(U - U_want) <= 1e-5 * abs(U_want)
    assert abs(h.energy() - U) == 0
    U64, g64 = o.energy_autograd(case["pos"], case["el"], ref, vol, case["mu"], case["lam"], dtype=np.float64)
    g = h.get(c.F_GRAD)
    assert rel(g, g64) <= 1e-5
    assert abs(U - U64) <= 1e-5 * abs(U64)
    h.explicit_grad()
    assert np.array_equal(h.get(c.F_FORCE), g)   # same kernel math: analytic == "autodiff" field


def _blocks(dim):
    if dim == 2:
        return [([0.8, 0.5], 0.21), ([0.2, 0.5], 0.21)]   # default.json:25-34
    return [([0.5, 0.5, 0.5], 0.0), ([0.35, 0.45, 0.35], 0.08)]


def test_kinematic_step_parity_fused_equals_unfused(case):
    c, h = capi(), case["h"]
    dim = case["pos"].shape[1]
    dt = 1e-4
    g_dir = [0, -1, 0][:dim]
    blocks = _blocks(dim)
    h.blocks_set([b[0] for b in blocks], [b[1] for b in blocks])
    rng = np.random.default_rng(7)
    vel0 = (0.5 * rng.standard_normal(case["pos"].shape)).astype(np.float32)
    h.set(c.F_POS, case["pos"]); h.set(c.F_VEL, vel0)
    mass = h.get(c.F_MASS)
    h.explicit_grad()
    force = h.get(c.F_FORCE)
    h.kinematic(dt, g_dir, 0)
    p_unf, v_unf = h.get(c.F_POS), h.get(c.F_VEL)
    assert np.all(h.get(c.F_FORCE) == 0)                       # kinematic.py:44-45
    pos, vel = case["pos"].copy(), vel0.copy()
    o.kinematic(pos, vel, force.copy(), mass, dt, g_dir, case["mat"]["damping"], blocks, auto_diff=0)
    assert rel(v_unf, vel) <= 2e-6 and rel(p_unf, pos) <= 1e-7
    # fused single pass: bit-identical to the two-call sequence
    h.set(c.F_POS, case["pos"]); h.set(c.F_VEL, vel0)
    h.explicit_step(dt, g_dir, 1)
    assert np.array_equal(h.get(c.F_POS), p_unf) and np.array_equal(h.get(c.F_VEL), v_unf)
    # auto_diff flavour: reads pos.grad, leaves force alone (kinematic.py:15-17,44)
    h.set(c.F_POS, case["pos"]); h.set(c.F_VEL, vel0)
    h.energy_grad()
    h.set(c.F_FORCE, np.ones_like(vel0))
    h.kinematic(dt, g_dir, 1)
    assert np.array_equal(h.get(c.F_POS), p_unf)
    assert np.all(h.get(c.F_FORCE) == 1)


@pytest.mark.parametrize("name", ["cube_multi_tile", "square_default_json"])
def test_trajectory_100_steps(name):
    """positions after 100 steps <= 1e-4 (north_star); graph-captured fused steps."""
    c = capi()
    (verts, el, center, pos0, pos), mat, T = CASES[name]()
    dim = pos.shape[1]
    h, mu, lam = make_handle(verts, el, center, mat, tile_elems=T)
    dt = 1e-4 if dim == 3 else 5e-4
    g_dir = [0, -1, 0][:dim]
    blocks = _blocks(dim)
    h.blocks_set([b[0] for b in blocks], [b[1] for b in blocks])
    h.set(c.F_POS, pos)
    ref, vol, mass = h.get(c.F_ELEM_REF), h.get(c.F_ELEM_VOLUME), h.get(c.F_MASS)
    p, v = pos.copy(), np.zeros_like(pos)
    f = np.zeros_like(pos)
    for _ in range(100):
        o.neo_hookean_1_grad(p, el, ref, vol, mu, lam, f)
        o.kinematic(p, v, f, mass, dt, g_dir, mat["damping"], blocks, auto_diff=0)
    h.explicit_step(dt, g_dir, 50)
    h.explicit_step(dt, g_dir, 50)      # second call replays the cached CUDA graph
    got_p, got_v = h.get(c.F_POS), h.get(c.F_VEL)
    assert rel(got_p, p) <= 1e-4
    assert rel(got_p - pos, p - pos) <= 1e-3      # the displacement itself, not just the O(1) coordinates
    assert rel(got_v, v) <= 1e-3
    h.close()


def test_no_cpu_fallback_and_launch_counter():
    c = capi()
    (verts, el, center, pos0, pos), mat, T = CASES["cube_one_tile"]()
    n0 = c.launch_count()
    h, *_ = make_handle(verts, el, center, mat, tile_elems=T)
    h.explicit_step(1e-4, [0, -1, 0], 3)
    h.sync()
    assert c.launch_count() - n0 >= 3 + 2      # setup + 3 steps
    with pytest.raises(RuntimeError, match="implicit"):
        h.advect_implicit(1e-4, [0, -1, 0])
    h.close()


def test_async_field_transfers_match_blocking():
    """femb_set_async / femb_get_async (pipelined host traffic) give the same state as the blocking calls, also when
    the same staging buffers are re-used back to back without a sync in between."""
    c = capi()
    verts, el, center, pos0, pos = cube_case(n=6)
    h, mu, lam = make_handle(verts, el, center, MAT3, tile_elems=128)
    g = [0, -1, 0]
    h.set(c.F_POS, pos)
    h.explicit_step(1e-4, g, 3)
    want = h.get(c.F_POS)
    vel0 = np.zeros_like(pos)
    outs = [np.empty_like(pos), np.empty_like(pos)]
    for k in range(4):                      # no sync between rounds: the events order the staging-buffer re-use
        h.set_async(c.F_POS, pos)
        h.set_async(c.F_VEL, vel0)
        h.explicit_step(1e-4, g, 3)
        h.get_async(c.F_POS, outs[k & 1])
    h.transfers_wait()
    assert np.array_equal(outs[0], want) and np.array_equal(outs[1], want)
    with pytest.raises(ValueError):
        h.set_async(c.F_POS, pos.astype(np.float64))
    h.close()


def test_device_diagnostics_count_inverted_and_nan(case):
    """femb_diagnostics (main.py --diag-every, object.Object.count_inverted): inverted elements counted on the
    device equal the fp64 count on the host; a NaN vertex shows up in all three counters."""
    c, h, el = capi(), case["h"], case["el"]
    pos0 = h.get(c.F_POS)
    h.energy_grad()
    assert h.diagnostics() == {"inverted_elements": 0, "nan_elements": 0, "nan_vertices": 0}
    # push one vertex far through its neighbours: some of its elements turn inside out
    pos = pos0.copy()
    v = int(el[len(el) // 2][0])
    pos[v] += 0.8
    h.set(c.F_POS, pos)

    def dets(p):
        return np.linalg.det((p[el[:, 1:]] - p[el[:, :1]]).astype(np.float64))
    want = int((dets(pos) * dets(pos0) <= 0).sum())
    assert want > 0
    h.energy_grad()
    d = h.diagnostics()
    assert d["inverted_elements"] == want and d["nan_vertices"] == 0
    assert d["nan_elements"] == int(np.isnan(h.get(c.F_ELEM_PHI)).sum()) > 0      # log of a negative determinant
    pos = pos0.copy()
    pos[v, 0] = np.nan
    h.set(c.F_POS, pos)
    h.energy_grad()
    d = h.diagnostics()
    touched = int((el == v).any(axis=1).sum())
    assert d == {"inverted_elements": touched, "nan_elements": touched, "nan_vertices": 1}
    h.set(c.F_POS, pos0)
