"""The C/OpenMP restatement (bench CPU baseline, large-size checker) against the numpy oracle."""
import numpy as np
import pytest

from helpers import MAT2, MAT3, cube_case, rel, square_case
from oracle import cref
from oracle import fem_oracle as o


@pytest.mark.parametrize("maker,mat", [(cube_case, MAT3), (square_case, MAT2)])
def test_c_port_matches_numpy_oracle(maker, mat):
    verts, el, center, pos0, pos = maker()
    mu, lam = o.lame(mat["E"], mat["nu"])
    ref, vol, mass = o.elements_init(pos0, el, mat["rho"])
    want = np.zeros_like(pos)
    o.neo_hookean_1_grad(pos, el, ref, vol, mu, lam, want)
    got = cref.explicit_grad(pos, el, ref, vol, mu, lam, np.zeros_like(pos))
    assert rel(got, want) < 2e-6
    dim = pos.shape[1]
    blocks = [([0.35, 0.45, 0.35][:dim], 0.08)]
    p1, v1, f1 = pos.copy(), np.zeros_like(pos), np.zeros_like(pos)
    for _ in range(10):
        o.neo_hookean_1_grad(p1, el, ref, vol, mu, lam, f1)
        o.kinematic(p1, v1, f1, mass, 1e-4, [0, -1, 0][:dim], mat["damping"], blocks)
    p2, v2, f2 = pos.copy(), np.zeros_like(pos), np.zeros_like(pos)
    cref.explicit_steps(p2, v2, f2, mass, el, ref, vol, mu, lam, 1e-4, [0, -1, 0][:dim], mat["damping"], blocks, 10)
    assert rel(p2, p1) < 1e-6 and rel(v2, v1) < 1e-4
    H = o.element_hessian_block(pos, el, ref, vol, mu, lam)
    Hc = cref.hessian_blocks(pos, el, ref, vol, mu, lam)
    assert rel(Hc, H) < 1e-5
    d = np.random.default_rng(2).standard_normal(pos.shape).astype(np.float32)
    for tr in (False, True):
        want = (o.apply_AT if tr else o.apply_A)(H, el, mass, 5e-4, d)
        assert rel(cref.apply(H, el, mass, 5e-4, d, tr), want) < 1e-5
    b = o.compute_linear_system_vector_b(pos, d * 0.1, el, ref, vol, mass, mu, lam, 1e-4)
    for pc in (0, 1):
        x1, it1, _ = o.cg_matrix_free(H, el, mass, 1e-4, b, pc)
        x2, it2, _ = cref.cg(H, el, mass, 1e-4, b, pc)
        assert abs(it1 - it2) <= 1 and rel(x2, x1) < 1e-3
    assert cref.threads() >= 1
