"""Committed golden vectors (tests/golden/fem_golden.npz, produced by tests/golden/make_golden.py from the numpy
oracle -- NOT reference outputs): the oracle must keep reproducing them (CPU), and the
CUDA path must match them through the C ABI (GPU)."""
import os

import numpy as np
import pytest

from helpers import MAT2, MAT3, rel
from oracle import fem_oracle as o

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fem_golden.npz"))
CASES = {"cube_": MAT3, "square_": MAT2}


def _blocks(g, pre):
    d = g[pre + "verts"].shape[1]
    return [(tuple(row[:d]), float(row[d])) for row in g[pre + "blocks"]]


@pytest.mark.parametrize("pre", list(CASES))
def test_oracle_reproduces_golden(pre):
    mat = CASES[pre]
    mu, lam = o.lame(mat["E"], mat["nu"])
    el, pos = G[pre + "elems"], G[pre + "pos"]
    pos0 = o.particles_init(G[pre + "verts"], G[pre + "center"])
    ref, vol, mass = o.elements_init(pos0, el, mat["rho"])
    assert np.array_equal(ref, G[pre + "ref"]) and np.array_equal(vol, G[pre + "vol"])
    assert rel(mass, G[pre + "mass"]) < 1e-6
    force = np.zeros_like(pos)
    o.neo_hookean_1_grad(pos, el, ref, vol, mu, lam, force)
    assert rel(force, G[pre + "force"]) < 1e-6
    phi, U = o.compute_energy(pos, el, ref, vol, mu, lam)
    assert rel(phi, G[pre + "phi"]) < 1e-5 and abs(U - G[pre + "U"]) <= 1e-5 * abs(G[pre + "U"])
    H = o.element_hessian_block(pos, el, ref, vol, mu, lam)
    assert rel(H, G[pre + "H"]) < 1e-6
    dt = float(G[pre + "dt"])
    assert rel(o.apply_A(H, el, mass, dt, G[pre + "d"]), G[pre + "Ad"]) < 1e-6
    assert rel(o.apply_AT(H, el, mass, dt, G[pre + "d"]), G[pre + "ATd"]) < 1e-6
    b = o.compute_linear_system_vector_b(pos, G[pre + "vel"], el, ref, vol, mass, mu, lam, dt)
    assert rel(b, G[pre + "b"]) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("pre", list(CASES))
def test_cuda_path_matches_golden(pre):
    from femb200 import _capi as c

    mat = CASES[pre]
    mu, lam = o.lame(mat["E"], mat["nu"])
    dt, g_dir = float(G[pre + "dt"]), G[pre + "g_dir"]
    h = c.Handle(G[pre + "verts"], G[pre + "elems"], G[pre + "center"], mat["rho"], mu, lam, mat["damping"],
                 implicit=True, tile_elems=64)
    assert np.array_equal(h.get(c.F_ELEM_REF), G[pre + "ref"])            # setup: bit-exact
    assert np.array_equal(h.get(c.F_ELEM_VOLUME), G[pre + "vol"])
    assert rel(h.get(c.F_MASS), G[pre + "mass"]) < 1e-6
    h.set(c.F_POS, G[pre + "pos"])
    h.explicit_grad()
    assert rel(h.get(c.F_FORCE), G[pre + "force"]) <= 1e-5               # north_star force tolerance
    U = h.energy_grad(fetch=True)
    assert abs(U - G[pre + "U"]) <= 1e-5 * abs(G[pre + "U"])
    assert rel(h.get(c.F_ELEM_PHI), G[pre + "phi"]) <= 1e-5
    assert rel(h.get(c.F_GRAD), G[pre + "force"]) <= 1e-5                # autodiff-path gradient == analytic force
    bl = _blocks(G, pre)
    h.blocks_set([b[0] for b in bl], [b[1] for b in bl])
    h.kinematic(dt, g_dir, 0)
    assert rel(h.get(c.F_POS), G[pre + "kin_pos"]) <= 1e-6 and rel(h.get(c.F_VEL), G[pre + "kin_vel"]) <= 1e-5
    # implicit pieces on the same state
    h.set(c.F_POS, G[pre + "pos"])
    h.set(c.F_VEL, G[pre + "vel"])
    h.implicit_assemble(dt)
    assert rel(h.get(c.F_ELEM_H), G[pre + "H"]) <= 1e-5 and rel(h.get(c.F_VEC_B), G[pre + "b"]) <= 1e-5
    assert rel(h.implicit_apply(dt, G[pre + "d"]), G[pre + "Ad"]) <= 1e-5
    assert rel(h.implicit_apply(dt, G[pre + "d"], transpose=True), G[pre + "ATd"]) <= 1e-5
    for precond, key in ((0, "x_cg"), (1, "x_ata")):
        h.set(c.F_VEL, G[pre + "vel"])
        h.implicit_solve(dt, 1, precond)
        assert rel(h.get(c.F_VEL), G[pre + key]) <= 2e-3                 # parity at convergence (r.r <= 1e-5)
    h.close()
