"""The oracle has no reference golden vectors to be pinned against (SURVEY.md
section 4: the reference ships none and taichi is not installable), so it is checked
against the independent invariants SURVEY.md section 8c lists."""
import numpy as np
import pytest

from oracle import fem_oracle as o

MU, LAM = o.lame(4e4, 0.4)


def _cube(n=3, h=0.1, jitter=0.1, seed=0, dtype=np.float32):
    verts, tets = o.kuhn_box(n, n, n, h)
    pos0 = o.particles_init(verts, [0.3, 0.5, 0.3], dtype)
    ref, vol, mass = o.elements_init(pos0, tets, 1000.0)
    rng = np.random.default_rng(seed)
    pos = (pos0 + (jitter * h * rng.uniform(-1, 1, pos0.shape)).astype(dtype)).astype(dtype)
    return pos0, pos, tets, ref, vol, mass


def _square(s=4, jitter=0.1, seed=1, dtype=np.float32):
    verts, faces = o.construct_2d_mesh(0.2, s)
    pos0 = o.particles_init(verts, [0.5, 0.8], dtype)
    ref, vol, mass = o.elements_init(pos0, faces, 500.0)
    rng = np.random.default_rng(seed)
    pos = (pos0 + (jitter * 0.2 / s * rng.uniform(-1, 1, pos0.shape)).astype(dtype)).astype(dtype)
    return pos0, pos, faces, ref, vol, mass


def rel(a, b):
    return np.linalg.norm(a.astype(np.float64) - b.astype(np.float64)) / max(np.linalg.norm(b.astype(np.float64)), 1e-300)


def test_2d_mesh_shape_default_json():
    verts, faces = o.construct_2d_mesh(0.2, 10)
    assert verts.shape == (121, 2) and faces.shape == (200, 3)  # default.json:18-19
    assert np.allclose(verts[1], [0.0, 0.02]) and np.allclose(verts[11], [0.02, 0.0])
    assert faces[0].tolist() == [0, 1, 12] and faces[1].tolist() == [0, 12, 11]


def test_kuhn_box_counts_and_volume():
    verts, tets = o.kuhn_box(4, 3, 2, 0.1)
    assert verts.shape == (5 * 4 * 3, 3) and tets.shape == (4 * 3 * 2 * 6, 4)
    _, vol, mass = o.elements_init(verts, tets, 1000.0)
    assert np.allclose(vol, 0.1 ** 3 / 6)
    assert np.isclose(mass.sum(), 1000.0 * vol.sum())  # invariant (v)
    # conforming: every interior face is shared by exactly two tets
    faces = {}
    for t in tets:
        for f in ((0, 1, 2), (0, 1, 3), (0, 2, 3), (1, 2, 3)):
            key = tuple(sorted(t[list(f)]))
            faces[key] = faces.get(key, 0) + 1
    assert set(faces.values()) == {1, 2}


@pytest.mark.parametrize("maker", [_cube, _square])
def test_rest_state_has_zero_gradient_and_energy(maker):
    pos0, _, el, ref, vol, _ = maker(dtype=np.float64)
    G = o.element_gradient(pos0, el, ref, vol, MU, LAM)
    assert np.abs(G).max() < 1e-9  # invariant (ii)
    phi, U = o.compute_energy(pos0, el, ref, vol, MU, LAM)
    assert abs(U) < 1e-12


@pytest.mark.parametrize("maker", [_cube, _square])
def test_analytic_gradient_equals_autodiff(maker):
    """invariant (i): explicit.py analytic force == d/dx of explicit_auto_diff.py energy."""
    _, pos, el, ref, vol, _ = maker(dtype=np.float64)
    f = np.zeros_like(pos)
    o.neo_hookean_1_grad(pos, el, ref, vol, MU, LAM, f)
    U, g = o.energy_autograd(pos, el, ref, vol, MU, LAM)
    assert rel(f, g) < 1e-11
    phi, U2 = o.compute_energy(pos, el, ref, vol, MU, LAM)
    assert np.isclose(U, U2, rtol=1e-12)
    # central differences on a few coordinates
    eps = 1e-6
    for (v, k) in [(0, 0), (5, 1), (pos.shape[0] - 1, pos.shape[1] - 1)]:
        p = pos.copy(); p[v, k] += eps
        m = pos.copy(); m[v, k] -= eps
        fd = (o.compute_energy(p, el, ref, vol, MU, LAM)[1] - o.compute_energy(m, el, ref, vol, MU, LAM)[1]) / (2 * eps)
        assert abs(fd - f[v, k]) < 1e-5 * max(1.0, abs(f[v, k]))


@pytest.mark.parametrize("maker", [_cube, _square])
def test_fp32_gradient_close_to_fp64_on_same_inputs(maker):
    _, pos, el, ref, vol, _ = maker(dtype=np.float32)
    f32 = np.zeros_like(pos)
    o.neo_hookean_1_grad(pos, el, ref, vol, MU, LAM, f32)
    f64 = np.zeros(pos.shape)
    o.neo_hookean_1_grad(pos.astype(np.float64), el, ref.astype(np.float64), vol.astype(np.float64), MU, LAM, f64)
    assert f32.dtype == np.float32
    assert rel(f32, f64) < 2e-5
    assert np.abs(f64.sum(axis=0)).max() < 1e-9 * np.abs(f64).max() * len(f64)  # invariant (iii)


def test_rigid_motion_gives_zero_gradient():
    pos0, _, el, ref, vol, _ = _cube(dtype=np.float64)
    th = 0.3
    Rm = np.array([[np.cos(th), -np.sin(th), 0], [np.sin(th), np.cos(th), 0], [0, 0, 1]])
    pos = pos0 @ Rm.T + np.array([0.1, -0.2, 0.05])
    G = o.element_gradient(pos, el, ref, vol, MU, LAM)
    assert np.abs(G).max() < 1e-8  # invariant (iv)


@pytest.mark.parametrize("maker", [_cube, _square])
def test_dense_assembly_equals_matrix_free_operator(maker):
    """invariants (vi) and (vii)."""
    _, pos, el, ref, vol, mass = maker(dtype=np.float64)
    dt = 5e-4
    A = o.compute_linear_system_matrix_a(pos, el, ref, vol, mass, MU, LAM, dt)
    nv, dim = pos.shape
    M = A.transpose(0, 2, 1, 3).reshape(nv * dim, nv * dim)
    H = o.element_hessian_block(pos, el, ref, vol, MU, LAM)
    rng = np.random.default_rng(3)
    d = rng.standard_normal((nv, dim))
    assert rel(o.apply_A(H, el, mass, dt, d).reshape(-1), M @ d.reshape(-1)) < 1e-13
    assert rel(o.apply_AT(H, el, mass, dt, d).reshape(-1), M.T @ d.reshape(-1)) < 1e-13
    assert rel(o.diag_A(H, el, mass, dt).reshape(-1), np.diag(M)) < 1e-13


def test_hessian_block_literal_transcription():
    """element_hessian_block uses the simplification dF = R_inv; check it against a
    literal per-(i,j) transcription of implicit.py:135-178 on one tet."""
    _, pos, el, ref, vol, mass = _cube(n=1, dtype=np.float64)
    e = 2
    idx = el[e]
    X = np.stack([pos[idx[j + 1]] - pos[idx[0]] for j in range(3)], axis=1)
    R = ref[e]
    F = X @ R
    Fi = np.linalg.inv(F)
    FiT = Fi.T
    logJ = np.log(max(np.linalg.det(F), 1e-4))
    nv = pos.shape[0]
    A = np.zeros((nv, nv, 3, 3))
    d00 = np.zeros((3, 3)); temp = np.zeros((3, 9))
    for i in range(3):
        di0 = np.zeros((3, 3))
        for j in range(3):
            dF = (np.eye(3) if i == j else np.zeros((3, 3))) @ R
            blk = MU * dF + (MU - LAM * logJ) * FiT @ dF.T @ FiT + LAM * np.trace(Fi @ dF) * FiT
            blk = -vol[e] * blk @ R.T
            d00 += blk
            A[idx[i + 1], idx[j + 1]] += blk
            di0 -= blk
            temp[:, j * 3:j * 3 + 3] -= blk
        A[idx[i + 1], idx[0]] += di0
    A[idx[0], idx[0]] += d00
    for i in range(3):
        A[idx[0], idx[i + 1]] += temp[:, i * 3:i * 3 + 3]
    H = o.element_hessian_block(pos, el[e:e + 1], ref[e:e + 1], vol[e:e + 1], MU, LAM)
    d = np.random.default_rng(0).standard_normal((nv, 3))
    y = np.einsum("ijkl,jl->ik", A, d)
    assert rel(o.apply_K(H, el[e:e + 1], d), y) < 1e-12


@pytest.mark.parametrize("precond", [0, 1])
def test_cg_dense_and_matrix_free_match_direct_solve(precond):
    """invariant (viii).  h=0.1, dt=5e-4 is the well-conditioned regime."""
    _, pos, el, ref, vol, mass = _cube(n=3, h=0.05, dtype=np.float64)
    nv, dim = pos.shape
    vel = np.random.default_rng(5).standard_normal((nv, dim)) * 0.1
    dt = 5e-4
    A = o.compute_linear_system_matrix_a(pos, el, ref, vol, mass, MU, LAM, dt)
    b = o.compute_linear_system_vector_b(pos, vel, el, ref, vol, mass, MU, LAM, dt)
    M = A.transpose(0, 2, 1, 3).reshape(nv * dim, nv * dim)
    x_direct = np.linalg.solve(M, b.reshape(-1)).reshape(nv, dim)
    x, it, delta = o.cg(A, b, precond, tol=1e-16)
    assert rel(x, x_direct) < 1e-6
    H = o.element_hessian_block(pos, el, ref, vol, MU, LAM)
    x2, it2, _ = o.cg_matrix_free(H, el, mass, dt, b, precond, tol=1e-16)
    assert rel(x2, x_direct) < 1e-6


def test_reference_jacobi_sweep_and_parallel_jacobi_agree_at_convergence():
    _, pos, el, ref, vol, mass = _cube(n=2, h=0.1, dtype=np.float64)
    nv, dim = pos.shape
    vel = np.random.default_rng(6).standard_normal((nv, dim)) * 0.1
    dt = 5e-4
    A = o.compute_linear_system_matrix_a(pos, el, ref, vol, mass, MU, LAM, dt)
    b = o.compute_linear_system_vector_b(pos, vel, el, ref, vol, mass, MU, LAM, dt)
    past = np.zeros_like(b)
    x, past1, it = o.jacobi_iter_field(A, b, past.copy())
    assert it > 0 and o.compute_error(A, b, x) < 1e-4
    H = o.element_hessian_block(pos, el, ref, vol, MU, LAM)
    x2, past2, it2 = o.jacobi_parallel_matrix_free(H, el, mass, dt, b, past.copy())
    M = A.transpose(0, 2, 1, 3).reshape(nv * dim, nv * dim)
    xd = np.linalg.solve(M, b.reshape(-1)).reshape(nv, dim)
    assert rel(x, xd) < 1e-3 and rel(x2, xd) < 1e-3


def test_kinematic_walls_blocks_and_force_clear():
    pos = np.array([[0.5, -0.01], [0.5, 1.01], [0.75, 0.5], [0.5, 0.5]], dtype=np.float32)
    vel = np.array([[0.0, -1.0], [0.0, 1.0], [1.0, 0.0], [0.0, 0.0]], dtype=np.float32)
    force = np.zeros_like(pos); mass = np.ones(4, dtype=np.float32)
    o.kinematic(pos, vel, force, mass, 1e-3, [0, 0], 0.0, [([0.8, 0.5], 0.21)], auto_diff=0)
    assert vel[0, 1] == 0 and vel[1, 1] == 0      # kinematic.py:25-30
    assert abs(vel[2, 0]) < 1e-7                  # inward normal velocity removed (32-41)
    assert np.all(force == 0)


def test_advect_implicit_upper_wall_keeps_vel_g():
    pos = np.array([[0.5, 1.5, 0.5], [0.5, -0.5, 0.5]], dtype=np.float32)
    vel = np.array([[0, 1.0, 0], [0, -1.0, 0]], dtype=np.float32)
    vel_g = np.array([[0, 2.0, 0], [0, -2.0, 0]], dtype=np.float32)
    o.advect_implicit(pos, vel, vel_g, 1e-3, [0, 0, 0], 0.0, [])
    assert vel[0, 1] == 0 and vel_g[0, 1] == 2.0   # implicit.py:420-423
    assert vel[1, 1] == 0 and vel_g[1, 1] == 0.0   # implicit.py:415-418
    assert pos[0, 1] == np.float32(1.5) and pos[1, 1] == np.float32(-0.5)


@pytest.mark.parametrize("maker", [_cube, _square])
def test_homogeneous_deformation_matches_closed_form_neo_hookean(maker):
    """Known-answer test that does not come from the restatement itself: under a homogeneous deformation
    x = F X of the whole body, linear elements reproduce F exactly, so with
        phi(F) = mu/2 (tr F^T F - dim) - mu ln J + lambda/2 (ln J)^2,   P(F) = mu (F - F^-T) + lambda ln J F^-T
    (the energy of explicit_auto_diff.py:26-28 and its first Piola stress, explicit.py:25-28)
        U = V_body phi(F),   sum_v g_v X_v^T = V_body P(F)   (virtual work, g = dU/dx),
    and every vertex strictly inside the body is in equilibrium (g_v = 0)."""
    pos0, _, el, ref, vol, _ = maker(dtype=np.float64)
    dim = pos0.shape[1]
    F = np.array([[1.20, 0.10, 0.00], [0.00, 0.90, 0.05], [0.03, 0.00, 1.10]])[:dim, :dim]
    pos = pos0 @ F.T
    J = np.linalg.det(F)
    FinvT = np.linalg.inv(F).T
    phi = MU / 2 * (np.trace(F.T @ F) - dim) - MU * np.log(J) + LAM / 2 * np.log(J) ** 2
    P = MU * (F - FinvT) + LAM * np.log(J) * FinvT
    V = vol.sum()
    _, U = o.compute_energy(pos, el, ref, vol, MU, LAM)
    assert abs(U - V * phi) <= 1e-12 * abs(V * phi)
    g = np.zeros_like(pos)
    o.neo_hookean_1_grad(pos, el, ref, vol, MU, LAM, g)            # "force" = +dU/dx (explicit.py:42-49)
    assert rel((g[:, :, None] * pos0[:, None, :]).sum(axis=0), V * P) <= 1e-11
    lo, hi = pos0.min(axis=0), pos0.max(axis=0)
    inside = np.all((pos0 > lo + 1e-9) & (pos0 < hi - 1e-9), axis=1)
    assert inside.sum() > 0 and np.abs(g[inside]).max() <= 1e-9 * np.abs(g).max()
    # the autodiff oracle agrees with the same closed form
    _, ga = o.energy_autograd(pos, el, ref, vol, MU, LAM, dtype=np.float64)
    assert rel((ga[:, :, None] * pos0[:, None, :]).sum(axis=0), V * P) <= 1e-11
    # fp32 on the same inputs: within the tolerance the GPU parity tests use
    p32, r32, v32 = pos.astype(np.float32), ref.astype(np.float32), vol.astype(np.float32)
    g32, g64 = np.zeros(pos.shape, np.float32), np.zeros(pos.shape)
    o.neo_hookean_1_grad(p32, el, r32, v32, MU, LAM, g32)
    o.neo_hookean_1_grad(p32.astype(np.float64), el, r32.astype(np.float64), v32.astype(np.float64), MU, LAM, g64)
    assert rel(g32, g64) <= 1e-5


def test_free_fall_with_damping_matches_closed_form():
    """Known answer for the integrator (kinematic.py:15-23, 43): with no elastic force the update is
    v <- (v + g dt) q,  x <- x + v dt,  q = exp(-dt c), so after n steps
    v_n = g dt q (1 - q^n) / (1 - q)   and   x_n = x_0 + dt * sum_k v_k  (geometric sums)."""
    dt, c, n = 5e-4, 14.5, 400
    g = np.array([0.0, -1.0])
    pos = np.array([[0.5, 0.9]], dtype=np.float64)
    vel = np.zeros_like(pos)
    mass = np.ones(1)
    for _ in range(n):
        o.kinematic(pos, vel, np.zeros_like(pos), mass, dt, g, c, [], auto_diff=0)
    q = np.exp(-dt * c)
    a = 9.8 * g * dt
    v_n = a * q * (1 - q ** n) / (1 - q)
    # sum_{k=1..n} v_k = a q / (1 - q) * (n - q (1 - q^n) / (1 - q))
    x_n = np.array([0.5, 0.9]) + dt * a * q / (1 - q) * (n - q * (1 - q ** n) / (1 - q))
    assert np.allclose(vel[0], v_n, rtol=1e-12, atol=0) or np.allclose(vel[0], v_n, atol=1e-15)
    assert np.allclose(pos[0], x_n, rtol=1e-12)
    # terminal velocity g / c is approached from below (damping 14.5 of default.json: 0.676 m/s)
    assert abs(vel[0, 1]) < 9.8 / c and abs(vel[0, 1]) > 0.9 * 9.8 / c * (1 - q ** n)
