"""CPU oracle: numpy restatement of the reference's Neo-Hookean FEM hot path.

TEST INFRASTRUCTURE ONLY.  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s cpu-baseline / ``--impl reference`` legs may import this package.
The product path (``finite-element-method_b200/``) never imports it and has no
CPU fallback.

PARITY UNPINNED BY THE REFERENCE: the reference (Jukgei/Finite-Element-Method)
ships no tests, golden vectors or fixtures and its runtime
(taichi==1.6.0, requirements.txt:1) is not installable here, so it cannot be
executed to produce vectors either.  This file is therefore a line-by-line
restatement of the reference *source*, cross-checked by independent invariants
(tests/test_oracle.py): analytic gradient == autograd / finite-difference
gradient of the restated energy, dense assembly == matrix-free operator,
transpose consistency, solver results vs numpy.linalg.solve.

Every function cites the reference file:line it follows (paths relative to the
reference root).  Arithmetic is done in the dtype of the inputs (float32 to
mirror ``ti.f32``, utils.py:19-20; float64 for the adjudicator) with the
reference's association order: differences first, then ``Ds @ Dm^-1``,
adjugate inverse, ``log(det F)``.
"""
from __future__ import annotations

import math

import numpy as np

JACOBIN_METHOD = 0  # constants.py:14
CONJUGATE_GRADIENT_METHOD = 1  # constants.py:15


# --------------------------------------------------------------------------
# material / meshes
# --------------------------------------------------------------------------
def lame(E, nu):
    """object.py:47-48 -- mu, s_lambda from Young's modulus / Poisson ratio."""
    mu = E / 2 / (1 + nu)
    s_lambda = E * nu / (1 + nu) / (1 - 2 * nu)
    return mu, s_lambda


def construct_2d_mesh(side_length, subdivisions):
    """object.py:95-126 -- regular (s+1)^2 grid, two triangles per cell."""
    s = subdivisions
    x = np.linspace(0, side_length, s + 1)
    y = np.linspace(0, side_length, s + 1)
    vertices = np.array(np.meshgrid(x, y)).T.reshape(-1, 2)
    faces = np.empty((s * s * 2, 3), dtype=np.int32)
    k = 0
    for i in range(s):
        for j in range(s):
            p1 = i * (s + 1) + j
            p2 = p1 + 1
            p3 = p1 + s + 1
            p4 = p3 + 1
            faces[k] = (p1, p2, p4)
            faces[k + 1] = (p1, p4, p3)
            k += 2
    return vertices, faces


_KUHN_PERMS = ((0, 1, 2), (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0))


def kuhn_box(nx, ny, nz, h, dtype=np.float64):
    """Structured Kuhn/Freudenthal box (not in the reference --
    the synthetic input generator of BASELINE configs 3 and 4).

    Vertex id (i*(ny+1)+j)*(nz+1)+k at (i,j,k)*h; every cell gives 6 tets
    [v000, v000+e_p1, v000+e_p1+e_p2, v111] over the permutations p.
    """
    i, j, k = np.meshgrid(np.arange(nx + 1), np.arange(ny + 1), np.arange(nz + 1), indexing="ij")
    verts = (np.stack([i, j, k], axis=-1).reshape(-1, 3) * h).astype(dtype)
    ci, cj, ck = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    ci, cj, ck = ci.ravel(), cj.ravel(), ck.ravel()
    stride = np.array([(ny + 1) * (nz + 1), nz + 1, 1], dtype=np.int64)
    v000 = ci * stride[0] + cj * stride[1] + ck * stride[2]
    tets = np.empty((v000.size, 6, 4), dtype=np.int64)
    for t, p in enumerate(_KUHN_PERMS):
        a = v000 + stride[p[0]]
        b = a + stride[p[1]]
        c = b + stride[p[2]]
        tets[:, t, 0] = v000
        tets[:, t, 1] = a
        tets[:, t, 2] = b
        tets[:, t, 3] = c
    return verts, tets.reshape(-1, 4).astype(np.int32)


def particles_init(raw_vertices, center, dtype=np.float32):
    """object.py:337-343 -- pos = raw + center (center is an OFFSET), ref_pos = pos."""
    return raw_vertices.astype(dtype) + np.asarray(center).astype(dtype)[None, :]


# --------------------------------------------------------------------------
# small-matrix helpers in the reference's association order (batched, last
# two axes are the matrix).  No BLAS: every op is an explicit elementwise op
# so float32 stays float32 and the summation order is fixed.
# --------------------------------------------------------------------------
def _matmul(a, b):
    """Taichi matrix '@': c[i,j] = sum_k a[i,k]*b[k,j], k ascending."""
    n = a.shape[-1]
    out = np.empty(a.shape[:-2] + (a.shape[-2], b.shape[-1]), dtype=a.dtype)
    for i in range(a.shape[-2]):
        for j in range(b.shape[-1]):
            acc = a[..., i, 0] * b[..., 0, j]
            for k in range(1, n):
                acc = acc + a[..., i, k] * b[..., k, j]
            out[..., i, j] = acc
    return out


def _det(m):
    """Cofactor expansion along the first row (taichi 1.6 Matrix.determinant)."""
    if m.shape[-1] == 2:
        return m[..., 0, 0] * m[..., 1, 1] - m[..., 0, 1] * m[..., 1, 0]
    a = m
    return (a[..., 0, 0] * (a[..., 1, 1] * a[..., 2, 2] - a[..., 2, 1] * a[..., 1, 2])
            - a[..., 1, 0] * (a[..., 0, 1] * a[..., 2, 2] - a[..., 2, 1] * a[..., 0, 2])
            + a[..., 2, 0] * (a[..., 0, 1] * a[..., 1, 2] - a[..., 1, 1] * a[..., 0, 2]))


def _inverse(m):
    """Closed-form adjugate inverse (ti.math.inverse for n<=3): adj * (1/det)."""
    out = np.empty_like(m)
    one = m.dtype.type(1.0)
    if m.shape[-1] == 2:
        inv_det = one / _det(m)
        out[..., 0, 0] = m[..., 1, 1] * inv_det
        out[..., 0, 1] = -m[..., 0, 1] * inv_det
        out[..., 1, 0] = -m[..., 1, 0] * inv_det
        out[..., 1, 1] = m[..., 0, 0] * inv_det
        return out
    inv_det = one / _det(m)
    for i in range(3):
        for j in range(3):
            i1, i2 = (i + 1) % 3, (i + 2) % 3
            j1, j2 = (j + 1) % 3, (j + 2) % 3
            # inverse[j,i] = cofactor(i,j) / det
            out[..., j, i] = inv_det * (m[..., i1, j1] * m[..., i2, j2] - m[..., i2, j1] * m[..., i1, j2])
    return out


def _transpose(m):
    return np.swapaxes(m, -1, -2)


def _edge_matrix(pos, elems):
    """X[:, j] = p_{j+1} - p_0  (explicit.py:11-16, object.py:350-355)."""
    dim = pos.shape[1]
    p0 = pos[elems[:, 0]]
    X = np.empty((elems.shape[0], dim, dim), dtype=pos.dtype)
    for j in range(dim):
        X[:, :, j] = pos[elems[:, j + 1]] - p0
    return X


# --------------------------------------------------------------------------
# setup (K2)
# --------------------------------------------------------------------------
def compute_volume(r):
    """object.py:375-392 -- |det|/2 in 2D, |p0.(p1 x p2)|/6 in 3D."""
    dt = r.dtype.type
    if r.shape[-1] == 2:
        p0, p1 = r[..., :, 0], r[..., :, 1]
        cross = p0[..., 0] * p1[..., 1] - p0[..., 1] * p1[..., 0]
        return np.abs(cross) / dt(2)
    p0, p1, p2 = r[..., :, 0], r[..., :, 1], r[..., :, 2]
    cx = p1[..., 1] * p2[..., 2] - p1[..., 2] * p2[..., 1]
    cy = p1[..., 2] * p2[..., 0] - p1[..., 0] * p2[..., 2]
    cz = p1[..., 0] * p2[..., 1] - p1[..., 1] * p2[..., 0]
    dot = p0[..., 0] * cx + p0[..., 1] * cy + p0[..., 2] * cz
    return dt(1 / 6) * np.abs(dot)


def elements_init(pos, elems, rho):
    """object.py:345-362 -- volume, lumped mass (+= V*rho/(dim+1) per corner), ref = Dm^-1."""
    dim = pos.shape[1]
    dt = pos.dtype.type
    r = _edge_matrix(pos, elems)
    volume = compute_volume(r)
    mass = np.zeros(pos.shape[0], dtype=pos.dtype)
    contrib = dt(1.0 / (dim + 1)) * volume * dt(rho)
    # element by element, corners j = 0..dim (the serial order of object.py:357-360; np.add.at is unbuffered)
    np.add.at(mass, elems.reshape(-1), np.repeat(contrib, dim + 1))
    ref = _inverse(r)
    return ref, volume, mass


# --------------------------------------------------------------------------
# explicit analytic gradient (K4) and energy (K5)
# --------------------------------------------------------------------------
def element_gradient(pos, elems, ref, volume, mu, s_lambda):
    """explicit.py:11-28 -- per-element G = V*(mu F - mu F^-T + lambda log(det F) F^-T) @ Dm^-T."""
    dt = pos.dtype.type
    X = _edge_matrix(pos, elems)
    F = _matmul(X, ref)
    F_inv_T = _transpose(_inverse(F))
    log_j = np.log(_det(F))
    P = dt(mu) * F - dt(mu) * F_inv_T + (dt(s_lambda) * log_j)[:, None, None] * F_inv_T
    G = _matmul(P, _transpose(ref))
    G = G * volume[:, None, None]
    return G


def scatter_gradient(G, elems, out):
    """explicit.py:42-49 -- out[v_{j+1}] += G[:, j]; out[v_0] += -sum_j G[:, j]."""
    dim = out.shape[1]
    f0 = np.zeros((elems.shape[0], dim), dtype=out.dtype)
    vals = np.empty((elems.shape[0], dim + 1, dim), dtype=out.dtype)
    for j in range(dim):
        f = G[:, :, j]
        vals[:, j] = f
        f0 = f0 - f
    vals[:, dim] = f0
    # element by element: corners 1..dim, then corner 0 (the serial order of explicit.py:42-49)
    order = list(range(1, dim + 1)) + [0]
    np.add.at(out, elems[:, order].reshape(-1), vals.reshape(-1, dim))
    return out


def neo_hookean_1_grad(pos, elems, ref, volume, mu, s_lambda, force):
    """explicit.py:8-49 -- accumulates +dE/dx into ``force`` (in place)."""
    G = element_gradient(pos, elems, ref, volume, mu, s_lambda)
    return scatter_gradient(G, elems, force)


def compute_energy(pos, elems, ref, volume, mu, s_lambda):
    """explicit_auto_diff.py:8-30 -- phi[i] = V*phi_i, U = sum (sequential order)."""
    dt = pos.dtype.type
    dim = pos.shape[1]
    X = _edge_matrix(pos, elems)
    F = _matmul(X, ref)
    log_j = np.log(_det(F))
    FtF = _matmul(_transpose(F), F)
    tr = FtF[:, 0, 0]
    for k in range(1, dim):
        tr = tr + FtF[:, k, k]
    phi = dt(mu) / dt(2) * (tr - dt(dim))
    phi = phi - dt(mu) * log_j
    phi = phi + dt(s_lambda) / dt(2) * log_j ** 2
    phi_v = phi * volume
    U = dt(0)
    for v in (volume * phi):  # U[None] += V * phi_i  (explicit_auto_diff.py:30)
        U = dt(U + v)
    return phi_v, U


def energy_autograd(pos, elems, ref, volume, mu, s_lambda, dtype=None):
    """The 'autodiff oracle' (main.py:107-108 ti.ad.Tape over compute_energy):
    reverse-mode gradient of the restated energy, here via torch.autograd on CPU.
    Returns (U, dU/dpos)."""
    import torch

    tdt = {np.float32: torch.float32, np.float64: torch.float64}[np.dtype(dtype or pos.dtype).type]
    x = torch.tensor(pos, dtype=tdt, requires_grad=True)
    R = torch.tensor(ref, dtype=tdt)
    V = torch.tensor(volume, dtype=tdt)
    idx = torch.tensor(elems.astype(np.int64))
    dim = pos.shape[1]
    p0 = x[idx[:, 0]]
    X = torch.stack([x[idx[:, j + 1]] - p0 for j in range(dim)], dim=2)
    F = X @ R
    log_j = torch.log(torch.linalg.det(F))
    phi = mu / 2 * ((F * F).sum(dim=(1, 2)) - dim) - mu * log_j + s_lambda / 2 * log_j ** 2
    U = (phi * V).sum()
    (g,) = torch.autograd.grad(U, x)
    return float(U.detach()), g.numpy()


# --------------------------------------------------------------------------
# explicit integrator (K6)
# --------------------------------------------------------------------------
def kinematic(pos, vel, grad, mass, dt, g_dir, damping, blocks, auto_diff=0):
    """kinematic.py:9-45.  ``grad`` is particles.force (auto_diff==0, zeroed on
    exit) or particles.pos.grad (auto_diff==1, left alone).  ``blocks`` is a
    list of (center, radius).  All arrays are updated in place."""
    ft = pos.dtype.type
    dim = pos.shape[1]
    dt = ft(dt)
    g = np.asarray(g_dir).astype(pos.dtype)
    # acc == 0 always (object.py:17): (acc + 9.8*g - grad/m) * dt
    vel += ((ft(9.8) * g)[None, :] - grad / mass[:, None]) * dt
    vel *= np.exp(-dt * ft(damping))
    for i in range(dim):  # walls: unit box, old positions (kinematic.py:25-30)
        lo = (pos[:, i] < 0) & (vel[:, i] < 0)
        vel[lo, i] = 0
        hi = (pos[:, i] > 1) & (vel[:, i] > 0)
        vel[hi, i] = 0
    for center, radius in blocks:  # kinematic.py:32-41
        if radius <= 0.0:
            continue
        c = np.asarray(center).astype(pos.dtype)
        disp = pos - c[None, :]
        norm = np.sqrt(_dot(disp, disp))
        inward = _dot(vel, -disp) > 0
        hit = (norm < ft(radius)) & inward
        # reference order: vel.dot(disp) * disp / disp.norm_sqr()
        with np.errstate(invalid="ignore", divide="ignore"):
            corr = (_dot(vel, disp)[:, None] * disp) / _dot(disp, disp)[:, None]
        vel[hit] -= corr[hit]
    pos += vel * dt
    if auto_diff == 0:
        grad[...] = 0
    return pos, vel


def _dot(a, b):
    acc = a[:, 0] * b[:, 0]
    for k in range(1, a.shape[1]):
        acc = acc + a[:, k] * b[:, k]
    return acc


# --------------------------------------------------------------------------
# implicit path -- literal dense restatement (K7) ...
# --------------------------------------------------------------------------
def element_hessian_block(pos, elems, ref, volume, mu, s_lambda):
    """implicit.py:116-147 -- the single dim x dim block H_e the reference
    scatters with a star pattern.  dF = I @ R_inv (only the i==j branch is
    non-zero, implicit.py:141-144); log_J = log(max(det F, 1e-4)) (131)."""
    ft = pos.dtype.type
    dim = pos.shape[1]
    X = _edge_matrix(pos, elems)
    F = _matmul(X, ref)
    F_inv = _inverse(F)
    log_j = np.log(np.maximum(_det(F), ft(1e-4)))
    F_inv_T = _transpose(F_inv)
    dF = ref
    dF_T = _transpose(dF)
    FiR = _matmul(F_inv, dF)
    tr = FiR[:, 0, 0]
    for k in range(1, dim):
        tr = tr + FiR[:, k, k]
    term = ft(mu) * dF \
        + _matmul(_matmul((ft(mu) - ft(s_lambda) * log_j)[:, None, None] * F_inv_T, dF_T), F_inv_T) \
        + (ft(s_lambda) * tr)[:, None, None] * F_inv_T
    H = _matmul((-volume)[:, None, None] * term, _transpose(ref))
    return H


def compute_linear_system_matrix_a(pos, elems, ref, volume, mass, mu, s_lambda, dt):
    """implicit.py:114-194 -- dense A[Nv,Nv,dim,dim] = I - dt^2 * (1/m_row) * K,
    K scattered as L_star (x) H_e: (v_i,v_i)+=H, (v_i,v_0)-=H, (v_0,v_i)-=H,
    (v_0,v_0)+=dim*H."""
    ft = pos.dtype.type
    dim = pos.shape[1]
    nv = pos.shape[0]
    H = element_hessian_block(pos, elems, ref, volume, mu, s_lambda)
    A = np.zeros((nv, nv, dim, dim), dtype=pos.dtype)
    v0 = elems[:, 0]
    ne = elems.shape[0]
    H00 = np.zeros_like(H)
    rows, cols, blks = [], [], []
    for i in range(dim):                   # element by element, in the order the reference's loop issues its adds
        vi = elems[:, i + 1]
        rows += [vi, vi]
        cols += [vi, v0]
        blks += [H, -H]                    # implicit.py:151 (i == j), 161 (dF_dxi0 = 0 - ... - H)
        H00 = H00 + H                      # implicit.py:149
    rows.append(v0)
    cols.append(v0)
    blks.append(H00)                       # implicit.py:168
    for i in range(dim):
        rows.append(v0)
        cols.append(elems[:, i + 1])
        blks.append(-H)                    # implicit.py:157,178 (temp block = 0 - ... - H)
    rows = np.stack(rows, 1).reshape(-1)
    cols = np.stack(cols, 1).reshape(-1)
    blks = np.stack(blks, 1).reshape(ne * len(blks), dim, dim)
    np.add.at(A, (rows, cols), blks)
    # implicit.py:183-187: (dt**2 [python double] * ((1/m) * eye)) @ A[i,j]  ==  f32(f32(dt^2) * (1/m)) * A[i,j]
    scale = ft(float(dt) ** 2) * (ft(1.0) / mass)
    A = scale[:, None, None, None] * A
    # implicit.py:189-194: A = I - A  (off-diagonal entries become 0 - a)
    eye = np.zeros((nv, nv, dim, dim), dtype=pos.dtype)
    eye[np.arange(nv), np.arange(nv)] = np.eye(dim, dtype=pos.dtype)
    return eye - A


def internal_force_implicit(pos, elems, ref, volume, mu, s_lambda):
    """implicit.py:72-88 -- f = -V (mu F - mu F^-T + (lambda/2) log det(F^T F) F^-T) Dm^-T."""
    ft = pos.dtype.type
    X = _edge_matrix(pos, elems)
    F = _matmul(X, ref)
    F_inv_T = _transpose(_inverse(F))
    log_term = np.log(_det(_matmul(_transpose(F), F)))
    P = ft(mu) * F - ft(mu) * F_inv_T + (ft(s_lambda) / ft(2) * log_term)[:, None, None] * F_inv_T
    G = _matmul(P, _transpose(ref))
    return G * (-volume)[:, None, None]


def compute_linear_system_vector_b(pos, vel, elems, ref, volume, mass, mu, s_lambda, dt):
    """implicit.py:65-105 -- b = vel; b[v] += dt * (1/m_v) * f_v.  No gravity."""
    ft = pos.dtype.type
    dim = pos.shape[1]
    b = vel.copy()
    G = internal_force_implicit(pos, elems, ref, volume, mu, s_lambda)
    f0 = np.zeros((elems.shape[0], dim), dtype=pos.dtype)
    vals = np.empty((elems.shape[0], dim + 1, dim), dtype=pos.dtype)
    for i in range(dim):
        f_n = G[:, :, i]
        f0 = f0 - f_n
        m_ = ft(1.0) / mass[elems[:, i + 1]]
        vals[:, i] = (ft(dt) * m_)[:, None] * f_n
    m_ = ft(1.0) / mass[elems[:, 0]]
    vals[:, dim] = (ft(dt) * m_)[:, None] * f0
    order = list(range(1, dim + 1)) + [0]              # element by element: corners 1..dim, then 0 (implicit.py:92-105)
    np.add.at(b, elems[:, order].reshape(-1), vals.reshape(-1, dim))
    return b


def _block_terms(A, x):
    """t[i, j, :] = A[i, j] @ x[j] with Taichi's matmul order (k ascending)."""
    t = A[..., 0] * x[None, :, None, 0]
    for k in range(1, x.shape[1]):
        t = t + A[..., k] * x[None, :, None, k]
    return t


def _dense_matvec(A, x):
    """v[i] = ((0 + A[i,0] @ x[0]) + A[i,1] @ x[1]) + ...  -- j ascending, one add per block, exactly the serial
    order of implicit.py:281-282, 347-349 (ufunc.accumulate is sequential)."""
    return np.add.accumulate(_block_terms(A, x), axis=1)[:, -1]


def _seq_sum(v):
    """((0 + v0) + v1) + ... in the dtype of v."""
    return np.add.accumulate(v)[-1]


def compute_error(A, b, x):
    """implicit.py:277-286 -- sqrt(sum_i ||b_i - (A x)_i||^2), each term as norm()**2 = sqrt(.)^2."""
    r = b - _dense_matvec(A, x)
    nrm = np.sqrt(_dot(r, r))
    return np.sqrt(_seq_sum(nrm * nrm))


def jacobi_iter_field_once(A, b, x, past_x):
    """implicit.py:391-404 -- ONE serial in-place sweep (rows see already-updated x: Gauss-Seidel order)."""
    ft = A.dtype.type
    nv, dim = b.shape
    omega = ft(0.75)
    for i in range(nv):
        t = _block_terms(A[i:i + 1], x)[0]                         # A[i,j] @ x[j] for every j, current x
        bi = np.subtract.accumulate(np.concatenate([b[i][None], t]), axis=0)[-1]   # b -= ... , j ascending
        for k in range(dim):
            a_ii = A[i, i, k, k]
            if abs(a_ii) < 1e-6:
                x[i, k] = 0.0
            else:
                bk = bi[k] + a_ii * x[i, k]
                x[i, k] = omega * bk / a_ii + (ft(1) - omega) * past_x[i, k]
    return x


def jacobi_iter_field(A, b, past_x, threshold=1e-5, max_iter=20000, trace=None):
    """implicit.py:227-262 + 391-404 -- the reference's 'Jacobi': a SERIAL
    in-place sweep (rows see already-updated x, i.e. Gauss-Seidel ordering),
    per-component diagonal solve, damped with omega=0.75 against past_vec_x,
    which is stale from the previous time step on the first iteration.
    Returns (x, past_x, iterations).  `trace` (a list) receives (err, x) per iteration."""
    ft = A.dtype.type
    x = ft(0.5) * b                                   # implicit.py:231-232
    err = compute_error(A, b, x)
    p_err = err
    it = 0
    past_x = past_x.copy()
    while err > threshold and it < max_iter:
        jacobi_iter_field_once(A, b, x, past_x)
        err = compute_error(A, b, x)
        it += 1
        if trace is not None:
            trace.append((float(err), x.copy()))
        if err >= p_err:                              # implicit.py:257-259
            x = past_x.copy()
            break
        p_err = err
        past_x = x.copy()                             # implicit.py:260-261
    return x, past_x, it


def cg(A, b, preconditioned, iter_max=500, tol=1e-5, trace=None):
    """implicit.py:311-386 -- textbook CG on ATA x = ATb (preconditioned==1:
    AT[i,j] = A[j,i]^T, ATb, ATA built block by block, implicit.py:289-299) or on A x = b (302-307); x0 = b;
    absolute stop delta_new <= 1e-5; no residual refresh.  Every sum runs in the reference's serial order.
    Returns (x, iterations, delta)."""
    nv, dim = b.shape
    if preconditioned == 1:
        AT = A.transpose(1, 0, 3, 2).copy()
        ATb = _dense_matvec(AT, b)                    # 0 + AT[i,0]@b[0] + ...
        ATA = np.zeros_like(A)
        for k in range(nv):                           # ATA[i,j] += AT[i,k] @ A[k,j], k ascending (296-299)
            a, c = AT[:, k], A[k]                     # [nv,dim,dim] (i), [nv,dim,dim] (j)
            blk = a[:, None, :, 0, None] * c[None, :, None, 0, :]
            for m in range(1, dim):
                blk = blk + a[:, None, :, m, None] * c[None, :, None, m, :]
            ATA = ATA + blk
    else:
        ATb = b.copy()
        ATA = A
    x = b.copy()
    # r[i] = ATb[i]; r[i] -= ATA[i,j] @ x[j], j ascending (324-327)
    r = np.subtract.accumulate(np.concatenate([ATb[:, None], _block_terms(ATA, x)], axis=1), axis=1)[:, -1]
    d = r.copy()
    delta_new = _seq_sum(_dot(r, r))
    it = 0
    while it < iter_max and delta_new > tol:
        q = _dense_matvec(ATA, d)
        alpha = delta_new / _seq_sum(_dot(d, q))
        x = x + alpha * d
        r = r - alpha * q
        delta_old = delta_new
        delta_new = _seq_sum(_dot(r, r))
        beta = delta_new / delta_old
        d = r + beta * d
        it += 1
        if trace is not None:
            trace.append((float(delta_new), x.copy()))
    return x, it, float(delta_new)


def implicit_solver_neo_hookean(pos, vel, elems, ref, volume, mass, mu, s_lambda, dt,
                                implicit_method, preconditioned, past_x):
    """implicit.py:201-223 -- assemble, solve, vel = x.  Returns (vel, past_x, info)."""
    A = compute_linear_system_matrix_a(pos, elems, ref, volume, mass, mu, s_lambda, dt)
    b = compute_linear_system_vector_b(pos, vel, elems, ref, volume, mass, mu, s_lambda, dt)
    if implicit_method == JACOBIN_METHOD:
        x, past_x, it = jacobi_iter_field(A, b, past_x)
        info = {"iters": it, "resid": float(compute_error(A, b, x))}
    else:
        x, it, delta = cg(A, b, preconditioned)
        info = {"iters": it, "resid": delta}
    return x, past_x, info


def advect_implicit(pos, vel, vel_g, dt, g_dir, damping, blocks):
    """implicit.py:407-437 -- gravity split, damping, walls (upper wall does NOT
    zero vel_g, 420-423), blocks project v, vel, vel_g separately, pos += v dt."""
    ft = pos.dtype.type
    dim = pos.shape[1]
    dt0 = dt
    dt = ft(dt)
    g = np.asarray(g_dir).astype(pos.dtype)
    vel_g += (ft(9.8) * g * dt)[None, :]
    # implicit.py:410-411: `ti.exp(-delta_time * obj.damping)` has only Python floats inside, so Taichi folds it in
    # DOUBLE at compile time and rounds the factor to f32 (kinematic.py:23 differs: there dt, damping are f32 args)
    damp = ft(math.exp(-float(dt0) * float(damping)))
    vel *= damp
    vel_g *= damp
    v = vel + vel_g
    for i in range(dim):
        lo = (pos[:, i] < 0) & (v[:, i] < 0)
        vel[lo, i] = 0
        vel_g[lo, i] = 0
        v[lo, i] = 0
        hi = (pos[:, i] > 1) & (v[:, i] > 0)
        vel[hi, i] = 0
        v[hi, i] = 0
    for center, radius in blocks:
        if radius <= 0.0:
            continue
        c = np.asarray(center).astype(pos.dtype)
        disp = pos - c[None, :]
        norm = np.sqrt(_dot(disp, disp))
        hit = (norm < ft(radius)) & (_dot(v, -disp) > 0)
        nsq = _dot(disp, disp)[:, None]
        with np.errstate(invalid="ignore", divide="ignore"):
            cv = (_dot(v, disp)[:, None] * disp) / nsq
            cvel = (_dot(vel, disp)[:, None] * disp) / nsq
            cvg = (_dot(vel_g, disp)[:, None] * disp) / nsq
        v[hit] -= cv[hit]
        vel[hit] -= cvel[hit]
        vel_g[hit] -= cvg[hit]
    pos += v * dt
    return pos, vel, vel_g


# --------------------------------------------------------------------------
# ... and the same operator matrix-free, proven equal to the
# dense restatement in tests/test_oracle.py.  This is what scales.
# --------------------------------------------------------------------------
def apply_K(H, elems, d, transpose=False):
    """(K d): e_j = d_{v_j} - d_{v_0}; y_{v_j} += H e_j; y_{v_0} -= sum_j H e_j."""
    dim = d.shape[1]
    Hm = _transpose(H) if transpose else H
    y = np.zeros_like(d)
    d0 = d[elems[:, 0]]
    s = np.zeros((elems.shape[0], dim), dtype=d.dtype)
    for j in range(dim):
        e = d[elems[:, j + 1]] - d0
        he = np.einsum("nkl,nl->nk", Hm, e)
        np.add.at(y, elems[:, j + 1], he)
        s = s + he
    np.add.at(y, elems[:, 0], -s)
    return y


def apply_A(H, elems, mass, dt, d):
    """A d = d - dt^2 M^-1 (K d)   (implicit.py:183-194 applied to a vector)."""
    ft = d.dtype.type
    return d - (ft(dt) ** 2 * (ft(1.0) / mass))[:, None] * apply_K(H, elems, d)


def apply_AT(H, elems, mass, dt, d):
    """A^T d = d - dt^2 K^T (M^-1 d)."""
    ft = d.dtype.type
    return d - ft(dt) ** 2 * apply_K(H, elems, (ft(1.0) / mass)[:, None] * d, transpose=True)


def diag_A(H, elems, mass, dt):
    """diag(A)_{v,k} = 1 - dt^2/m_v * sum_e c_{e,v} H_e[k,k], c = dim for local vertex 0 else 1."""
    ft = H.dtype.type
    dim = H.shape[-1]
    hd = np.stack([H[:, k, k] for k in range(dim)], axis=1)
    acc = np.zeros((mass.shape[0], dim), dtype=H.dtype)
    for j in range(dim):
        np.add.at(acc, elems[:, j + 1], hd)
    np.add.at(acc, elems[:, 0], ft(dim) * hd)
    return ft(1.0) - (ft(dt) ** 2 * (ft(1.0) / mass))[:, None] * acc


def cg_matrix_free(H, elems, mass, dt, b, preconditioned, iter_max=500, tol=1e-5):
    """implicit.py:311-386 with the dense products replaced by operator passes."""
    ft = b.dtype.type

    def op(v):
        av = apply_A(H, elems, mass, dt, v)
        return apply_AT(H, elems, mass, dt, av) if preconditioned == 1 else av

    def dot(a, c):
        return ft((a * c).sum(dtype=np.float64))

    rhs = apply_AT(H, elems, mass, dt, b) if preconditioned == 1 else b.copy()
    x = b.copy()
    r = rhs - op(x)
    d = r.copy()
    delta_new = dot(r, r)
    it = 0
    while it < iter_max and delta_new > tol:
        q = op(d)
        alpha = delta_new / dot(d, q)
        x = x + alpha * d
        r = r - alpha * q
        delta_old = delta_new
        delta_new = dot(r, r)
        beta = delta_new / delta_old
        d = r + beta * d
        it += 1
    return x, it, float(delta_new)


def jacobi_parallel_matrix_free(H, elems, mass, dt, b, past_x, threshold=1e-5, max_iter=20000):
    """The GPU build's solver for implicit_method==0 restated on the CPU: a TRUE
    (simultaneous) weighted Jacobi with the reference's omega, guard, stale
    past_x, threshold and revert/break rule (implicit.py:227-262, 391-404), but
    with the residual of the *current* iterate taken from the same operator
    pass that drives the update.  Not the reference's iterate path (that one is
    a serial sweep); agreement with jacobi_iter_field is at convergence only."""
    ft = b.dtype.type
    omega = ft(0.75)
    diag = diag_A(H, elems, mass, dt)
    x = ft(0.5) * b
    p_err = None
    it = 0
    while True:
        ax = apply_A(H, elems, mass, dt, x)
        r = b - ax
        err = ft(np.sqrt((r * r).sum(dtype=np.float64)))
        if p_err is not None:
            if err >= p_err:
                x = past_x.copy()
                break
            past_x = x.copy()
        p_err = err
        if not (err > threshold and it < max_iter):
            break
        small = np.abs(diag) < 1e-6
        with np.errstate(divide="ignore", invalid="ignore"):
            xn = omega * (r + diag * x) / diag + (ft(1) - omega) * past_x
        x = np.where(small, ft(0), xn).astype(b.dtype)
        it += 1
    return x, past_x, it
