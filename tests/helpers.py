"""Shared fixtures for the parity tests: seeded meshes/states and error norms."""
import numpy as np

from oracle import fem_oracle as o


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


MAT3 = dict(E=4e4, nu=0.4, rho=1000.0, damping=10.0)      # config/demo_3d.json:16-21
MAT2 = dict(E=4e4, nu=0.2, rho=500.0, damping=14.5)       # default.json:16-23


def cube_case(n=8, h=0.4 / 55, center=(0.3, 0.5, 0.3), jitter=0.1, seed=0, dims=None):
    nx, ny, nz = dims or (n, n, n)
    verts, el = o.kuhn_box(nx, ny, nz, h)
    verts = verts.astype(np.float32)
    pos0 = o.particles_init(verts, center)
    rng = np.random.default_rng(seed)
    pos = (pos0 + (jitter * h * rng.uniform(-1, 1, pos0.shape)).astype(np.float32)).astype(np.float32)
    return verts, el, np.asarray(center, np.float32), pos0, pos


def square_case(s=10, side=0.2, center=(0.5, 0.8), jitter=0.1, seed=1):
    verts, el = o.construct_2d_mesh(side, s)
    verts = verts.astype(np.float32)
    pos0 = o.particles_init(verts, center)
    rng = np.random.default_rng(seed)
    pos = (pos0 + (jitter * side / s * rng.uniform(-1, 1, pos0.shape)).astype(np.float32)).astype(np.float32)
    return verts, el, np.asarray(center, np.float32), pos0, pos


def make_handle(verts, el, center, mat, implicit=False, tile_elems=0):
    from femb200 import _capi

    mu, lam = o.lame(mat["E"], mat["nu"])
    h = _capi.Handle(verts, el, center, mat["rho"], mu, lam, mat["damping"], implicit=implicit, tile_elems=tile_elems)
    return h, mu, lam
