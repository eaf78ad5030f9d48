# coding=utf-8
"""Module-level configuration globals, same names as the reference's constants.py:4-15.
Filled by utils.sys_init(config).  The solver wrappers read them at CALL time (the
reference snapshots them at import time, main.py:38-49 -- both orders work here)."""
import numpy as np

dim = None
auto_diff = None
use_explicit_method = None
implicit_method = None
preconditioned = None
g_dir = None
vec = None      # callable: vec(x) -> float32 array of length dim   (reference: ti.types.vector)
mat = None      # callable: mat(x) -> float32 [dim, dim]            (reference: ti.types.matrix)
index = None    # callable: index(x) -> int32 array of length dim+1
delta_time = None
# implicit_method 0 only (extension, config key 'jacobi_order' or env FEMB_JACOBI_ORDER):
#   'simultaneous' (default) -- weighted Jacobi, all rows at once: the fast GPU form, parity at convergence;
#   'reference'              -- the reference's serial in-place sweep (implicit.py:391-404) iterate for iterate,
#                               run by dependency levels (femb_implicit_solve method 2); single GPU only
jacobi_order = 'simultaneous'
JACOBIN_METHOD = 0
CONJUGATE_GRADIENT_METHOD = 1


def _make_types(d):
    def vec_(x=0.0):
        return np.broadcast_to(np.asarray(x, dtype=np.float32), (d,)).copy()

    def mat_(x=0.0):
        return np.broadcast_to(np.asarray(x, dtype=np.float32), (d, d)).copy()

    def index_(x=0):
        return np.broadcast_to(np.asarray(x, dtype=np.int32), (d + 1,)).copy()

    return vec_, mat_, index_
