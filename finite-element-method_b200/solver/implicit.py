# coding=utf-8
"""implicit_solver_neo_hookean(obj) and advect_implicit(obj, circle_blocks)
(reference solver/implicit.py:201-223, 407-437), matrix-free on the GPU.  The dense
Nv x Nv matrix_A / matrix_AT / matrix_ATA fields of the reference are never formed."""
import os

import constants

FEMB_METHOD_SWEEP = 2   # include/femb200.h


def _method():
    m = constants.implicit_method
    if m == constants.JACOBIN_METHOD:
        order = os.environ.get("FEMB_JACOBI_ORDER") or getattr(constants, "jacobi_order", None) or "simultaneous"
        if order not in ("simultaneous", "reference"):
            raise ValueError("jacobi_order must be 'simultaneous' or 'reference', not %r" % (order,))
        if order == "reference":
            return FEMB_METHOD_SWEEP
    return m


def implicit_solver_neo_hookean(obj):
    it, resid = obj._implicit_solve(constants.delta_time, _method(), 1 if constants.preconditioned == 1 else 0)
    obj.last_solve = {"iters": it, "resid": resid}


def advect_implicit(obj, circle_blocks):
    obj._sync_blocks(circle_blocks)
    obj._h.advect_implicit(constants.delta_time, constants.g_dir)
