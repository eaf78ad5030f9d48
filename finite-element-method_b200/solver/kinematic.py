# coding=utf-8
"""kinematic(obj, circle_blocks): symplectic-Euler vertex update with damping, unit-box
walls and circle/sphere blocks (reference solver/kinematic.py:9-45)."""
import constants


def kinematic(obj, circle_blocks):
    obj._sync_blocks(circle_blocks)
    dt, g = constants.delta_time, constants.g_dir
    if constants.auto_diff == 1:
        if obj._halo is not None:
            raise RuntimeError("the auto_diff path is not available on a slab-partitioned object (use the analytic "
                               "explicit step: identical gradient, fused with its halo exchange)")
        obj._materialize_force()
        obj._h.kinematic(dt, g, 1)             # reads pos.grad, leaves force alone (kinematic.py:44)
    elif obj._pending_grad and obj._force_clean:
        obj._explicit_step(dt, g, 1)           # fused gradient + update (+ halo exchange on a slab)
        obj._pending_grad = False
    else:
        obj._materialize_force()
        obj._h.kinematic(dt, g, 0)             # reads force, zeroes it
        obj._force_clean = True
