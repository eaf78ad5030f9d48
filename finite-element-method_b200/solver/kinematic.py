# coding=utf-8
"""kinematic(obj, circle_blocks): symplectic-Euler vertex update with damping, unit-box
walls and circle/sphere blocks (reference solver/kinematic.py:9-45)."""
import constants


def kinematic(obj, circle_blocks):
    obj._sync_blocks(circle_blocks)
    dt, g = constants.delta_time, constants.g_dir
    if constants.auto_diff == 1:
        obj._materialize_force()
        obj._h.kinematic(dt, g, 1)             # reads pos.grad, leaves force alone (kinematic.py:44)
    elif obj._pending_grad and obj._force_clean:
        obj._h.explicit_step(dt, g, 1)         # fused gradient + update
        obj._pending_grad = False
    else:
        obj._materialize_force()
        obj._h.kinematic(dt, g, 0)             # reads force, zeroes it
        obj._force_clean = True
