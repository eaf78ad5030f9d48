# coding=utf-8
"""neo_hookean_1_grad(obj): accumulate +dE/dx of the Neo-Hookean energy into
obj.particles.force (reference solver/explicit.py:8-49), on the GPU through
libfemb200.  The launch is deferred by one call: when kinematic() follows (the
reference's step loop, main.py:105-110) both run as ONE fused kernel pass so the
gradient never round-trips through HBM; reading particles.force in between
materialises it with the stand-alone kernel."""


def neo_hookean_1_grad(obj):
    if obj._pending_grad:          # two calls in a row: force accumulates twice, as in the reference
        obj._materialize_force()
    obj._pending_grad = True
