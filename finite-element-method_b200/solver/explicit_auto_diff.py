# coding=utf-8
"""compute_energy(obj): per-element energy phi, total U and dU/dx into
obj.particles.pos.grad (reference solver/explicit_auto_diff.py:8-30 run under
ti.ad.Tape, main.py:107-108).  The gradient is the closed-form derivative of the
same energy (verified against autograd in tests), produced in the same pass."""


def compute_energy(obj):
    obj._h.energy_grad()
