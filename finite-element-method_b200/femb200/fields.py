"""Field proxies: the subset of the Taichi field API the reference's callers use
(`.to_numpy()`, `.from_numpy()`, `[i]`, `.grad`, `U[None]`), backed by device
buffers behind the C-ABI handle."""
import numpy as np

from . import _capi


class VertexField:
    def __init__(self, obj, field, grad_field=None, writable=True):
        self._obj, self._field, self._writable = obj, field, writable
        if grad_field is not None:
            self.grad = VertexField(obj, grad_field)

    def to_numpy(self):
        if self._field == _capi.F_FORCE:
            self._obj._materialize_force()
        return self._obj._h.get(self._field)

    def from_numpy(self, arr):
        if not self._writable:
            raise RuntimeError("field is read-only")
        if self._field == _capi.F_FORCE:
            self._obj._materialize_force()
            self._obj._force_clean = False
        self._obj._h.set(self._field, arr)

    def __getitem__(self, i):
        return self.to_numpy()[i]

    @property
    def shape(self):
        return (self._obj.particle_cnt,)


class Particles:
    """obj.particles.<member> with the reference's Particle struct names (object.py:12-23)."""

    def __init__(self, obj):
        c = _capi
        self.pos = VertexField(obj, c.F_POS, grad_field=c.F_GRAD)
        self.vel = VertexField(obj, c.F_VEL)
        self.force = VertexField(obj, c.F_FORCE)
        self.mass = VertexField(obj, c.F_MASS)
        self.ref_pos = VertexField(obj, c.F_REF_POS, writable=False)
        if obj.implicit:
            self.vel_g = VertexField(obj, c.F_VEL_G)


class ElementField:
    def __init__(self, obj, field=None, host=None):
        self._obj, self._field, self._host = obj, field, host

    def to_numpy(self):
        return self._host.copy() if self._host is not None else self._obj._h.get(self._field)

    def __getitem__(self, i):
        return self.to_numpy()[i]


class Elements:
    """obj.elements.<member> with the reference's Element struct names (object.py:32-36)."""

    def __init__(self, obj, indices):
        self.vertex_indices = ElementField(obj, host=indices)
        self.volume = ElementField(obj, _capi.F_ELEM_VOLUME)
        self.ref = ElementField(obj, _capi.F_ELEM_REF)


class Meshs:
    """obj.meshs.p0/p1/p2 (object.py:25-30, render.py:70)."""

    def __init__(self, obj, faces):
        self.p0 = ElementField(obj, host=faces[:, 0].copy())
        self.p1 = ElementField(obj, host=faces[:, 1].copy())
        self.p2 = ElementField(obj, host=faces[:, 2].copy())


class EnergyScalar:
    """obj.U -- `obj.U[None]` reads the total energy of the last compute_energy (object.py:69)."""

    def __init__(self, obj):
        self._obj = obj

    def __getitem__(self, _):
        return self._obj._h.energy()

    def to_numpy(self):
        return np.float32(self._obj._h.energy())


class Tape:
    """Stand-in for `ti.ad.Tape(loss=obj.U)` (main.py:107): compute_energy here produces
    phi, U and dU/dx in one pass, so the context manager has nothing left to do."""

    def __init__(self, loss=None):
        self.loss = loss

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False
