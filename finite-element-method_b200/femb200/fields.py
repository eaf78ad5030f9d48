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

    def to_numpy(self, out=None, blocking=True):
        """Taichi's field.to_numpy().  Extensions: `out` (a float32 array, page-locked for speed) receives the copy;
        blocking=False only enqueues it (device->host copy stream) -- call obj.sync() before reading `out`."""
        if self._field == _capi.F_FORCE:
            self._obj._materialize_force()
        if not blocking:
            if out is None:
                raise ValueError("to_numpy(blocking=False) needs out=")
            return self._obj._h.get_async(self._field, out)
        return self._obj._h.get(self._field, out=out)

    def from_numpy(self, arr, blocking=True):
        """Taichi's field.from_numpy().  Extension: blocking=False enqueues the copy on the host->device copy stream
        (it overlaps the kernels of earlier calls); `arr` must stay untouched until obj.sync()."""
        if not self._writable:
            raise RuntimeError("field is read-only")
        if not blocking:
            self._obj._materialize_force()
            if self._field == _capi.F_FORCE:
                self._obj._force_clean = False
            return self._obj._h.set_async(self._field, arr)
        # a requested-but-deferred neo_hookean_1_grad must see the state of ITS call time (explicit.py:8): run it
        # before any vertex field changes under it
        self._obj._materialize_force()
        if self._field == _capi.F_FORCE:
            self._obj._force_clean = False
        self._obj._h.set(self._field, arr)

    def __getitem__(self, i):
        return self.to_numpy()[i]

    @property
    def shape(self):
        return (self._obj.particle_cnt,)


class DenseOperator:
    """Debug densifier for obj.matrix_A / matrix_AT / matrix_ATA (reference object.py:82,91,92; read in the
    commented debug block of main.py:14-28).  The solver never forms these Nv x Nv block matrices (108 Nv^2
    bytes); for SMALL meshes `.to_numpy()` builds [Nv, Nv, dim, dim] column by column from the matrix-free
    operator at the CURRENT positions (femb_implicit_assemble + femb_implicit_apply on unit vectors):
    block [i, j][:, k] = (A e_{j,k})_i.  matrix_ATA follows implicit.py:289-307: A^T A when
    constants.preconditioned == 1, else A itself."""

    MAX_VERTICES = 2000

    def __init__(self, obj, kind):
        self._obj, self._kind = obj, kind

    @property
    def shape(self):
        return (self._obj.particle_cnt, self._obj.particle_cnt)

    def to_numpy(self, dt=None):
        import constants
        obj, nv = self._obj, self._obj.particle_cnt
        if nv > self.MAX_VERTICES:
            raise RuntimeError("dense {} is a debug aid for <= {} vertices (this object has {}): the operator is "
                               "matrix-free".format(self._kind, self.MAX_VERTICES, nv))
        dt = constants.delta_time if dt is None else dt
        dim = obj._h.dim
        obj._h.implicit_assemble(dt)
        A = np.empty((nv, nv, dim, dim), np.float32)
        e = np.zeros((nv, dim), np.float32)

        def column(j, k):
            e[j, k] = 1.0
            if self._kind == 'matrix_AT':
                y = obj._h.implicit_apply(dt, e, transpose=True)
            else:
                y = obj._h.implicit_apply(dt, e)
                if self._kind == 'matrix_ATA' and constants.preconditioned == 1:
                    y = obj._h.implicit_apply(dt, y, transpose=True)
            e[j, k] = 0.0
            return y

        for j in range(nv):
            for k in range(dim):
                A[:, j, :, k] = column(j, k)
        return A

    def __getitem__(self, ij):
        return self.to_numpy()[ij]


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
