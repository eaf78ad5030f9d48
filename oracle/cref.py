"""ctypes view of oracle/libfem_ref.so (C/OpenMP restatement; TEST INFRASTRUCTURE ONLY --
see fem_oracle.py for who may import this)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "libfem_ref.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
        _lib = C.CDLL(_PATH)
        _lib.fem_ref_threads.restype = C.c_int
        _lib.fem_ref_cg.restype = C.c_int
        _lib.fem_ref_jacobi.restype = C.c_int
    return _lib


def _f(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _i(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def threads():
    return lib().fem_ref_threads()


def set_threads(n=None):
    """Use `n` (default: every host core) OpenMP threads, whatever OMP_NUM_THREADS the launcher exported
    (torchrun sets it to 1)."""
    lib().fem_ref_set_threads(C.c_int(int(n or os.cpu_count() or 1)))
    return threads()


def elements_init(pos, elems, rho):
    """object.py:345-362 in C: (ref, volume, mass); mass by atomic adds (order differs from the serial oracle)."""
    nv, d = pos.shape
    ref = np.empty((len(elems), d, d), dtype=np.float32)
    vol = np.empty(len(elems), dtype=np.float32)
    mass = np.empty(nv, dtype=np.float32)
    lib().fem_ref_elements_init(C.c_int(d), C.c_int64(nv), C.c_int64(len(elems)), _i(elems), _f(pos), C.c_float(rho),
                                _f(ref), _f(vol), _f(mass))
    return ref, vol, mass


def _blocks(blocks, d):
    n = len(blocks)
    c = np.ascontiguousarray([b[0] for b in blocks], dtype=np.float32).reshape(n, d) if n else np.zeros((1, d), np.float32)
    r = np.ascontiguousarray([b[1] for b in blocks], dtype=np.float32) if n else np.zeros(1, np.float32)
    return n, c, r


def explicit_grad(pos, elems, ref, vol, mu, lam, force):
    d = pos.shape[1]
    lib().fem_ref_explicit_grad(C.c_int(d), C.c_int64(len(elems)), _i(elems), _f(pos), _f(ref), _f(vol),
                                C.c_float(mu), C.c_float(lam), _f(force))
    return force


def explicit_steps(pos, vel, force, mass, elems, ref, vol, mu, lam, dt, g_dir, damping, blocks, n_steps):
    d = pos.shape[1]
    g = np.ascontiguousarray(g_dir, dtype=np.float32)
    n, c, r = _blocks(blocks, d)
    lib().fem_ref_explicit_steps(C.c_int(d), C.c_int64(len(pos)), C.c_int64(len(elems)), _i(elems), _f(pos), _f(vel),
                                 _f(force), _f(mass), _f(ref), _f(vol), C.c_float(mu), C.c_float(lam), C.c_float(dt),
                                 _f(g), C.c_float(damping), C.c_int(n), _f(c), _f(r), C.c_int(n_steps))


def hessian_blocks(pos, elems, ref, vol, mu, lam):
    d = pos.shape[1]
    H = np.empty((len(elems), d, d), dtype=np.float32)
    lib().fem_ref_hessian_blocks(C.c_int(d), C.c_int64(len(elems)), _i(elems), _f(pos), _f(ref), _f(vol),
                                 C.c_float(mu), C.c_float(lam), _f(H))
    return H


def apply(H, elems, mass, dt, x, transpose=False):
    nv, d = x.shape
    y = np.empty_like(x)
    scratch = np.empty(2 * nv * d, dtype=np.float32)
    lib().fem_ref_apply(C.c_int(d), C.c_int64(nv), C.c_int64(len(elems)), _i(elems), _f(H), _f(mass), C.c_float(dt),
                        C.c_int(int(transpose)), _f(x), _f(y), _f(scratch))
    return y


def cg(H, elems, mass, dt, b, preconditioned, iter_max=500, tol=1e-5):
    nv, d = b.shape
    x = np.empty_like(b)
    work = np.empty(8 * nv * d, dtype=np.float32)
    delta = C.c_float()
    it = lib().fem_ref_cg(C.c_int(d), C.c_int64(nv), C.c_int64(len(elems)), _i(elems), _f(H), _f(mass), C.c_float(dt),
                          C.c_int(int(preconditioned)), _f(b), _f(x), C.c_int(iter_max), C.c_float(tol), _f(work),
                          C.byref(delta))
    return x, it, delta.value


def jacobi(H, elems, mass, dt, b, past_x, iter_max=20000, tol=1e-5):
    """Simultaneous weighted Jacobi with the matrix-free operator (timing baseline; see fem_ref.c)."""
    nv, d = b.shape
    x = np.empty_like(b)
    work = np.empty(4 * nv * d, dtype=np.float32)
    err = C.c_float()
    it = lib().fem_ref_jacobi(C.c_int(d), C.c_int64(nv), C.c_int64(len(elems)), _i(elems), _f(H), _f(mass), C.c_float(dt),
                              _f(b), _f(x), _f(past_x), C.c_int(iter_max), C.c_float(tol), _f(work), C.byref(err))
    return x, it, err.value
