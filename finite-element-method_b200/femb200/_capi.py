"""ctypes binding of libfemb200.so (include/femb200.h).

There is no CPU fallback: if the shared library is missing it is built once
with ``make`` (nvcc cross-compiles without a GPU); if that fails, or a compute
entry point is called without a CUDA device, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FEMB200_LIB") or os.path.join(_HERE, "libfemb200.so")  # env override: tuning builds
CSRC = os.path.join(os.path.dirname(_HERE), "csrc")

# femb_field
F_POS, F_VEL, F_VEL_G, F_FORCE, F_GRAD, F_MASS, F_REF_POS, F_VEC_B, F_VEC_X, F_PAST_X, F_DIAG_A = range(11)
F_ELEM_VOLUME, F_ELEM_REF, F_ELEM_PHI, F_ELEM_H = 32, 33, 34, 35
# femb_comm
COMM_NONE, COMM_HALO, COMM_ALLREDUCE, COMM_ITER = 0, 1, 2, 3

EXPORTS = [
    "femb_object_create", "femb_object_destroy", "femb_object_info", "femb_set_stream", "femb_blocks_set",
    "femb_explicit_grad", "femb_energy_grad", "femb_kinematic", "femb_explicit_step", "femb_explicit_step_timed", "femb_implicit_solve",
    "femb_halo_bind", "femb_halo_mass", "femb_explicit_step_phase",
    "femb_comm_bytes", "femb_peer_bind", "femb_peer_mass", "femb_reduce_bind", "femb_implicit_solve_begin", "femb_implicit_solve_advance", "femb_implicit_solve_poll",
    "femb_implicit_solve_finish",
    "femb_advect_implicit", "femb_implicit_assemble", "femb_implicit_apply", "femb_implicit_apply_timed", "femb_render", "femb_get", "femb_set",
    "femb_set_async", "femb_get_async", "femb_transfers_wait", "femb_set_host_threads",
    "femb_get_energy", "femb_diagnostics", "femb_trace_enable", "femb_trace_get", "femb_sync", "femb_last_error", "femb_launch_count", "femb_version",
    "femb_tiling_build", "femb_tiling_build_ex", "femb_tiling_destroy", "femb_tiling_info", "femb_tiling_get",
]


class Desc(C.Structure):
    _fields_ = [
        ("dim", C.c_int32), ("n_vert", C.c_int32), ("n_elem", C.c_int32),
        ("verts", C.POINTER(C.c_float)), ("elems", C.POINTER(C.c_int32)),
        ("center", C.c_float * 3),
        ("rho", C.c_float), ("mu", C.c_float), ("lam", C.c_float), ("damping", C.c_float),
        ("implicit", C.c_int32), ("tile_elems", C.c_int32), ("device", C.c_int32),
        ("n_halo", C.c_int32), ("halo_verts", C.POINTER(C.c_int32)),
        ("halo_owned", C.POINTER(C.c_uint8)),
    ]


class PeerDesc(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("left", C.c_int32), ("right", C.c_int32),
                ("n_left", C.c_int32), ("left_dst_off", C.c_int32), ("right_dst_off", C.c_int32),
                ("recv_stride", C.c_int32), ("bufs", C.POINTER(C.c_void_p))]


class View(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("eye", C.c_float * 3), ("lookat", C.c_float * 3),
                ("up", C.c_float * 3), ("fov_deg", C.c_float), ("prims_version", C.c_int32)]


class Info(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "dim", "n_vert", "n_elem", "n_tiles", "tile_elems", "max_tile_verts", "max_rounds",
        "n_interior", "n_boundary", "n_partials", "smem_bytes", "block_threads", "stage_stride", "n_halo_tiles")] + [("device_bytes", C.c_int64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def build(force=False):
    """Compile libfemb200.so in-tree (sm_100a).  Raises on failure."""
    if force or not os.path.exists(LIB_PATH):
        r = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("building libfemb200.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(LIB_PATH)
    vp, i32, f32 = C.c_void_p, C.c_int32, C.c_float
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    sig = {
        "femb_object_create": (C.c_int, [C.POINTER(Desc), C.POINTER(vp)]),
        "femb_object_destroy": (C.c_int, [vp]),
        "femb_object_info": (C.c_int, [vp, C.POINTER(Info)]),
        "femb_set_stream": (C.c_int, [vp, vp]),
        "femb_blocks_set": (C.c_int, [vp, i32, fp, fp]),
        "femb_explicit_grad": (C.c_int, [vp]),
        "femb_energy_grad": (C.c_int, [vp, fp]),
        "femb_kinematic": (C.c_int, [vp, f32, fp, i32]),
        "femb_explicit_step": (C.c_int, [vp, f32, fp, i32]),
        "femb_explicit_step_timed": (C.c_int, [vp, f32, fp, i32, fp, fp]),
        "femb_halo_bind": (C.c_int, [vp, vp, vp]),
        "femb_halo_mass": (C.c_int, [vp, i32]),
        "femb_explicit_step_phase": (C.c_int, [vp, f32, fp, i32]),
        "femb_implicit_solve": (C.c_int, [vp, f32, i32, i32, i32, f32, ip, fp]),
        "femb_comm_bytes": (C.c_int, [i32, C.POINTER(C.c_size_t), ip]),
        "femb_peer_bind": (C.c_int, [vp, C.POINTER(PeerDesc)]),
        "femb_peer_mass": (C.c_int, [vp]),
        "femb_reduce_bind": (C.c_int, [vp, vp]),
        "femb_implicit_solve_begin": (C.c_int, [vp, f32, i32, i32, i32, f32]),
        "femb_implicit_solve_advance": (C.c_int, [vp]),
        "femb_implicit_solve_poll": (C.c_int, [vp, ip, ip, fp]),
        "femb_implicit_solve_finish": (C.c_int, [vp, ip, fp]),
        "femb_advect_implicit": (C.c_int, [vp, f32, fp]),
        "femb_implicit_assemble": (C.c_int, [vp, f32]),
        "femb_implicit_apply": (C.c_int, [vp, f32, i32, fp, fp]),
        "femb_implicit_apply_timed": (C.c_int, [vp, f32, i32, i32, fp, fp]),
        "femb_render": (C.c_int, [vp, C.POINTER(View), ip, i32, C.POINTER(C.c_uint8)]),
        "femb_get": (C.c_int, [vp, i32, vp, C.c_size_t]),
        "femb_set": (C.c_int, [vp, i32, vp, C.c_size_t]),
        "femb_set_async": (C.c_int, [vp, i32, vp, C.c_size_t]),
        "femb_get_async": (C.c_int, [vp, i32, vp, C.c_size_t]),
        "femb_transfers_wait": (C.c_int, [vp]),
        "femb_set_host_threads": (C.c_int, [i32]),
        "femb_get_energy": (C.c_int, [vp, fp]),
        "femb_diagnostics": (C.c_int, [vp, C.POINTER(C.c_int64)]),
        "femb_trace_enable": (C.c_int, [vp, C.c_int32]),
        "femb_trace_get": (C.c_int, [vp, C.POINTER(C.c_uint64), C.c_int32, C.POINTER(C.c_int32), C.c_char_p, C.c_size_t]),
        "femb_sync": (C.c_int, [vp]),
        "femb_last_error": (C.c_char_p, []),
        "femb_launch_count": (C.c_uint64, []),
        "femb_version": (C.c_char_p, []),
        "femb_tiling_build": (C.c_int, [i32, i32, i32, fp, ip, i32, C.POINTER(vp)]),
        "femb_tiling_build_ex": (C.c_int, [i32, i32, i32, fp, ip, i32, i32, ip, i32, C.POINTER(vp)]),
        "femb_tiling_destroy": (C.c_int, [vp]),
        "femb_tiling_info": (C.c_int, [vp, C.POINTER(Info)]),
        "femb_tiling_get": (C.c_int64, [vp, i32, vp, C.c_int64]),
    }
    for name, (res, args) in sig.items():
        if os.environ.get("FEMB200_ALLOW_OLD_LIB") and not hasattr(L, name):
            continue           # tools/gpu_ab_old.sh only: same-box A/B against a library built from an older commit
        fn = getattr(L, name)  # AttributeError here == missing export: fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise RuntimeError("libfemb200: %s (status %d)" % (lib().femb_last_error().decode(), rc))


def _fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _iptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def gvec(g_dir, dim):
    g = np.zeros(3, dtype=np.float32)
    g[:dim] = np.asarray(g_dir, dtype=np.float32)[:dim]
    return g


class Handle:
    """Owns one femb_object."""

    def __init__(self, verts, elems, center, rho, mu, lam, damping, implicit=False, tile_elems=0, device=-1,
                 halo_verts=None, halo_owned=None):
        halo = np.ascontiguousarray(halo_verts if halo_verts is not None else [], dtype=np.int32)
        self.n_halo = int(halo.size)
        owned = None
        if halo_owned is not None:
            owned = np.ascontiguousarray(halo_owned, dtype=np.uint8)
            if owned.size != halo.size:
                raise ValueError("halo_owned must have one entry per halo vertex")
        verts = np.ascontiguousarray(verts, dtype=np.float32)
        elems = np.ascontiguousarray(elems, dtype=np.int32)
        self.dim = int(verts.shape[1])
        if elems.ndim != 2 or elems.shape[1] != self.dim + 1:
            raise ValueError("elements must be [n_elem, dim+1]")
        self.n_vert, self.n_elem = int(verts.shape[0]), int(elems.shape[0])
        d = Desc()
        d.dim, d.n_vert, d.n_elem = self.dim, self.n_vert, self.n_elem
        d.verts, d.elems = _fptr(verts), _iptr(elems)
        c = np.zeros(3, dtype=np.float32)
        c[:self.dim] = np.asarray(center, dtype=np.float32)[:self.dim]
        d.center = (C.c_float * 3)(*c.tolist())
        d.rho, d.mu, d.lam, d.damping = float(rho), float(mu), float(lam), float(damping)
        d.implicit, d.tile_elems, d.device = int(bool(implicit)), int(tile_elems), int(device)
        d.n_halo, d.halo_verts = self.n_halo, _iptr(halo)
        d.halo_owned = owned.ctypes.data_as(C.POINTER(C.c_uint8)) if owned is not None else None
        self._h = C.c_void_p()
        check(lib().femb_object_create(C.byref(d), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().femb_object_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    # -- fields
    def _shape(self, field):
        d, nv, ne = self.dim, self.n_vert, self.n_elem
        if field == F_MASS:
            return (nv,)
        if field in (F_ELEM_VOLUME, F_ELEM_PHI):
            return (ne,)
        if field in (F_ELEM_REF, F_ELEM_H):
            return (ne, d, d)
        return (nv, d)

    def get(self, field, out=None):
        shape = self._shape(field)
        if out is None:
            out = np.empty(shape, dtype=np.float32)
        assert out.dtype == np.float32 and out.flags.c_contiguous and out.shape == shape
        check(lib().femb_get(self._h, field, out.ctypes.data_as(C.c_void_p), out.nbytes))
        return out

    def set(self, field, arr):
        a = np.ascontiguousarray(arr, dtype=np.float32)
        if a.shape != self._shape(field):
            raise ValueError("field shape mismatch: %r vs %r" % (a.shape, self._shape(field)))
        check(lib().femb_set(self._h, field, a.ctypes.data_as(C.c_void_p), a.nbytes))

    def set_async(self, field, arr):
        """femb_set_async: `arr` (float32, C-contiguous, ideally page-locked) must stay alive until transfers_wait()."""
        if arr.dtype != np.float32 or not arr.flags.c_contiguous or arr.shape != self._shape(field):
            raise ValueError("set_async needs a C-contiguous float32 array of shape %r" % (self._shape(field),))
        check(lib().femb_set_async(self._h, field, arr.ctypes.data_as(C.c_void_p), arr.nbytes))

    def get_async(self, field, out):
        if out.dtype != np.float32 or not out.flags.c_contiguous or out.shape != self._shape(field):
            raise ValueError("get_async needs a C-contiguous float32 array of shape %r" % (self._shape(field),))
        check(lib().femb_get_async(self._h, field, out.ctypes.data_as(C.c_void_p), out.nbytes))
        return out

    def transfers_wait(self):
        check(lib().femb_transfers_wait(self._h))

    def info(self):
        i = Info()
        check(lib().femb_object_info(self._h, C.byref(i)))
        return i.as_dict()

    # -- operations
    def set_stream(self, cuda_stream):
        check(lib().femb_set_stream(self._h, C.c_void_p(int(cuda_stream))))

    def blocks_set(self, centers, radii):
        n = len(radii)
        c = np.ascontiguousarray(np.asarray(centers, dtype=np.float32).reshape(n, self.dim) if n else np.zeros((0, self.dim), np.float32))
        r = np.ascontiguousarray(radii, dtype=np.float32)
        check(lib().femb_blocks_set(self._h, n, _fptr(c), _fptr(r)))

    def explicit_grad(self):
        check(lib().femb_explicit_grad(self._h))

    def energy_grad(self, fetch=False):
        if fetch:
            u = C.c_float()
            check(lib().femb_energy_grad(self._h, C.byref(u)))
            return u.value
        check(lib().femb_energy_grad(self._h, None))

    def energy(self):
        u = C.c_float()
        check(lib().femb_get_energy(self._h, C.byref(u)))
        return u.value

    def diagnostics(self):
        """{inverted_elements, nan_elements, nan_vertices}, counted on the device."""
        out = (C.c_int64 * 3)()
        check(lib().femb_diagnostics(self._h, out))
        return {"inverted_elements": int(out[0]), "nan_elements": int(out[1]), "nan_vertices": int(out[2])}

    def trace_enable(self, on=True):
        check(lib().femb_trace_enable(self._h, int(bool(on))))

    def trace_get(self, max_slots=512):
        """[(name, start_ns, end_ns, wait_end_ns, flags_released_ns, sender_go_ns, sender_stored_ns, sender_fenced_ns, 0)]
        of the traced launches, in launch order (include/femb200.h: femb_trace_enable)."""
        words = (C.c_uint64 * (8 * max_slots))()
        n = C.c_int32()
        names = C.create_string_buffer(64 * max_slots)
        check(lib().femb_trace_get(self._h, words, max_slots, C.byref(n), names, len(names)))
        nm = names.value.decode().split("\n")
        return [(nm[i],) + tuple(int(words[8 * i + j]) for j in range(8)) for i in range(n.value)]

    def kinematic(self, dt, g_dir, use_grad):
        g = gvec(g_dir, self.dim)
        check(lib().femb_kinematic(self._h, float(dt), _fptr(g), int(use_grad)))

    def explicit_step(self, dt, g_dir, n_steps=1):
        g = gvec(g_dir, self.dim)
        check(lib().femb_explicit_step(self._h, float(dt), _fptr(g), int(n_steps)))

    # -- multi-GPU halo
    def halo_bind(self, send_ptr, recv_ptr):
        check(lib().femb_halo_bind(self._h, C.c_void_p(int(send_ptr)), C.c_void_p(int(recv_ptr))))

    def halo_mass(self, phase):
        check(lib().femb_halo_mass(self._h, int(phase)))

    def explicit_step_phase(self, dt, g_dir, phase):
        g = gvec(g_dir, self.dim)
        check(lib().femb_explicit_step_phase(self._h, float(dt), _fptr(g), int(phase)))

    def explicit_step_timed(self, dt, g_dir, n_steps):
        g = gvec(g_dir, self.dim)
        t, b = C.c_float(), C.c_float()
        check(lib().femb_explicit_step_timed(self._h, float(dt), _fptr(g), int(n_steps), C.byref(t), C.byref(b)))
        return t.value, b.value

    def implicit_solve(self, dt, method, preconditioned, max_iter=0, tol=-1.0, fetch=True):
        if fetch:
            it, res = C.c_int32(), C.c_float()
            check(lib().femb_implicit_solve(self._h, float(dt), int(method), int(preconditioned), int(max_iter),
                                            float(tol), C.byref(it), C.byref(res)))
            return it.value, res.value
        check(lib().femb_implicit_solve(self._h, float(dt), int(method), int(preconditioned), int(max_iter),
                                        float(tol), None, None))

    # -- multi-GPU over peer memory: halo sums and dot products inside the library's kernels
    def peer_bind(self, rank, world, left, right, n_left, left_dst_off, right_dst_off, recv_stride, buf_ptrs):
        d = PeerDesc()
        d.rank, d.world, d.left, d.right = int(rank), int(world), int(left), int(right)
        d.n_left, d.left_dst_off, d.right_dst_off = int(n_left), int(left_dst_off), int(right_dst_off)
        d.recv_stride = int(recv_stride)
        arr = (C.c_void_p * world)(*[C.c_void_p(int(p)) for p in buf_ptrs])
        d.bufs = C.cast(arr, C.POINTER(C.c_void_p))
        check(lib().femb_peer_bind(self._h, C.byref(d)))

    def peer_mass(self):
        check(lib().femb_peer_mass(self._h))

    # -- the solve as segments between communication points (multi-GPU callers own the collectives)
    def reduce_bind(self, dev_ptr):
        check(lib().femb_reduce_bind(self._h, C.c_void_p(int(dev_ptr))))

    def solve_begin(self, dt, method, preconditioned, max_iter=0, tol=-1.0):
        check(lib().femb_implicit_solve_begin(self._h, float(dt), int(method), int(preconditioned), int(max_iter),
                                              float(tol)))

    def solve_advance(self):
        rc = lib().femb_implicit_solve_advance(self._h)
        if rc < 0:
            check(rc)
        return rc

    def solve_poll(self):
        done, it, res = C.c_int32(), C.c_int32(), C.c_float()
        check(lib().femb_implicit_solve_poll(self._h, C.byref(done), C.byref(it), C.byref(res)))
        return bool(done.value), it.value, res.value

    def last_solve_iters(self):
        return self.solve_poll()[1]

    def solve_finish(self, fetch=True):
        if fetch:
            it, res = C.c_int32(), C.c_float()
            check(lib().femb_implicit_solve_finish(self._h, C.byref(it), C.byref(res)))
            return it.value, res.value
        check(lib().femb_implicit_solve_finish(self._h, None, None))

    def advect_implicit(self, dt, g_dir):
        g = gvec(g_dir, self.dim)
        check(lib().femb_advect_implicit(self._h, float(dt), _fptr(g)))

    def implicit_assemble(self, dt):
        check(lib().femb_implicit_assemble(self._h, float(dt)))

    def implicit_apply(self, dt, vec, transpose=False):
        a = np.ascontiguousarray(vec, dtype=np.float32)
        out = np.empty_like(a)
        check(lib().femb_implicit_apply(self._h, float(dt), int(bool(transpose)), _fptr(a), _fptr(out)))
        return out

    def implicit_apply_timed(self, dt, transpose, n):
        t, b = C.c_float(), C.c_float()
        check(lib().femb_implicit_apply_timed(self._h, float(dt), int(bool(transpose)), int(n), C.byref(t), C.byref(b)))
        return t.value, b.value

    def render(self, width=640, height=640, prims=None, prims_version=0, eye=(-6.36, 3.49, 2.44),
               lookat=(-5.40, 3.19, 2.43), up=(0.0, 1.0, 0.0), fov_deg=45.0):
        """One headless frame (render/render.py:64-98) -> uint8 [height, width, 3], row 0 at the top.
        prims: 3D surface triangles [n, 3] in the caller's vertex numbering; None in 2D (the elements)."""
        v = View(int(width), int(height), (C.c_float * 3)(*eye), (C.c_float * 3)(*lookat), (C.c_float * 3)(*up),
                 float(fov_deg), int(prims_version))
        n, pp = 0, None
        if prims is not None:
            prims = np.ascontiguousarray(prims, np.int32).reshape(-1, 3)
            n, pp = len(prims), prims.ctypes.data_as(C.POINTER(C.c_int32))
        img = np.empty((int(height), int(width), 3), np.uint8)
        check(lib().femb_render(self._h, C.byref(v), pp, n, img.ctypes.data_as(C.POINTER(C.c_uint8))))
        return img

    def sync(self):
        check(lib().femb_sync(self._h))


def comm_bytes(max_halo):
    """(bytes, recv_stride) of the comm buffer every rank must expose for halos of up to max_halo vertices."""
    n, stride = C.c_size_t(), C.c_int32()
    check(lib().femb_comm_bytes(int(max_halo), C.byref(n), C.byref(stride)))
    return int(n.value), int(stride.value)


def set_host_threads(n=0):
    """OpenMP threads of the library's host-side set-up (0: only query)."""
    return int(lib().femb_set_host_threads(int(n)))


def launch_count():
    return int(lib().femb_launch_count())


class HostTiling:
    """Host-only view of the tile layout (no CUDA calls) for tests and tools."""

    def __init__(self, verts, elems, tile_elems, halo_verts=None, wave=0):
        verts = np.ascontiguousarray(verts, dtype=np.float32)
        elems = np.ascontiguousarray(elems, dtype=np.int32)
        halo = np.ascontiguousarray(halo_verts if halo_verts is not None else [], dtype=np.int32)
        self.dim = verts.shape[1]
        self._h = C.c_void_p()
        check(lib().femb_tiling_build_ex(self.dim, verts.shape[0], elems.shape[0], _fptr(verts), _iptr(elems),
                                         int(tile_elems), int(halo.size), _iptr(halo), int(wave), C.byref(self._h)))
        i = Info()
        check(lib().femb_tiling_info(self._h, C.byref(i)))
        self.info = i.as_dict()

    def array(self, what):
        dtype = {4: np.uint16, 5: np.uint32, 6: np.uint32}.get(what, np.int32)
        n = lib().femb_tiling_get(self._h, what, None, 0)
        out = np.empty(n, dtype=dtype)
        lib().femb_tiling_get(self._h, what, out.ctypes.data_as(C.c_void_p), n)
        return out

    def __del__(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().femb_tiling_destroy(self._h)
            self._h = C.c_void_p()
