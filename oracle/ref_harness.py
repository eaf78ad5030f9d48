"""Runs the UNMODIFIED reference sources (/root/reference) on the CPU under oracle/ti_shim.

TEST INFRASTRUCTURE ONLY.  Used by tests/golden/make_ref_golden.py (fixture generator) and by the
`not gpu` tests that re-derive fixtures when /root/reference is present (this container only; the GPU box
has no /root/reference and only reads the committed tests/golden/ref_*.npz).

Recipe = SURVEY.md section 8c: put the shim's `taichi` and the empty trimesh/pyvista/tetgen stubs on sys.path,
`utils.sys_init(config)` FIRST, then import constants / solver.* / object / circle_blocks in main.py:42-49's order
(those modules snapshot `constants.*` at import), never construct Render (window), subclass Object only to inject
(vertices, faces, tets) in 3D (object.py:128-161 needs TetGen), and drive the kernels exactly as main.py:101-113.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

import numpy as np

REFERENCE = os.environ.get("FEMB_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
_REF_MODULES = ("constants", "utils", "object", "circle_blocks", "solver", "solver.explicit", "solver.explicit_auto_diff",
                "solver.kinematic", "solver.implicit", "main", "render", "render.render")


def available():
    return os.path.exists(os.path.join(REFERENCE, "solver", "implicit.py"))


def load_reference(config, real=None):
    """Fresh import of the reference for one top-level JSON config (dict).  real=np.float64 runs it in double."""
    for p in (os.path.join(HERE, "ti_shim", "stubs"), os.path.join(HERE, "ti_shim"), REFERENCE):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    for m in _REF_MODULES:
        sys.modules.pop(m, None)
    import taichi as ti
    assert getattr(ti, "__file__", "").startswith(HERE), "the real taichi is importable: use it instead of the shim"
    ti.force_real(real)
    ti._GRAD_FIELDS.clear()
    utils = importlib.import_module("utils")
    assert utils.__file__.startswith(REFERENCE), utils.__file__
    utils.sys_init(config)                                     # main.py:38
    ti.init(ti.cpu, debug=False)                               # main.py:40 (arch: cpu)
    ref = types.SimpleNamespace(ti=ti, utils=utils, config=config)
    ref.constants = importlib.import_module("constants")      # main.py:42
    ref.kinematic = importlib.import_module("solver.kinematic")           # main.py:44
    ref.implicit = importlib.import_module("solver.implicit")             # main.py:45
    ref.explicit_auto_diff = importlib.import_module("solver.explicit_auto_diff")  # main.py:46
    ref.explicit = importlib.import_module("solver.explicit")             # main.py:47
    ref.object = importlib.import_module("object")                        # main.py:48
    ref.circle_blocks = importlib.import_module("circle_blocks")          # main.py:49
    ref.main = importlib.import_module("main")                 # only for fem(), main.py:8-12
    for name, mod in (("constants", ref.constants), ("neo_hookean_1_grad", ref.explicit.neo_hookean_1_grad),
                      ("implicit_solver_neo_hookean", ref.implicit.implicit_solver_neo_hookean)):
        setattr(ref.main, name, mod)                           # the names main.py's __main__ block imports
    for m in (ref.constants, ref.kinematic, ref.implicit, ref.explicit, ref.object, ref.circle_blocks, ref.main):
        assert m.__file__.startswith(REFERENCE), m.__file__
    return ref


def make_object(ref, obj_cfg, verts=None, tets=None):
    """object.Object(obj_cfg); in 3D the tet mesh is injected in place of TetGen, the surface comes from the
    reference's own extract_surface (object.py:255-308)."""
    ti = ref.ti
    Object = ref.object.Object
    if ref.constants.dim == 2:
        return Object(obj_cfg)
    tet = types.SimpleNamespace(elem=np.asarray(tets), node=np.asarray(verts, np.float64))
    surface, _ = Object.extract_surface(tet)

    class InjectedObject(Object):
        def load_obj(self, config):
            self.center = ti.Vector(config.get("center"))      # object.py:160
            return tet.node, np.asarray(surface), tet.elem, 4

    return InjectedObject(obj_cfg)


def substep(ref, obj, blocks):
    """One iteration of main.py:103-112."""
    c = ref.constants
    if c.auto_diff == 0:
        ref.main.fem(obj)                                      # main.py:105 -> main.py:8-12
    else:
        with ref.ti.ad.Tape(loss=obj.U):                       # main.py:107
            ref.explicit_auto_diff.compute_energy(obj)         # main.py:108
    if c.use_explicit_method or c.auto_diff == 1:
        ref.kinematic.kinematic(obj, blocks)                   # main.py:110
    else:
        ref.implicit.advect_implicit(obj, blocks)              # main.py:112


def fem_only(ref, obj):
    """The force / energy / solve half of a substep (main.py:104-108), without the vertex update."""
    c = ref.constants
    if c.auto_diff == 0:
        ref.main.fem(obj)
    else:
        with ref.ti.ad.Tape(loss=obj.U):
            ref.explicit_auto_diff.compute_energy(obj)


def advance_only(ref, obj, blocks):
    c = ref.constants
    if c.use_explicit_method or c.auto_diff == 1:
        ref.kinematic.kinematic(obj, blocks)
    else:
        ref.implicit.advect_implicit(obj, blocks)
