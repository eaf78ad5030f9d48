"""Host logic of the headless driver (finite-element-method_b200/main.py) without a GPU: the step dispatch of the
reference's loop (main.py:8-12, 103-112), checkpoint save / restore and read_config's exit(3)."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finite-element-method_b200"))


class _Field:
    def __init__(self, a):
        self.a = np.asarray(a, np.float32)

    def to_numpy(self):
        return self.a.copy()

    def from_numpy(self, a):
        self.a = np.asarray(a, np.float32).copy()


class _Obj:
    def __init__(self, n, implicit, seed):
        r = np.random.default_rng(seed)
        self.implicit = implicit
        self.particles = type("P", (), {})()
        self.particles.pos = _Field(r.standard_normal((n, 3)))
        self.particles.vel = _Field(r.standard_normal((n, 3)))
        if implicit:
            self.particles.vel_g = _Field(r.standard_normal((n, 3)))
            self.past_vec_x = _Field(r.standard_normal((n, 3)))


@pytest.fixture()
def main_mod():
    import constants
    m = importlib.import_module("main")
    calls = []
    m.constants = constants
    m.neo_hookean_1_grad = lambda o: calls.append("grad")
    m.implicit_solver_neo_hookean = lambda o: calls.append("solve")
    m.compute_energy = lambda o: calls.append("energy")
    m.advect_implicit = lambda o, b: calls.append("advect")
    m.ki = type("K", (), {"kinematic": staticmethod(lambda o, b: calls.append("kinematic"))})
    return m, constants, calls


def test_step_dispatch_follows_the_reference_loop(main_mod):
    m, constants, calls = main_mod
    for use_explicit, auto_diff, want in [
        (True, 0, ["grad", "kinematic"]),          # main.py:103-110 analytic explicit
        (True, 1, ["energy", "kinematic"]),        # main.py:107-110 autodiff (Tape around compute_energy)
        (False, 0, ["solve", "advect"]),           # main.py:105, 112 implicit
        (False, 1, ["energy", "kinematic"]),       # auto_diff wins over the implicit switch (main.py:104, 109)
    ]:
        constants.use_explicit_method, constants.auto_diff = use_explicit, auto_diff
        del calls[:]
        m.step(object(), object())
        assert calls == want


def test_checkpoint_roundtrip_and_mismatch(main_mod, tmp_path):
    m, _, _ = main_mod
    objs = [_Obj(5, False, 1), _Obj(7, True, 2)]
    path = str(tmp_path / "ck.npz")
    m.save_checkpoint(path, objs, 42, 0.21, ply_cnt=13, output_frame_cnt=12)
    fresh = [_Obj(5, False, 11), _Obj(7, True, 12)]
    frame, vt, counters = m.load_checkpoint(path, fresh)
    assert counters == (13, 12)
    assert (frame, vt) == (42, 0.21)
    for a, b in zip(objs, fresh):
        assert np.array_equal(a.particles.pos.a, b.particles.pos.a) and np.array_equal(a.particles.vel.a, b.particles.vel.a)
    assert np.array_equal(objs[1].particles.vel_g.a, fresh[1].particles.vel_g.a)
    assert np.array_equal(objs[1].past_vec_x.a, fresh[1].past_vec_x.a)      # the stale Jacobi state travels too
    with pytest.raises(ValueError, match="objects"):
        m.load_checkpoint(path, fresh[:1])
    # an explicit checkpoint restored into an implicit object leaves vel_g / past_x alone
    m.save_checkpoint(path, [_Obj(7, False, 3)], 1, 0.0)
    target = _Obj(7, True, 4)
    before = target.particles.vel_g.a.copy()
    m.load_checkpoint(path, [target])
    assert np.array_equal(target.particles.vel_g.a, before)


def test_read_config_failure_prints_and_exits_3(tmp_path, capsys):
    import utils
    bad = tmp_path / "broken.json"
    bad.write_text("{ not json")
    with pytest.raises(SystemExit) as e:
        utils.read_config(str(bad))
    assert e.value.code == 3 and "Parsing config file error" in capsys.readouterr().out     # utils.py:29-32
    with pytest.raises(SystemExit):
        utils.read_config(str(tmp_path / "missing.json"))


def test_sys_init_fills_constants_like_the_reference():
    import constants
    import utils
    cfg = utils.read_config(os.path.join(ROOT, "finite-element-method_b200", "config", "demo_3d_implicit_blob.json"))
    utils.sys_init(cfg)
    assert constants.dim == 3 and constants.use_explicit_method is False and constants.delta_time == cfg["delta_time"]
    assert constants.implicit_method == constants.CONJUGATE_GRADIENT_METHOD == 1 and constants.JACOBIN_METHOD == 0
    assert constants.vec(2.0).shape == (3,) and constants.mat(0.0).shape == (3, 3) and constants.index(0).shape == (4,)
    assert constants.vec(1.0).dtype == np.float32 and constants.index(0).dtype == np.int32


def test_jacobi_order_selects_the_solver_method(monkeypatch):
    """implicit_method 0 runs the simultaneous Jacobi unless the config (extension key `jacobi_order`) or the
    environment asks for the reference's serial sweep order (femb_implicit_solve method 2)."""
    import constants
    import utils
    from solver import implicit

    class Obj:
        def _implicit_solve(self, dt, method, pre):
            self.call = (dt, method, pre)
            return 3, 1e-6

    base = {"dim": 2, "delta_time": 5e-4, "g_dir": [0, -1], "implicit_method": 0, "preconditioned": 0}
    monkeypatch.delenv("FEMB_JACOBI_ORDER", raising=False)
    for cfg, env, want in [({}, None, 0), ({"jacobi_order": "reference"}, None, 2), ({}, "reference", 2),
                           ({"jacobi_order": "reference"}, "simultaneous", 0),
                           ({"implicit_method": 1, "jacobi_order": "reference", "preconditioned": 1}, None, 1)]:
        utils.sys_init(dict(base, **cfg))
        if env is None:
            monkeypatch.delenv("FEMB_JACOBI_ORDER", raising=False)
        else:
            monkeypatch.setenv("FEMB_JACOBI_ORDER", env)
        obj = Obj()
        implicit.implicit_solver_neo_hookean(obj)
        assert obj.call == (5e-4, want, 1 if cfg.get("preconditioned") == 1 else 0)
        assert obj.last_solve == {"iters": 3, "resid": 1e-6}
    utils.sys_init(dict(base, jacobi_order="gauss"))
    monkeypatch.delenv("FEMB_JACOBI_ORDER", raising=False)
    with pytest.raises(ValueError):
        implicit.implicit_solver_neo_hookean(Obj())
    utils.sys_init(base)
    assert constants.jacobi_order == "simultaneous"
