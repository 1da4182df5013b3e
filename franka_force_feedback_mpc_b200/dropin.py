"""Run the reference's runners unchanged on the B200 engine.

`src/run/run_classical.py` / `run_force_feedback.py` import, unconditionally,

    import pinocchio as pin                                                   (run_classical.py:11)
    from src.mpc.crocoddyl_classical import ClassicalCrocoddylMPC, ClassicalMPCConfig      (:19)

and use `pin` only for the kinematic alignment check (`pin.forwardKinematics(mpc.model, mpc.data, q, v)`,
`pin.updateFramePlacements(...)`, `mpc.data.oMf[mpc.ee_fid].translation / .rotation`, :153-158).  `install()` puts into
`sys.modules`

  * `src.mpc.crocoddyl_classical` / `src.mpc.crocoddyl_force_feedback`  -> this package's controller classes and config
    dataclasses (same names), so the runner's import line binds the engine-backed controllers;
  * `pinocchio` (only when the real one is not importable) -> the handful of functions above over `PandaModel`;
  * `crocoddyl` / `example_robot_data` (only when absent) -> empty placeholders: nothing on the runner's path touches them
    once the controllers are replaced.

MuJoCo is NOT shimmed: it is the plant (`src/sim/franka_sim.py`), outside the hot path, and unavailable on this box.

    python -m franka_force_feedback_mpc_b200.dropin src/run/run_classical.py --scenario flat --time 20 --no-viewer

runs the given script (from the reference checkout, cwd = its root) with the substitutions installed.
"""
from __future__ import annotations

import importlib.util
import runpy
import sys
import types

import numpy as np


class _Placement:
    """The two members of `pinocchio.SE3` the runners read."""

    def __init__(self, translation: np.ndarray, rotation: np.ndarray):
        self.translation = np.asarray(translation, dtype=float)
        self.rotation = np.asarray(rotation, dtype=float)


class _KinData:
    """Stand-in for `pinocchio.Data`: frame placements of the last forwardKinematics call."""

    def __init__(self):
        self.q = None
        self.oMf = {}


def _pinocchio_shim() -> types.ModuleType:
    pin = types.ModuleType("pinocchio")
    pin.__doc__ = "franka_force_feedback_mpc_b200 stand-in: forward kinematics of PandaModel behind pinocchio's call names"
    pin.LOCAL, pin.WORLD, pin.LOCAL_WORLD_ALIGNED = 0, 1, 2

    def forwardKinematics(model, data, q, v=None, a=None):
        data.q = np.asarray(q, dtype=float).reshape(-1)[: model.nv].copy()

    def updateFramePlacements(model, data):
        if data.q is None:
            raise RuntimeError("updateFramePlacements before forwardKinematics")
        p, R = model.fk(data.q)
        data.oMf = {0: _Placement(p, R)}     # PandaModel carries one operational frame (panda_link8), id 0

    def framesForwardKinematics(model, data, q):
        forwardKinematics(model, data, q)
        updateFramePlacements(model, data)

    pin.forwardKinematics, pin.updateFramePlacements, pin.framesForwardKinematics = forwardKinematics, updateFramePlacements, framesForwardKinematics
    pin.SE3 = _Placement
    return pin


def _controller_module(name: str, **members) -> types.ModuleType:
    mod = types.ModuleType(name)
    mod.__doc__ = "engine-backed replacement installed by franka_force_feedback_mpc_b200.dropin"
    for k, v in members.items():
        setattr(mod, k, v)
    return mod


def _importable(name: str) -> bool:
    if name in sys.modules:
        return True
    try:
        return importlib.util.find_spec(name) is not None
    except (ImportError, ValueError):
        return False


def install(solver=None) -> dict:
    """Install the substitutions; returns {module name: module} of what was put into `sys.modules`.  `solver`
    (optional) is forwarded to every controller the runner builds (tests inject a CPU solver; the default is the CUDA
    engine, which fails loudly without a GPU)."""
    from . import config as _config, controllers as _ctl

    done = {}
    if not _importable("pinocchio"):
        done["pinocchio"] = sys.modules["pinocchio"] = _pinocchio_shim()
    for name in ("crocoddyl", "example_robot_data"):
        if not _importable(name):
            done[name] = sys.modules[name] = types.ModuleType(name)

    def bind(cls):
        if solver is None:
            return cls

        class Bound(cls):
            def __init__(self, *a, **kw):
                kw.setdefault("solver", solver)
                super().__init__(*a, **kw)

        Bound.__name__, Bound.__qualname__ = cls.__name__, cls.__qualname__
        return Bound

    # controllers expose `data` for the runner's alignment check: give every new controller a kinematics scratch
    def with_data(cls):
        class WithData(cls):
            def __init__(self, *a, **kw):
                super().__init__(*a, **kw)
                if self.data is None:
                    self.data = _KinData()

        WithData.__name__, WithData.__qualname__ = cls.__name__, cls.__qualname__
        return WithData

    classical = with_data(bind(_ctl.ClassicalCrocoddylMPC))
    ff = with_data(bind(_ctl.ForceFeedbackCrocoddylMPC))
    for pkg in ("src", "src.mpc"):
        if pkg not in sys.modules and not _importable(pkg):
            done[pkg] = sys.modules[pkg] = types.ModuleType(pkg)
            sys.modules[pkg].__path__ = []
    done["src.mpc.crocoddyl_classical"] = sys.modules["src.mpc.crocoddyl_classical"] = _controller_module(
        "src.mpc.crocoddyl_classical", ClassicalCrocoddylMPC=classical, ClassicalMPCConfig=_config.ClassicalMPCConfig)
    done["src.mpc.crocoddyl_force_feedback"] = sys.modules["src.mpc.crocoddyl_force_feedback"] = _controller_module(
        "src.mpc.crocoddyl_force_feedback", ForceFeedbackCrocoddylMPC=ff, ForceFeedbackMPCConfig=_config.ForceFeedbackMPCConfig)
    return done


def main(argv=None) -> None:
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv:
        raise SystemExit("usage: python -m franka_force_feedback_mpc_b200.dropin <runner.py> [runner arguments]")
    install()
    sys.argv = argv
    if "" not in sys.path and "." not in sys.path:
        sys.path.insert(0, "")
    runpy.run_path(argv[0], run_name="__main__")


if __name__ == "__main__":
    main()
