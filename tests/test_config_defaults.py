"""One table of defaults, three holders: the reference's dataclasses, their mirror in config.py, and ffmpc_default_config in
the C library (plus the oracle's copy, checked against the library in tests/test_abi.py).  tests/golden/config_defaults.json
is dumped from the REFERENCE's own dataclasses by tools/make_golden.py::gen_defaults (crocoddyl_classical.py:12-110,
crocoddyl_force_feedback.py:12-146)."""
import ctypes as C
import dataclasses
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "config_defaults.json")))


@pytest.mark.parametrize("name", ["ClassicalMPCConfig", "ForceFeedbackMPCConfig"])
def test_config_mirror_has_the_reference_defaults(name):
    from franka_force_feedback_mpc_b200 import config

    mine = getattr(config, name)()
    ref = GOLD[name]
    fields = {f.name for f in dataclasses.fields(mine)}
    assert fields == set(ref), (sorted(fields - set(ref)), sorted(set(ref) - fields))
    for k, v in ref.items():
        got = getattr(mine, k)
        if isinstance(v, list):
            assert np.array_equal(np.asarray(got, float), np.asarray(v, float)), k
        else:
            assert got == v and type(got) is type(v), (k, got, v)


@pytest.mark.parametrize("ff", [False, True])
def test_c_default_config_is_the_lowered_dataclass_default(built, ff):
    """ffmpc_default_config(variant) == config.to_abi(<dataclass defaults>) on every field the dataclass determines
    (the solver constants of Crocoddyl -- reg_*, th_*, n_alphas, boxqp_* -- have no dataclass field: both sides hold the
    library's own table)."""
    from franka_force_feedback_mpc_b200 import _abi, config, engine, panda

    lib = engine.load_library()
    c = _abi.Config()
    lib.ffmpc_default_config(C.byref(c), _abi.VARIANT_FF if ff else _abi.VARIANT_CLASSICAL)
    m = panda.default_model()
    d = config.to_abi((config.ForceFeedbackMPCConfig if ff else config.ClassicalMPCConfig)(), m, panda.Q_NEUTRAL)
    derived = {"q_soft_ref", "q_soft_lb", "q_soft_ub", "R_des", "w_q_soft_limits"}   # depend on the model / nominal pose: lowered per handle
    diff = []
    for name, typ in _abi.Config._fields_:
        if name in derived or name.startswith("reserved"):
            continue
        a, b = getattr(c, name), getattr(d, name)
        if hasattr(a, "__len__"):
            if not np.array_equal(np.ctypeslib.as_array(a), np.ctypeslib.as_array(b)):
                diff.append(name)
        elif a != b:
            diff.append((name, a, b))
    assert not diff, diff
