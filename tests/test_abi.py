"""The C-ABI drop-in boundary (include/ffmpc.h): the product library loads, exports every declared
symbol, the ctypes mirror matches the compiled struct layout, and -- on a box without a GPU -- the
engine FAILS LOUDLY instead of falling back to a CPU path.  No compute calls here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "ffmpc.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(ffmpc_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_the_seam():
    syms = _declared_symbols()
    for s in ("ffmpc_create", "ffmpc_destroy", "ffmpc_solve_batch", "ffmpc_solve_batch_host", "ffmpc_node_eval_host",
              "ffmpc_last_error", "ffmpc_default_config", "ffmpc_sync"):
        assert s in syms, s


def test_library_exports_every_declared_symbol(built):
    from franka_force_feedback_mpc_b200 import _abi, engine

    lib = engine.load_library()
    for s in _declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/ffmpc.h but not exported by libffmpc_b200.so"
    assert set(_abi.EXPORTED_SYMBOLS) >= set(_declared_symbols())
    assert b"sm_100a" in lib.ffmpc_version()


def test_ctypes_mirror_matches_compiled_structs(built):
    from franka_force_feedback_mpc_b200 import _abi, engine

    lib = engine.load_library()
    sizes = (C.c_int * 10)()
    lib.ffmpc_abi_sizes(sizes)
    assert sizes[0] == C.sizeof(_abi.Model)
    assert sizes[1] == C.sizeof(_abi.Config)
    assert sizes[2] == C.sizeof(_abi.BatchIn)
    assert sizes[3] == C.sizeof(_abi.BatchOut)
    assert sizes[4] == C.sizeof(_abi.NodeOut)
    assert sizes[5] == C.sizeof(_abi.TickPolicy)
    assert sizes[6] == C.sizeof(_abi.TickIn)
    assert sizes[7] == C.sizeof(_abi.TickOut)
    assert sizes[8] == C.sizeof(_abi.Options)
    assert sizes[9] == C.sizeof(_abi.Info)


def test_default_config_matches_oracle_default(built, orc):
    """Both implementations of the header fill the same defaults (crocoddyl solver constants + dataclass defaults)."""
    from franka_force_feedback_mpc_b200 import _abi, engine

    lib = engine.load_library()
    for variant in (_abi.VARIANT_CLASSICAL, _abi.VARIANT_FF):
        a = _abi.Config()
        lib.ffmpc_default_config(C.byref(a), variant)
        b = orc.default_config(variant)
        assert bytes(a) == bytes(b)
        assert a.th_stop == 5e-5 and a.n_alphas == 10 and a.reg_min == 1e-9 and a.boxqp_maxiter == 100


def test_bad_arguments_return_codes(built, model):
    from franka_force_feedback_mpc_b200 import _abi, config, engine, panda

    lib = engine.load_library()
    cfg = config.to_abi(config.benchmark_ff_config(horizon=10), model, panda.Q_NEUTRAL)
    h = C.c_void_p()
    assert lib.ffmpc_create(None, C.byref(model.to_abi()), 0, 4, C.byref(h)) == _abi.ERR_ARG
    bad = _abi.Config.from_buffer_copy(bytes(cfg))
    bad.contact_nc = 2  # the reference raises ValueError for an unknown contact model (crocoddyl_classical.py:141-142)
    assert lib.ffmpc_create(C.byref(bad), C.byref(model.to_abi()), 0, 4, C.byref(h)) == _abi.ERR_ARG
    assert b"contact_nc" in lib.ffmpc_last_error()
    bad = _abi.Config.from_buffer_copy(bytes(cfg))
    bad.horizon = 0
    assert lib.ffmpc_create(C.byref(bad), C.byref(model.to_abi()), 0, 4, C.byref(h)) == _abi.ERR_ARG


def test_no_cpu_fallback_without_a_device(built, model):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from franka_force_feedback_mpc_b200 import _abi, config, engine, panda

    cfg = config.to_abi(config.benchmark_ff_config(horizon=10), model, panda.Q_NEUTRAL)
    with pytest.raises(engine.FFMPCError) as e:
        engine.FFMPCEngine(cfg, model.to_abi(), device=0, max_batch=4)
    assert "no CPU fallback" in str(e.value)
    lib = engine.load_library()
    tf = C.c_double()
    assert lib.ffmpc_dfma_peak(0, C.byref(tf), None) == _abi.ERR_NO_DEVICE


def test_product_package_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under the product package may import or load it."""
    pkg = os.path.join(ROOT, "franka_force_feedback_mpc_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "libffmpc_oracle" not in txt, f
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "ffmpc_oracle_" not in txt, f
