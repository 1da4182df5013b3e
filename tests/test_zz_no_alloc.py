"""ffmpc_node_eval_static (include/ffmpc.h): the knot evaluation with its scratch in a static device array, no allocation
inside the library.  No counterpart in the reference.  Checked through the
C ABI against the existing entry point ffmpc_node_eval_host on the same inputs, two calls in a row (the second call finds
the first one's records in the shared scratch), with different sizes and variants so that stale contents would show.
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from ffmpc_cases import make_case, rel_err

NODE_KEYS = ("xnext", "cost", "force", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu")
TOL = 1e-10   # the bar test_node_calc_and_calcdiff_match_oracle uses for these arrays (same kernels: equal in practice)


def _points(inp, ff, B, N, per, seed):
    rng = np.random.default_rng(seed)
    inst = np.repeat(np.arange(B), per).astype(np.int32)
    knot = np.tile(([0, 7, N - 1, N] * per)[:per], B).astype(np.int32)
    x = inp.x0[inst] + rng.normal(0, 0.03, (inst.size, inp.nx))
    u = rng.normal(0, 3.0, (inst.size, 7)) + (inp.x0[inst][:, 14:21] if ff else 0.0)
    u[::5] *= 40.0
    return inst, knot, x, u


def _compare(engine, cfg, am, inp, pts, B):
    eng = engine.FFMPCEngine(cfg, am, device=0, max_batch=B)
    ref = eng.node_eval(inp, *pts)
    eng.close()
    got = engine.node_eval_static(cfg, am, inp, *pts, device=0)
    contact = inp.mode[pts[0]] == 1
    for k in NODE_KEYS:
        a, b = got[k], ref[k]
        if k == "force":   # undefined (not written) in free mode
            a, b = a[contact], b[contact]
            if a.size == 0:
                continue
        assert np.isfinite(a).all(), k
        assert rel_err(a, b) <= TOL, (k, rel_err(a, b))


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_node_eval_host_twice(built):
    from franka_force_feedback_mpc_b200 import _abi, engine

    N = 20
    # call 1: force-feedback, point3d, the full 512 points (fills the whole record scratch)
    B1 = 128
    cfg1, am, inp1 = make_case(True, 3, B1, N, 11, 0.5)
    _compare(engine, cfg1, am, inp1, _points(inp1, True, B1, N, 4, 3), B1)
    # call 2, right after: classical, normal_1d, fewer points -- finds call 1's records in the shared array
    B2 = 48
    cfg2, am2, inp2 = make_case(False, 1, B2, N, 12, 0.75)
    _compare(engine, cfg2, am2, inp2, _points(inp2, False, B2, N, 3, 4), B2)
    # and the first case again: same answer as before
    _compare(engine, cfg1, am, inp1, _points(inp1, True, B1, N, 4, 3), B1)


@pytest.mark.gpu
def test_static_scratch_entry_point_rejects_sizes_beyond_the_maximum(built):
    from franka_force_feedback_mpc_b200 import _abi, engine

    lib = engine.load_library()
    N = 10
    cfg, am, inp = make_case(True, 1, 8, N, 5, 0.5)
    s_in, s_out = inp.to_abi(), _abi.NodeOut()
    for name, _ in _abi.NodeOut._fields_:
        setattr(s_out, name, 8)   # never dereferenced: the size check comes first
    dummy = C.c_void_p(8)
    rc = lib.ffmpc_node_eval_static(C.byref(cfg), C.byref(am), 0, C.byref(s_in), _abi.STATIC_MAX_POINTS + 1, dummy, dummy, dummy, dummy,
                                    C.byref(s_out), None)
    assert rc == _abi.ERR_CAPACITY
    assert b"FFMPC_STATIC_MAX_POINTS" in lib.ffmpc_last_error()
    s_in.B = _abi.STATIC_MAX_BATCH + 1
    rc = lib.ffmpc_node_eval_static(C.byref(cfg), C.byref(am), 0, C.byref(s_in), 4, dummy, dummy, dummy, dummy, C.byref(s_out), None)
    assert rc == _abi.ERR_CAPACITY
    assert b"FFMPC_STATIC_MAX_BATCH" in lib.ffmpc_last_error()
