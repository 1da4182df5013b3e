"""numpy/ctypes front-end of the CPU ORACLE (``oracle/libffmpc_oracle.so``).

TEST INFRASTRUCTURE ONLY: may be imported from ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` -- never from the product package.
See the header of ``ffmpc_oracle.cpp`` for what the oracle restates and how it is pinned.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional

import numpy as np

from franka_force_feedback_mpc_b200 import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libffmpc_oracle.so")
_FAST_PATH = os.path.join(_HERE, "libffmpc_oracle_fast.so")
_QUAD_PATH = os.path.join(_HERE, "libffmpc_oracle_quad.so")
_lib = None
_fast = None
_quad = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "ffmpc_oracle.cpp")
    hdr = os.path.join(_HERE, "..", "include", "ffmpc.h")
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.exists(p) and os.path.getmtime(p) > os.path.getmtime(_LIB_PATH) for p in (src, hdr)
    )
    stale = stale or not os.path.exists(_FAST_PATH) or os.path.getmtime(_FAST_PATH) < os.path.getmtime(_LIB_PATH)
    stale = stale or not os.path.exists(_QUAD_PATH) or os.path.getmtime(_QUAD_PATH) < os.path.getmtime(_LIB_PATH)
    cnt = os.path.join(_HERE, "libffmpc_oracle_count.so")
    stale = stale or not os.path.exists(cnt) or os.path.getmtime(cnt) < os.path.getmtime(_LIB_PATH)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "all"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.ffmpc_oracle_last_error.restype = C.c_char_p
    return _lib


from franka_force_feedback_mpc_b200.batch import (  # noqa: E402,F401  (shared host-side containers)
    BatchInputs, alloc_node_outputs, alloc_outputs, node_outputs_abi, outputs_abi, f64 as _f64, ptr as _ptr)


def quad_lib():
    """The same oracle source with every scalar in IEEE binary128 (libquadmath): the higher-precision arbiter of the
    parity tests.  fp64 at the C boundary, ~50x slower than the fp64 build.  Use as ``solve_batch(..., _use=quad_lib())``."""
    global _quad
    if _quad is None:
        build()
        _quad = C.CDLL(_QUAD_PATH)
        _quad.ffmpc_oracle_last_error.restype = C.c_char_p
    return _quad


def count_flops(cfg, model, inp, max_iters: int, fixed_iters: bool = True, nthreads: int = 1) -> float:
    """Scalar operations (add / sub / mul / div / sqrt / trig = 1 each) the oracle executes per solve of this batch:
    the same source on a counting scalar (libffmpc_oracle_count.so).  It counts THIS restatement (forward-mode duals,
    dense Riccati products), not the engine."""
    build()
    L = C.CDLL(os.path.join(_HERE, "libffmpc_oracle_count.so"))
    L.ffmpc_oracle_flops.restype = C.c_longlong
    L.ffmpc_oracle_flops(1)
    solve_batch(cfg, model, inp, max_iters, fixed_iters, nthreads=nthreads, _use=L)
    return L.ffmpc_oracle_flops(1) / float(inp.B)


def _cpu_has(*flags) -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    have = set(line.split(":", 1)[1].split())
                    return all(fl in have for fl in flags)
    except OSError:
        pass
    return False


def timing_lib():
    """Library for the TIMED cpu_baseline of bench.py: the -O3/AVX2/FMA build of the same source when the
    host CPU supports it, else the plain build.  Returns (ctypes lib, description)."""
    global _fast
    if _fast is None:
        if os.path.exists(_FAST_PATH) and _cpu_has("avx2", "fma"):
            _fast = (C.CDLL(_FAST_PATH), "g++ -O3 -mavx2 -mfma")
            _fast[0].ffmpc_oracle_last_error.restype = C.c_char_p
        else:
            _fast = (lib(), "g++ -O2")
    return _fast


def _check(rc):
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}: {lib().ffmpc_oracle_last_error().decode()}")


def default_config(variant: int) -> _abi.Config:
    c = _abi.Config()
    lib().ffmpc_oracle_default_config(C.byref(c), int(variant))
    return c


def solve_batch(cfg: _abi.Config, model: _abi.Model, inp: BatchInputs, max_iters: int, fixed_iters: bool = True,
                nthreads: int = 1, _use=None) -> dict:
    out = alloc_outputs(inp.B, inp.N, inp.nx)
    s_in, s_out = inp.to_abi(), outputs_abi(out)
    _check((_use or lib()).ffmpc_oracle_solve_batch(C.byref(cfg), C.byref(model), C.byref(s_in), int(max_iters),
                                          int(bool(fixed_iters)), C.byref(s_out), int(nthreads)))
    return out


def node_eval(cfg: _abi.Config, model: _abi.Model, inp: BatchInputs, inst, knot, x, u, _use=None) -> dict:
    inst = np.ascontiguousarray(inst, dtype=np.int32)
    knot = np.ascontiguousarray(knot, dtype=np.int32)
    M = inst.size
    x = _f64(x, (M, inp.nx))
    u = _f64(u, (M, 7))
    out = alloc_node_outputs(M, inp.nx)
    s_in, s_out = inp.to_abi(), node_outputs_abi(out)
    _check((_use or lib()).ffmpc_oracle_node_eval(C.byref(cfg), C.byref(model), C.byref(s_in), M, _ptr(inst), _ptr(knot),
                                        _ptr(x), _ptr(u), C.byref(s_out)))
    return out


def node_calc(cfg: _abi.Config, model: _abi.Model, inp: BatchInputs, inst, knot, x, u):
    inst = np.ascontiguousarray(inst, dtype=np.int32)
    knot = np.ascontiguousarray(knot, dtype=np.int32)
    M = inst.size
    x = _f64(x, (M, inp.nx))
    u = _f64(u, (M, 7))
    xnext, cost, force = np.zeros((M, inp.nx)), np.zeros(M), np.zeros((M, 3))
    s_in = inp.to_abi()
    _check(lib().ffmpc_oracle_node_calc(C.byref(cfg), C.byref(model), C.byref(s_in), M, _ptr(inst), _ptr(knot),
                                        _ptr(x), _ptr(u), _ptr(xnext), _ptr(cost), _ptr(force)))
    return xnext, cost, force


def fk(model: _abi.Model, q):
    q = _f64(q, (7,))
    p, R = np.zeros(3), np.zeros((3, 3))
    lib().ffmpc_oracle_fk(C.byref(model), _ptr(q), _ptr(p), _ptr(R))
    return p, R


def rnea(model: _abi.Model, q, v, a):
    q, v, a = _f64(q, (7,)), _f64(v, (7,)), _f64(a, (7,))
    tau = np.zeros(7)
    lib().ffmpc_oracle_rnea(C.byref(model), _ptr(q), _ptr(v), _ptr(a), _ptr(tau))
    return tau


def crba(model: _abi.Model, q):
    q = _f64(q, (7,))
    M = np.zeros((7, 7))
    lib().ffmpc_oracle_crba(C.byref(model), _ptr(q), _ptr(M))
    return M


def frame_jacobian(model: _abi.Model, q):
    q = _f64(q, (7,))
    J = np.zeros((6, 7))
    lib().ffmpc_oracle_frame_jacobian(C.byref(model), _ptr(q), _ptr(J))
    return J


def rollout_stats(reset: bool = False, _use=None):
    """(rollouts, iterations, per-step-length trial histogram) counted by the oracle's line search since the last reset:
    how many forward passes the FDDP iterations really ran (feeds the FLOP accounting of bench.py)."""
    L = _use if _use is not None else lib()
    r, i = C.c_longlong(), C.c_longlong()
    h = (C.c_longlong * 32)()
    L.ffmpc_oracle_rollout_stats.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    L.ffmpc_oracle_rollout_stats.restype = None
    L.ffmpc_oracle_rollout_stats(C.byref(r), C.byref(i), h, int(bool(reset)))
    return r.value, i.value, list(h)
