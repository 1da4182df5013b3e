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
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "ffmpc_oracle.cpp")
    hdr = os.path.join(_HERE, "..", "include", "ffmpc.h")
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.exists(p) and os.path.getmtime(p) > os.path.getmtime(_LIB_PATH) for p in (src, hdr)
    )
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "libffmpc_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.ffmpc_oracle_last_error.restype = C.c_char_p
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64(a, shape=None):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


class BatchInputs:
    """Host-side bundle matching ``ffmpc_batch_in``."""

    def __init__(self, nx: int, N: int, mode, x0, p_ref, v_ref, x_reg_ref, tau_ref=None, xs_init=None, us_init=None):
        self.nx, self.N = int(nx), int(N)
        self.mode = np.ascontiguousarray(mode, dtype=np.int32).reshape(-1)
        B = self.B = self.mode.size
        self.x0 = _f64(x0, (B, nx))
        self.p_ref = _f64(p_ref, (B, N + 1, 3))
        self.v_ref = _f64(v_ref, (B, N + 1, 3))
        self.x_reg_ref = _f64(x_reg_ref, (B, 14))
        self.tau_ref = _f64(tau_ref, (B, 7))
        self.xs_init = _f64(xs_init, (B, N + 1, nx))
        self.us_init = _f64(us_init, (B, N, 7))

    def to_abi(self) -> _abi.BatchIn:
        s = _abi.BatchIn()
        s.B = self.B
        s.mode = _ptr(self.mode)
        s.x0 = _ptr(self.x0)
        s.p_ref = _ptr(self.p_ref)
        s.v_ref = _ptr(self.v_ref)
        s.x_reg_ref = _ptr(self.x_reg_ref)
        s.tau_ref = _ptr(self.tau_ref)
        s.xs_init = _ptr(self.xs_init)
        s.us_init = _ptr(self.us_init)
        return s

    def slice(self, idx):
        idx = np.asarray(idx)
        g = lambda a: None if a is None else a[idx]
        return BatchInputs(self.nx, self.N, self.mode[idx], self.x0[idx], self.p_ref[idx], self.v_ref[idx],
                           self.x_reg_ref[idx], g(self.tau_ref), g(self.xs_init), g(self.us_init))


def alloc_outputs(B: int, N: int, nx: int) -> dict:
    return dict(
        xs=np.zeros((B, N + 1, nx)), us=np.zeros((B, N, 7)), K=np.zeros((B, N, 7, nx)), k=np.zeros((B, N, 7)),
        cost=np.zeros(B), stop=np.zeros(B), gaps=np.zeros((B, N + 1, nx)), f_pred=np.zeros((B, 2, 3)),
        iters=np.zeros(B, dtype=np.int32), ok=np.zeros(B, dtype=np.int32),
    )


def outputs_abi(out: dict) -> _abi.BatchOut:
    s = _abi.BatchOut()
    for name, _ in _abi.BatchOut._fields_:
        setattr(s, name, _ptr(out.get(name)))
    return s


def alloc_node_outputs(M: int, nx: int) -> dict:
    return dict(
        xnext=np.zeros((M, nx)), cost=np.zeros(M), force=np.zeros((M, 3)), Fx=np.zeros((M, nx, nx)),
        Fu=np.zeros((M, nx, 7)), Lx=np.zeros((M, nx)), Lu=np.zeros((M, 7)), Lxx=np.zeros((M, nx, nx)),
        Lxu=np.zeros((M, nx, 7)), Luu=np.zeros((M, 7, 7)),
    )


def node_outputs_abi(out: dict) -> _abi.NodeOut:
    s = _abi.NodeOut()
    for name, _ in _abi.NodeOut._fields_:
        setattr(s, name, _ptr(out.get(name)))
    return s


def _check(rc):
    if rc != 0:
        raise RuntimeError(f"oracle error {rc}: {lib().ffmpc_oracle_last_error().decode()}")


def default_config(variant: int) -> _abi.Config:
    c = _abi.Config()
    lib().ffmpc_oracle_default_config(C.byref(c), int(variant))
    return c


def solve_batch(cfg: _abi.Config, model: _abi.Model, inp: BatchInputs, max_iters: int, fixed_iters: bool = True,
                nthreads: int = 1) -> dict:
    out = alloc_outputs(inp.B, inp.N, inp.nx)
    s_in, s_out = inp.to_abi(), outputs_abi(out)
    _check(lib().ffmpc_oracle_solve_batch(C.byref(cfg), C.byref(model), C.byref(s_in), int(max_iters),
                                          int(bool(fixed_iters)), C.byref(s_out), int(nthreads)))
    return out


def node_eval(cfg: _abi.Config, model: _abi.Model, inp: BatchInputs, inst, knot, x, u) -> dict:
    inst = np.ascontiguousarray(inst, dtype=np.int32)
    knot = np.ascontiguousarray(knot, dtype=np.int32)
    M = inst.size
    x = _f64(x, (M, inp.nx))
    u = _f64(u, (M, 7))
    out = alloc_node_outputs(M, inp.nx)
    s_in, s_out = inp.to_abi(), node_outputs_abi(out)
    _check(lib().ffmpc_oracle_node_eval(C.byref(cfg), C.byref(model), C.byref(s_in), M, _ptr(inst), _ptr(knot),
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
