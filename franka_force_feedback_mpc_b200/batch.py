"""Host-side containers for the batched solve call: numpy mirrors of ``ffmpc_batch_in`` /
``ffmpc_batch_out`` / ``ffmpc_node_out`` (include/ffmpc.h)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def f64(a, shape=None):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


class BatchInputs:
    """Inputs of one batched solve (host memory). Shapes as in ``ffmpc_batch_in``."""

    def __init__(self, nx: int, N: int, mode, x0, p_ref, v_ref, x_reg_ref, tau_ref=None, xs_init=None, us_init=None):
        self.nx, self.N = int(nx), int(N)
        self.mode = np.ascontiguousarray(mode, dtype=np.int32).reshape(-1)
        B = self.B = self.mode.size
        self.x0 = f64(x0, (B, nx))
        self.p_ref = f64(p_ref, (B, N + 1, 3))
        self.v_ref = f64(v_ref, (B, N + 1, 3))
        self.x_reg_ref = f64(x_reg_ref, (B, 14))
        self.tau_ref = f64(tau_ref, (B, 7))
        self.xs_init = f64(xs_init, (B, N + 1, nx))
        self.us_init = f64(us_init, (B, N, 7))

    @classmethod
    def from_samples(cls, s: dict, nx: int, N: int, **kw):
        return cls(nx, N, s["mode"], s["x0"], s["p_ref"], s["v_ref"], s["x_reg_ref"], **kw)

    def to_abi(self) -> _abi.BatchIn:
        s = _abi.BatchIn()
        s.B = self.B
        for name in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref", "tau_ref", "xs_init", "us_init"):
            setattr(s, name, ptr(getattr(self, name)))
        return s

    def slice(self, idx):
        idx = np.asarray(idx)
        g = lambda a: None if a is None else a[idx]
        return BatchInputs(self.nx, self.N, self.mode[idx], self.x0[idx], self.p_ref[idx], self.v_ref[idx],
                           self.x_reg_ref[idx], g(self.tau_ref), g(self.xs_init), g(self.us_init))

    def h2d_bytes(self) -> int:
        return sum(a.nbytes for a in (self.mode, self.x0, self.p_ref, self.v_ref, self.x_reg_ref, self.tau_ref,
                                      self.xs_init, self.us_init) if a is not None)


OUTPUT_NAMES = tuple(n for n, _ in _abi.BatchOut._fields_)


def alloc_outputs(B: int, N: int, nx: int, want=OUTPUT_NAMES) -> dict:
    shapes = dict(xs=(B, N + 1, nx), us=(B, N, 7), K=(B, N, 7, nx), k=(B, N, 7), cost=(B,), stop=(B,),
                  gaps=(B, N + 1, nx), f_pred=(B, 2, 3), iters=(B,), ok=(B,))
    return {n: np.zeros(shapes[n], dtype=np.int32 if n in ("iters", "ok") else np.float64) for n in want}


def outputs_abi(out: dict) -> _abi.BatchOut:
    s = _abi.BatchOut()
    for name in OUTPUT_NAMES:
        setattr(s, name, ptr(out.get(name)))
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
        setattr(s, name, ptr(out.get(name)))
    return s
