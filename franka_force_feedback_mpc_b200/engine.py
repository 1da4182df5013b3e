"""ctypes loader and thin Python front-end of ``libffmpc_b200.so`` (the CUDA engine).

There is no fallback: if the shared library is missing, or no CUDA device is visible, construction
raises.  PyTorch is used only for device buffers / streams (``solve_device``); numpy for the host API.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

from . import _abi
from .batch import (BatchInputs, OUTPUT_NAMES, alloc_node_outputs, alloc_outputs, node_outputs_abi, outputs_abi, ptr)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FFMPC_LIB") or os.path.join(_HERE, "csrc", "libffmpc_b200.so")  # FFMPC_LIB: A/B builds of the same engine
_lib = None


class FFMPCError(RuntimeError):
    pass


def _declare(lib):
    lib.ffmpc_last_error.restype = C.c_char_p
    lib.ffmpc_version.restype = C.c_char_p
    lib.ffmpc_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    lib.ffmpc_create_ex.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
    lib.ffmpc_default_options.argtypes = [C.c_void_p]
    lib.ffmpc_default_options.restype = None
    lib.ffmpc_destroy.argtypes = [C.c_void_p]
    lib.ffmpc_destroy.restype = None
    lib.ffmpc_solve_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.ffmpc_solve_batch_host.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.ffmpc_solve_batch_host_async.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.ffmpc_wait.argtypes = [C.c_void_p]
    lib.ffmpc_node_eval_host.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ffmpc_tick_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.ffmpc_tick_batch_host.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.ffmpc_last_timing.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int)]
    lib.ffmpc_sync.argtypes = [C.c_void_p]
    lib.ffmpc_get_info.argtypes = [C.c_void_p, C.c_void_p]
    lib.ffmpc_set_profiling.argtypes = [C.c_void_p, C.c_int]
    if hasattr(lib, "ffmpc_node_eval_static"):  # absent from the test-only host emulator build
        lib.ffmpc_node_eval_static.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p]
    if hasattr(lib, "ffmpc_dfma_peak"):  # absent from the test-only host emulator build
        lib.ffmpc_dfma_peak.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.ffmpc_get_profile.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int)]
    return lib


def load_library(path: Optional[str] = None):
    """Load the CUDA engine. Raises ``FFMPCError`` if it has not been built (``__graft_entry__.build()``)."""
    global _lib
    if path is None:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise FFMPCError(
                f"{LIB_PATH} not found: build the CUDA extension first (python -c 'import __graft_entry__ as g; g.build()'). "
                "The engine has no CPU fallback."
            )
        _lib = _declare(C.CDLL(LIB_PATH))
        return _lib
    return _declare(C.CDLL(path))


def dfma_peak_tflops(device: int = 0) -> float:
    """Measured fp64 FMA-pipe peak of `device` in TFLOP/s (micro-benchmark inside the engine library)."""
    lib = load_library()
    tf, ms = C.c_double(), C.c_double()
    rc = lib.ffmpc_dfma_peak(int(device), C.byref(tf), C.byref(ms))
    if rc != 0:
        raise FFMPCError(f"ffmpc_dfma_peak failed with code {rc}")
    return tf.value


def node_eval_static(cfg, model, inp: BatchInputs, inst, knot, x, u, device: int = 0, stream=None) -> dict:
    """``ffmpc_node_eval_static``: the knot evaluation without any allocation inside the library (scratch in a static
    device array; limits ``_abi.STATIC_MAX_POINTS`` / ``STATIC_MAX_BATCH``).  The buffers the CALLER owns are torch device
    tensors made here; the result comes back as numpy arrays.  Calls that overlap on the GPU race on the shared scratch."""
    import torch

    lib = load_library()
    dev = torch.device("cuda", device)
    nx = 21 if cfg.variant == _abi.VARIANT_FF else 14
    M = int(np.asarray(inst).size)
    up = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a, dtype=dt)).to(dev)
    keep = dict(mode=up(inp.mode, np.int32), x0=up(inp.x0, np.float64), p_ref=up(inp.p_ref, np.float64), v_ref=up(inp.v_ref, np.float64),
                x_reg_ref=up(inp.x_reg_ref, np.float64))
    for name in ("tau_ref",):
        v = getattr(inp, name, None)
        if v is not None:
            keep[name] = up(v, np.float64)
    s_in = _abi.BatchIn()
    s_in.B = inp.B
    for name, t in keep.items():
        setattr(s_in, name, t.data_ptr())
    d_inst, d_knot = up(inst, np.int32), up(knot, np.int32)
    d_x, d_u = up(np.asarray(x, dtype=np.float64).reshape(M, nx), np.float64), up(np.asarray(u, dtype=np.float64).reshape(M, 7), np.float64)
    shapes = dict(xnext=(M, nx), cost=(M,), force=(M, 3), Fx=(M, nx, nx), Fu=(M, nx, 7), Lx=(M, nx), Lu=(M, 7), Lxx=(M, nx, nx),
                  Lxu=(M, nx, 7), Luu=(M, 7, 7))
    outs = {k: torch.full(sh, float("nan"), dtype=torch.float64, device=dev) for k, sh in shapes.items()}
    s_out = _abi.NodeOut()
    for k, t in outs.items():
        setattr(s_out, k, t.data_ptr())
    st = stream if stream is not None else torch.cuda.current_stream(dev)
    rc = lib.ffmpc_node_eval_static(C.byref(cfg), C.byref(model), int(device), C.byref(s_in), M, d_inst.data_ptr(), d_knot.data_ptr(),
                                    d_x.data_ptr(), d_u.data_ptr(), C.byref(s_out), C.c_void_p(st.cuda_stream))
    if rc != 0:
        raise FFMPCError(f"ffmpc_node_eval_static failed ({rc}): {lib.ffmpc_last_error().decode()}")
    st.synchronize()
    return {k: t.cpu().numpy() for k, t in outs.items()}


class FFMPCEngine:
    """One handle per GPU (``ffmpc_create_ex``).  ``line_search``: "auto" | "queue" | "spec" | "grouped" | "tail" | "fill"
    (the schedule of step-length groups, ``FFMPC_LS_*``); ``sub_batches``: into how many
    concurrently scheduled sub-batches (own workspace + stream) a batch is split (0 = auto).  Both only schedule work:
    results are bit-identical."""

    def __init__(self, cfg: _abi.Config, model: _abi.Model, device: int = 0, max_batch: int = 1, _lib=None,
                 line_search: str = "auto", sub_batches: int = 0):
        self.lib = _lib if _lib is not None else load_library()
        self.cfg, self.model = cfg, model
        self.N = int(cfg.horizon)
        self.nx = 21 if cfg.variant == _abi.VARIANT_FF else 14
        self.max_batch = int(max_batch)
        self.device = int(device)
        self._h = C.c_void_p()
        opt = _abi.Options()
        self.lib.ffmpc_default_options(C.byref(opt))
        opt.line_search = _abi.LINE_SEARCH_MODES[line_search]
        opt.sub_batches = int(sub_batches)
        rc = self.lib.ffmpc_create_ex(C.byref(cfg), C.byref(model), self.device, self.max_batch, C.byref(opt), C.byref(self._h))
        self._check(rc)

    def _check(self, rc):
        if rc != 0:
            raise FFMPCError(f"ffmpc error {rc}: {self.lib.ffmpc_last_error().decode()}")

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.ffmpc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- host API: numpy in / numpy out, copies inside ------------------------------------------
    def solve_host(self, inp: BatchInputs, max_iters: int, fixed_iters: bool = True, want=OUTPUT_NAMES, out: Optional[dict] = None) -> dict:
        if out is None:
            out = alloc_outputs(inp.B, self.N, self.nx, want)
        s_in, s_out = inp.to_abi(), outputs_abi(out)
        self._check(self.lib.ffmpc_solve_batch_host(self._h, C.byref(s_in), int(max_iters), int(bool(fixed_iters)), C.byref(s_out)))
        return out

    def solve_host_ptrs(self, s_in: _abi.BatchIn, s_out: _abi.BatchOut, max_iters: int, fixed_iters: bool = True):
        """Same with pre-built pointer structs (e.g. pinned torch tensors), no allocation per call."""
        self._check(self.lib.ffmpc_solve_batch_host(self._h, C.byref(s_in), int(max_iters), int(bool(fixed_iters)), C.byref(s_out)))

    def solve_host_ptrs_async(self, s_in: _abi.BatchIn, s_out: _abi.BatchOut, max_iters: int, fixed_iters: bool = True):
        """``ffmpc_solve_batch_host_async``: returns after enqueueing; buffers must stay valid until ``wait()``."""
        self._check(self.lib.ffmpc_solve_batch_host_async(self._h, C.byref(s_in), int(max_iters), int(bool(fixed_iters)), C.byref(s_out)))

    def wait(self):
        self._check(self.lib.ffmpc_wait(self._h))

    # ---- device API: torch CUDA tensors as buffers, stream-ordered, no sync ----------------------
    def solve_device(self, t_in: dict, max_iters: int, fixed_iters: bool = True, t_out: Optional[dict] = None, stream=None) -> dict:
        import torch

        B = int(t_in["mode"].shape[0])
        dev = t_in["x0"].device
        if t_out is None:
            shapes = dict(xs=(B, self.N + 1, self.nx), us=(B, self.N, 7), K=(B, self.N, 7, self.nx), k=(B, self.N, 7),
                          cost=(B,), stop=(B,), gaps=(B, self.N + 1, self.nx), f_pred=(B, 2, 3), iters=(B,), ok=(B,))
            t_out = {n: torch.empty(shapes[n], device=dev, dtype=torch.int32 if n in ("iters", "ok") else torch.float64)
                     for n in OUTPUT_NAMES}
        s_in = _abi.BatchIn()
        s_in.B = B
        for name in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref", "tau_ref", "xs_init", "us_init"):
            t = t_in.get(name)
            if t is not None:
                assert t.is_cuda and t.is_contiguous(), name
                assert t.dtype == (torch.int32 if name == "mode" else torch.float64), name
            setattr(s_in, name, None if t is None else t.data_ptr())
        s_out = _abi.BatchOut()
        for name in OUTPUT_NAMES:
            t = t_out.get(name)
            setattr(s_out, name, None if t is None else t.data_ptr())
        if stream is None:
            # order the solve against the producers / consumers of these tensors: torch's current stream on their device
            # (a NULL stream would mean the handle's own stream, ffmpc.h)
            stream = torch.cuda.current_stream(dev)
        sp = C.c_void_p(stream.cuda_stream)
        self._check(self.lib.ffmpc_solve_batch(self._h, C.byref(s_in), int(max_iters), int(bool(fixed_iters)), C.byref(s_out), sp))
        return t_out

    # ---- receding-horizon tick for the whole batch (ffmpc_tick_batch_host): shift, solve, policy, guard, filter ----
    def tick_host(self, mode, x_now, p_ref, v_ref, x_reg_ref, tau_bias, policy: _abi.TickPolicy, max_iters: int,
                  fixed_iters: bool = False, reset: bool = False, want_plan: bool = False, tau_ref=None) -> dict:
        mode = np.ascontiguousarray(mode, dtype=np.int32).reshape(-1)
        B, N, nx = mode.size, self.N, self.nx
        arrs = dict(x_now=(B, nx), p_ref=(B, N + 1, 3), v_ref=(B, N + 1, 3), x_reg_ref=(B, 14), tau_bias=(B, 7))
        vals = dict(x_now=x_now, p_ref=p_ref, v_ref=v_ref, x_reg_ref=x_reg_ref, tau_bias=tau_bias)
        if tau_ref is not None:
            arrs["tau_ref"], vals["tau_ref"] = (B, 7), tau_ref
        s_in = _abi.TickIn()
        s_in.B, s_in.reset, s_in.mode = B, int(bool(reset)), ptr(mode)
        keep = [mode]
        for k, shp in arrs.items():
            a = np.ascontiguousarray(vals[k], dtype=np.float64).reshape(shp)
            keep.append(a)
            setattr(s_in, k, ptr(a))
        out = dict(tau_cmd=np.zeros((B, 7)), tau_raw=np.zeros((B, 7)), cost=np.zeros(B), fn_pred=np.zeros(B),
                   iters=np.zeros(B, np.int32), ok=np.zeros(B, np.int32), unstable=np.zeros(B, np.int32))
        if want_plan:
            out["xs"], out["us"] = np.zeros((B, N + 1, nx)), np.zeros((B, N, 7))
        s_out = _abi.TickOut()
        for k, a in out.items():
            setattr(s_out, k, ptr(a))
        self._check(self.lib.ffmpc_tick_batch_host(self._h, C.byref(s_in), C.byref(policy), int(max_iters), int(bool(fixed_iters)), C.byref(s_out)))
        return out

    def info(self) -> dict:
        i = _abi.Info()
        self._check(self.lib.ffmpc_get_info(self._h, C.byref(i)))
        return dict(sub_batches=i.sub_batches, sub_batch_capacity=i.sub_batch_capacity, trial_slots=i.trial_slots,
                    line_search={v: k for k, v in _abi.LINE_SEARCH_MODES.items()}[i.line_search])

    def sync(self):
        self._check(self.lib.ffmpc_sync(self._h))

    def set_profiling(self, on: bool):
        self._check(self.lib.ffmpc_set_profiling(self._h, int(bool(on))))

    def get_profile(self) -> dict:
        """{family: (device ms, kernel launches)} of the last solve (waits for it)."""
        ms = (C.c_double * _abi.NFAM)()
        nl = (C.c_int * _abi.NFAM)()
        self._check(self.lib.ffmpc_get_profile(self._h, ms, nl))
        return {name: (ms[i], nl[i]) for i, name in enumerate(_abi.FAMILY_NAMES)}

    def last_timing(self):
        ms, n = C.c_double(), C.c_int()
        self._check(self.lib.ffmpc_last_timing(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    # ---- ActionModel.calc / calcDiff at arbitrary points (parity-test hook) ----------------------
    def node_eval(self, inp: BatchInputs, inst, knot, x, u) -> dict:
        inst = np.ascontiguousarray(inst, dtype=np.int32)
        knot = np.ascontiguousarray(knot, dtype=np.int32)
        M = inst.size
        x = np.ascontiguousarray(x, dtype=np.float64).reshape(M, self.nx)
        u = np.ascontiguousarray(u, dtype=np.float64).reshape(M, 7)
        out = alloc_node_outputs(M, self.nx)
        s_in, s_out = inp.to_abi(), node_outputs_abi(out)
        self._check(self.lib.ffmpc_node_eval_host(self._h, C.byref(s_in), M, ptr(inst), ptr(knot), ptr(x), ptr(u), C.byref(s_out)))
        return out
