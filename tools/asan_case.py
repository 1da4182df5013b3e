"""CPU sanitizer run (compute-sanitizer is closed on this pool): the kernel SOURCES under the fiber
emulator and the oracle, both built with -fsanitize=address,undefined, solving small problems of every instantiation.
    tools/asan.sh   (builds into /tmp, runs this under LD_PRELOAD=libasan)"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from ffmpc_cases import make_case  # noqa: E402
from franka_force_feedback_mpc_b200 import engine  # noqa: E402
from oracle import oracle as orc  # noqa: E402

emu = engine.load_library(os.environ["EMU_ASAN"])
ora = C.CDLL(os.environ["ORACLE_ASAN"])
ora.ffmpc_oracle_last_error.restype = C.c_char_p
B, N, iters = 5, 6, 3
for ff, nc, kw in ((True, 1, {}), (True, 3, dict(w_friction_cone=2.0e2, mu=0.3)), (False, 1, {}), (False, 3, {})):
    cfg, am, inp = make_case(ff, nc, B, N, 7, 0.6, **kw)
    ref = orc.solve_batch(cfg, am, inp, iters, True, nthreads=2, _use=ora)
    for ls in ("queue", "spec"):
        eng = engine.FFMPCEngine(cfg, am, device=0, max_batch=B, _lib=emu, line_search=ls, sub_batches=2)
        out = eng.solve_host(inp, iters, True)
        eng.close()
        assert np.array_equal(out["iters"], ref["iters"]) and np.abs(out["cost"] - ref["cost"]).max() < 1e-8 * np.abs(ref["cost"]).max()
    print("ok", "ff" if ff else "classical", nc, flush=True)
print("asan/ubsan run finished")
