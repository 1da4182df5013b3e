"""Parity table for DESIGN.md section 6: per variant, the error of the CUDA engine (fp64) and of the fp64 oracle against
the SAME oracle source built in IEEE binary128 (oracle/libffmpc_oracle_quad.so, the arbiter), at a fixed iteration count.
Per-instance error = max |a - b| over the array / max |b|; reported: median and max over the instances.

    python tools/arbiter_report.py [--B 64] [--iters 10] [--emu]     (--emu: the CPU emulator build instead of the GPU)
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from ffmpc_cases import make_case  # noqa: E402
from franka_force_feedback_mpc_b200 import engine  # noqa: E402
from oracle import oracle as orc  # noqa: E402

VARIANTS = [(False, 1, 1.0), (False, 1, 0.0), (False, 3, 0.5), (True, 1, 1.0), (True, 1, 0.0), (True, 3, 0.5)]


def inst_rel(a, b):
    a = np.asarray(a, float).reshape(a.shape[0], -1)
    b = np.asarray(b, float).reshape(b.shape[0], -1)
    return np.abs(a - b).max(axis=1) / np.maximum(np.abs(b).max(axis=1), 1e-300)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=64)
    ap.add_argument("--N", type=int, default=50)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--emu", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    lib = engine.load_library(os.path.join(ROOT, "tests", "host_emul", "libffmpc_emu.so")) if a.emu else None
    nt = os.cpu_count() or 1
    rows = []
    for ff, nc, cf in VARIANTS:
        cfg, am, inp = make_case(ff, nc, a.B, a.N, 5, cf)
        eng = engine.FFMPCEngine(cfg, am, device=0, max_batch=a.B, _lib=lib)
        out = eng.solve_host(inp, a.iters, True)
        eng.close()
        ref = orc.solve_batch(cfg, am, inp, a.iters, True, nthreads=nt)
        quad = orc.solve_batch(cfg, am, inp, a.iters, True, nthreads=nt, _use=orc.quad_lib())
        row = dict(variant="ff" if ff else "classical", nc=nc, contact_fraction=cf, B=a.B, N=a.N, iters=a.iters,
                   iters_ok_equal=bool(np.array_equal(out["iters"], ref["iters"]) and np.array_equal(out["ok"], ref["ok"])))
        for k in ("K", "k", "cost", "xs", "us", "gaps"):
            if k == "gaps":
                eg = np.abs(out[k] - quad[k]).reshape(a.B, -1).max(axis=1)
                eo = np.abs(ref[k] - quad[k]).reshape(a.B, -1).max(axis=1)
            else:
                eg, eo = inst_rel(out[k], quad[k]), inst_rel(ref[k], quad[k])
            row[k] = dict(gpu_vs_quad_median=float(np.median(eg)), gpu_vs_quad_max=float(eg.max()),
                          oracle_vs_quad_median=float(np.median(eo)), oracle_vs_quad_max=float(eo.max()))
        rows.append(row)
        print(f"{row['variant']:9s} nc={nc} cf={cf}: iters/ok equal {row['iters_ok_equal']} | " +
              " | ".join(f"{k}: gpu {row[k]['gpu_vs_quad_max']:.1e} oracle {row[k]['oracle_vs_quad_max']:.1e}" for k in ("K", "k", "cost", "gaps")),
              flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
