#!/usr/bin/env python
"""BASELINE.json configs 1 and 2 on this box: one controller, a stream of warm-started receding-horizon ticks.

The reference runs `ClassicalCrocoddylMPC` (benchmark preset, N = 36, `run_classical.py:269-315`) /
`ForceFeedbackCrocoddylMPC` (N = 40, `run_force_feedback.py:272-330`) for 4000 ticks of 5 ms against MuJoCo; Crocoddyl and
MuJoCo are unavailable on this box, so the stream is a kinematic replay of the reference
trajectory: x_now(k) = IK of the trajectory point at t_k (+ finite-difference velocity, gravity torque as torque state), the
warm start is the previous plan shifted by one knot.

GPU arm: `ffmpc_tick_batch_host` with B = 1 (shift + solve + policy on the device, wall clock per tick including the host
copies).  CPU arm: the oracle port on one core, the same x_now / references / shifted warm start (taken from the GPU plan of
the previous tick), the same iteration cap and early exit.  Prints one JSON line.

    python tools/tick_stream.py --variant classical --ticks 1000
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def gpu_stream(variant: str, ticks: int, t0_s: float = 0.0, keep_plans: int = 0, device: int = 0):
    """The GPU arm: `ticks` warm-started ticks of one controller through ffmpc_tick_batch_host (B = 1).
    Returns (result dict, context for the CPU replay)."""
    from franka_force_feedback_mpc_b200 import config, engine, panda, synthetic

    ff = variant == "ff"
    N = 40 if ff else 36
    cfgd = (config.benchmark_ff_config if ff else config.benchmark_classical_config)(horizon=N)
    m = panda.default_model()
    cfg, am = config.to_abi(cfgd, m, panda.Q_NEUTRAL), m.to_abi()
    dt, dt_ocp, nx = 0.005, float(cfgd.dt_ocp), 21 if ff else 14
    traj = synthetic.benchmark_trajectory(m)
    K = ticks
    tk = t0_s + dt * np.arange(K + 2)
    p_mj, _, surf = traj(tk)
    q = synthetic.batched_ik(m, p_mj @ panda.R_MJ_FROM_PIN, panda.reference_R_des(m), panda.Q_NEUTRAL, iters=60)
    dq = np.gradient(q, dt, axis=0)
    g = synthetic.batched_gravity_torque(m, q)
    x_all = np.concatenate([q, dq] + ([g] if ff else []), axis=1)
    x_reg = np.concatenate([panda.Q_NEUTRAL, np.zeros(7)])[None]
    pol = config.tick_policy(cfgd, dt_mpc=dt)

    eng = engine.FFMPCEngine(cfg, am, device=device, max_batch=1)
    lat, iters, plans = [], [], []
    prev_mode = None
    for k in range(K):
        p_ref, v_ref = synthetic.refs_pin(traj, tk[k:k + 1], N, dt_ocp)
        mode = np.array([int(surf[k])], np.int32)
        t0 = time.perf_counter()
        out = eng.tick_host(mode, x_all[k:k + 1], p_ref, v_ref, x_reg, g[k:k + 1], pol, int(cfgd.max_iters), fixed_iters=False,
                            reset=(k == 0), want_plan=bool(keep_plans))
        lat.append(time.perf_counter() - t0)
        iters.append(int(out["iters"][0]))
        if keep_plans and k < keep_plans:
            plans.append((mode.copy(), p_ref, v_ref, out["xs"][0].copy(), out["us"][0].copy(), prev_mode is not None and prev_mode == int(mode[0])))
        prev_mode = int(mode[0])
    eng.close()
    lat = np.array(lat[min(20, K // 4):]) * 1e3

    res = {"config": f"BASELINE config {'2' if ff else '1'}: {'ForceFeedback' if ff else 'Classical'}CrocoddylMPC benchmark preset, N={N}, "
                     f"max_iters={int(cfgd.max_iters)} with early exit, dt=5 ms, kinematic replay of the flat scenario",
           "ticks": K, "gpu_tick_ms": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)), "mean": float(lat.mean())},
           "gpu_what": "ffmpc_tick_batch_host, B=1, wall clock per tick (H2D + shift + solve + policy + D2H)",
           "iters_mean": float(np.mean(iters)), "realtime_budget_ms": 5.0}
    return res, dict(plans=plans, cfg=cfg, am=am, cfgd=cfgd, x_all=x_all, x_reg=x_reg, nx=nx, N=N)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", default="classical", choices=["classical", "ff"])
    ap.add_argument("--ticks", type=int, default=1000)
    ap.add_argument("--t0", type=float, default=0.0)
    ap.add_argument("--cpu-ticks", type=int, default=300, help="ticks replayed on the CPU port (0 = skip)")
    args = ap.parse_args()

    from franka_force_feedback_mpc_b200.batch import BatchInputs

    res, ctx = gpu_stream(args.variant, args.ticks, args.t0, keep_plans=args.cpu_ticks)
    plans, cfg, am, cfgd, x_all, x_reg, nx, N = (ctx[k] for k in ("plans", "cfg", "am", "cfgd", "x_all", "x_reg", "nx", "N"))
    if args.cpu_ticks:
        from oracle import oracle as orc

        lib, how = orc.timing_lib()
        cl, ci = [], []
        for k in range(1, len(plans)):
            mode, p_ref, v_ref, _, _, _ = plans[k]
            _, _, _, xs_p, us_p, _ = plans[k - 1]
            warm = plans[k][5]
            x0 = x_all[k:k + 1]
            kw = {}
            if warm:   # `_shift_guess` (crocoddyl_classical.py:733-757)
                kw = dict(xs_init=np.concatenate([x0, xs_p[1:]])[None], us_init=np.concatenate([us_p[1:], us_p[-1:]])[None])
            inp = BatchInputs(nx, N, mode, x0, p_ref, v_ref, x_reg, **kw)
            t0 = time.perf_counter()
            o = orc.solve_batch(cfg, am, inp, int(cfgd.max_iters), False, nthreads=1, _use=lib)
            cl.append(time.perf_counter() - t0)
            ci.append(int(o["iters"][0]))
        cl = np.array(cl) * 1e3
        res["cpu_tick_ms"] = {"p50": float(np.percentile(cl, 50)), "p99": float(np.percentile(cl, 99)), "mean": float(cl.mean()), "ticks": int(cl.size)}
        res["cpu_what"] = f"oracle port ({how}), 1 core, same inputs and shifted warm start, solve only"
        res["cpu_iters_mean"] = float(np.mean(ci))
    print(json.dumps(res))


if __name__ == "__main__":
    main()
