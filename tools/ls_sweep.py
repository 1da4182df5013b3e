"""Development sweep on the GPU: device throughput (inputs resident, CUDA events) of one batched solve per line-search
schedule (ffmpc_options.line_search) and batch size.  Feeds the AUTO thresholds of set_schedule (csrc/ffmpc_engine.cu).

    python tools/ls_sweep.py [--variant ff|classical] [--batches 1,256,...] [--modes queue,spec,grouped,tail,auto] [--out file.json]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from franka_force_feedback_mpc_b200 import config, engine, panda, synthetic  # noqa: E402
from franka_force_feedback_mpc_b200.batch import BatchInputs  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", default="ff")
    ap.add_argument("--batches", default="1,64,256,512,1024,2048,4096,8192,12288,16384,32768,65536")
    ap.add_argument("--modes", default="queue,spec,grouped,tail,auto")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--horizon", type=int, default=50)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--sub-batches", type=int, default=0)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    ff = a.variant == "ff"
    m = panda.default_model()
    am = m.to_abi()
    cfgd = config.benchmark_ff_config(horizon=a.horizon) if ff else config.benchmark_classical_config(horizon=a.horizon)
    cfg = config.to_abi(cfgd, m, panda.Q_NEUTRAL)
    dev = torch.device("cuda", 0)
    st = torch.cuda.current_stream(dev)
    rows = []
    for B in [int(x) for x in a.batches.split(",")]:
        s = synthetic.sample_instances(B, a.horizon, ff, seed=5, contact_fraction=1.0, scenario="actuation_uncertainty")
        inp = BatchInputs.from_samples(s, 21 if ff else 14, a.horizon)
        t_in = {k: torch.as_tensor(getattr(inp, k)).to(dev) for k in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref")}
        t_out = None
        ref_cost = None
        for mode in a.modes.split(","):
            if mode == "spec" and B > 16384:
                continue
            eng = engine.FFMPCEngine(cfg, am, 0, B, line_search=mode, sub_batches=a.sub_batches)
            best = 1e30
            for r in range(a.reps + 1):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                t_out = eng.solve_device(t_in, a.iters, True, t_out=t_out, stream=st)
                e1.record(st)
                e1.synchronize()
                if r > 0:
                    best = min(best, e0.elapsed_time(e1))
            _, nl = eng.last_timing()
            cost = t_out["cost"].cpu().numpy().copy()
            if ref_cost is None:
                ref_cost = cost
            same = bool(np.array_equal(cost, ref_cost))
            info = eng.info()
            eng.close()
            rows.append(dict(B=B, mode=mode, resolved=info["line_search"], sub_batches=info["sub_batches"], ms=best, solves_per_s=B / best * 1e3,
                             launches=nl, cost_bitwise_equal=same))
            print(f"B={B:6d} {mode:8s} -> {info['line_search']:8s} S={info['sub_batches']} {best:9.3f} ms {B / best * 1e3:10.0f} solves/s "
                  f"launches={nl} same={same}", flush=True)
        del t_in, t_out
        torch.cuda.empty_cache()
    if a.out:
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
