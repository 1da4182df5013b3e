"""Development sweep on the GPU: per-kernel-family device time for a list of (B, env overrides)."""
import os, sys, time, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from franka_force_feedback_mpc_b200 import config, engine, panda, synthetic
from franka_force_feedback_mpc_b200.batch import BatchInputs

m = panda.default_model(); am = m.to_abi()
_cache = {}
def inputs(ff, B, N, nc, cf):
    key = (ff, B, N, nc, cf)
    if key not in _cache:
        s = synthetic.sample_instances(B, N, ff, seed=5, contact_fraction=cf, scenario="actuation_uncertainty")
        _cache[key] = BatchInputs.from_samples(s, 21 if ff else 14, N)
    return _cache[key]

def run(ff=True, B=16384, N=50, iters=10, nc=1, cf=1.0, env=None, reps=2, tag=""):
    for k, v in (env or {}).items():
        os.environ[k] = str(v)
    kw = dict(horizon=N, contact_model="point3d" if nc == 3 else "normal_1d")
    cfgd = config.benchmark_ff_config(**kw) if ff else config.benchmark_classical_config(**kw)
    c = config.to_abi(cfgd, m, panda.Q_NEUTRAL)
    inp = inputs(ff, B, N, nc, cf)
    eng = engine.FFMPCEngine(c, am, 0, B)
    eng.set_profiling(True)
    best = None
    for r in range(reps):
        out = eng.solve_host(inp, iters, True, want=("cost", "iters"))
        ms, nl = eng.last_timing()
        prof = eng.get_profile()
        if best is None or ms < best[0]:
            best = (ms, prof)
    ms, prof = best
    print(f"{tag} ff={ff} nc={nc} B={B} iters={iters} env={env}: {ms:8.2f} ms  {B / ms * 1e3:9.0f} solves/s  | " +
          " ".join(f"{k}={v[0]:.2f}" for k, v in prof.items()) + f" | mean cost {out['cost'].mean():.6f}", flush=True)
    eng.close()
    for k in (env or {}):
        os.environ.pop(k, None)

if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "fwd"
    if what == "fwd":
        for B in (4096, 16384, 65536):
            for A in (1, 2, 4, 8, 16):
                run(B=B, env={"FFMPC_FWD_LANES": A})
        for A in (1, 4, 16):
            run(B=1, env={"FFMPC_FWD_LANES": A}, reps=3)
        run(ff=False, B=4096)
        run(ff=True, nc=3, B=16384, iters=20, cf=0.75)
    elif what == "default":
        for B in (1, 256, 4096, 16384, 65536):
            run(B=B)
        run(ff=False, B=4096)
        run(ff=False, B=65536)
        run(ff=True, nc=3, B=16384, iters=20, cf=0.75)
