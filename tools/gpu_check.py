"""Ad-hoc GPU check used during development: parity vs the oracle for every variant/mode and a
quick throughput probe.  (The judged versions of these live in tests/ and bench.py.)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from franka_force_feedback_mpc_b200 import config, engine, panda, synthetic
from franka_force_feedback_mpc_b200.batch import BatchInputs
from oracle import oracle as orc

m = panda.default_model(); am = m.to_abi()
def rel(a, b): return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))

def parity(ff, nc, cf, B=32, N=50, iters=10):
    kw = dict(horizon=N, contact_model="point3d" if nc == 3 else "normal_1d")
    cfgd = config.benchmark_ff_config(**kw) if ff else config.benchmark_classical_config(**kw)
    c = config.to_abi(cfgd, m, panda.Q_NEUTRAL)
    nx = 21 if ff else 14
    s = synthetic.sample_instances(B, N, ff, seed=5, contact_fraction=cf)
    inp = BatchInputs.from_samples(s, nx, N)
    eng = engine.FFMPCEngine(c, am, 0, B)
    out = eng.solve_host(inp, iters, True)
    ref = orc.solve_batch(c, am, inp, iters, True, nthreads=os.cpu_count())
    print(f"parity ff={ff} nc={nc} cf={cf}: " + " ".join(f"{k}:{rel(out[k], ref[k]):.1e}" for k in ("xs", "us", "K", "k", "cost")),
          "iters_eq", np.array_equal(out["iters"], ref["iters"]), "gap", np.abs(out["gaps"]).max(), flush=True)
    eng.close()

def speed(ff, B, N=50, iters=10, nc=1, cf=1.0, reps=3, sub_batches=0):
    kw = dict(horizon=N, contact_model="point3d" if nc == 3 else "normal_1d")
    cfgd = config.benchmark_ff_config(**kw) if ff else config.benchmark_classical_config(**kw)
    c = config.to_abi(cfgd, m, panda.Q_NEUTRAL)
    nx = 21 if ff else 14
    s = synthetic.sample_instances(B, N, ff, seed=5, contact_fraction=cf)
    inp = BatchInputs.from_samples(s, nx, N)
    eng = engine.FFMPCEngine(c, am, 0, B, sub_batches=sub_batches)
    for r in range(reps):
        t = time.time(); out = eng.solve_host(inp, iters, True, want=("xs", "us", "K", "cost", "iters", "ok")); wall = time.time() - t
        ms, nl = eng.last_timing()
        print(f"speed ff={ff} nc={nc} B={B} N={N} iters={iters}: kernels {ms:.2f} ms ({B / ms * 1e3:.0f} solves/s), e2e {wall * 1e3:.1f} ms ({B / wall:.0f} solves/s), launches {nl}, mean cost {out['cost'].mean():.4f}", flush=True)
    eng.close()

if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("all", "parity"):
        for ff in (False, True):
            for nc, cf in ((1, 1.0), (1, 0.0), (1, 0.5), (3, 1.0)):
                parity(ff, nc, cf)
    if what in ("all", "speed"):
        speed(True, 4096); speed(False, 4096); speed(True, 16384); speed(True, 1, reps=5)
    if what == "prof":   # PROF_B instances in PROF_S sub-batches (default: 4096 in 2, i.e. 2048 per launch)
        speed(True, int(os.environ.get("PROF_B", 4096)), iters=2, reps=1, sub_batches=int(os.environ.get("PROF_S", 0)))
    if what == "lat":
        speed(True, 1, reps=6)
