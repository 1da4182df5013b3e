#!/usr/bin/env python
"""bench.py -- FDDP MPC solves/sec (fp64, batched) on 1..8 B200, one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch B] ...
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json config 4, the configuration the 1e5 solves/s/GPU target is quoted on):
ForceFeedbackCrocoddylMPC OCP (nx=21, nu=7), normal_1d contact, horizon 50, dt_ocp 0.01, Box-FDDP with
a FIXED 10 iterations (converged early exit disabled), cold start, synthetic x0 / reference
trajectories: every GPU solves `--batch` instances (default 65536 = the whole of config 4, which fits one
GPU: 4 scenarios x 16384) made of the four sweep scenarios tilted_5 / tilted_10 / tilted_15 /
actuation_uncertainty in equal parts.  One step = one batched solve of the whole per-GPU batch.  Instances are independent:
ranks shard them, there is no collective on the solve path (weak scaling).

Printed JSON line (rank 0): see the keys below; `value` = device-resident throughput (CUDA events on
the solve stream), `e2e` = through the host-pointer C-ABI call with pinned host buffers, H2D of the
inputs and D2H of xs/us/K/k/cost/iters/ok inside the timed region.

`--impl reference` times the CPU restatement of the reference path (oracle/, -O3 AVX2 build, all host
threads) on a bounded sample of the same workload.  The reference's own implementation is Crocoddyl /
Pinocchio called from Python: not installed here or on the GPU box, so the oracle port
is the CPU arm.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "fddp_mpc_solves_per_sec"
UNIT = "solves/s"
SCENARIOS = ("tilted_5", "tilted_10", "tilted_15", "actuation_uncertainty")

# Algorithmic fp64 work, SURVEY.md section 8(d): dense-algebra operation counts of the reference algorithm
# per knot (fma = 2), no credit for symmetry / sparsity.  FLOP per launch = per-knot figure x knots.
F_DIFF = {0: 16000.0, 1: 20000.0, 3: 22000.0}   # action-model derivatives, by contact dimension
F_CALC = {0: 5000.0, 1: 6500.0, 3: 6500.0}      # action-model calc (one rollout knot)
F_EI = {21: 2000.0, 14: 1000.0}                  # gaps / expected-improvement terms per knot


def f_bwd(nx: int, nu: int = 7) -> float:
    return 4 * nx**3 + 6 * nx**2 * nu + 4 * nx * nu**2 + nu**3 / 3.0 + 6 * nx**2 + 4 * nx * nu + 4 * nu**2


def flop_per_solve(nx: int, nc: int, N: int, iters: int, rollouts_per_iter: float = 1.0) -> float:
    return iters * N * (F_DIFF[nc] + f_bwd(nx) + F_EI[nx] + rollouts_per_iter * F_CALC[nc])


def io_bytes_per_solve(nx: int, N: int) -> int:
    """Compulsory I/O of one solve: inputs x0, refs, warm start; outputs xs, us, K, k, scalars."""
    n_in = (N + 1) * nx + N * 7 + (N + 1) * 6 + nx
    n_out = (N + 1) * nx + N * 7 + N * 7 * nx + N * 7 + 4
    return 8 * (n_in + n_out)


# ------------------------------------------------------------------------------------------------
def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=65536, help="instances per GPU per step (default: all of BASELINE config 4)")
    ap.add_argument("--horizon", type=int, default=50)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--variant", default="ff", choices=["ff", "classical"])
    ap.add_argument("--contact-model", default="normal_1d", choices=["normal_1d", "point3d"])
    ap.add_argument("--contact-fraction", type=float, default=1.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--cpu-sample", type=int, default=0, help="instances of the CPU baseline sample (0 = auto)")
    ap.add_argument("--sub-batches", type=int, default=0, help="ffmpc_options.sub_batches (0 = auto)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak (default, the driver's contract): --batch instances on EVERY GPU, and a `strong` block beside it; "
                         "strong: --batch instances IN TOTAL sharded over the GPUs (BASELINE configs[3] as written) as the line's value")
    ap.add_argument("--tick-stream-ticks", type=int, default=600, help="ticks of the B = 1 latency streams (configs 1, 2)")
    ap.add_argument("--no-also", action="store_true", help="skip the short BASELINE config 3 / config 5 runs")
    a = ap.parse_args()
    a.n_alphas_report = 10
    return a


def make_problem(args, rank: int, B: int):
    from franka_force_feedback_mpc_b200 import config, panda, synthetic
    from franka_force_feedback_mpc_b200.batch import BatchInputs

    ff = args.variant == "ff"
    model = panda.default_model()
    kw = dict(horizon=args.horizon, contact_model=args.contact_model, max_iters=args.iters)
    cfgd = config.benchmark_ff_config(**kw) if ff else config.benchmark_classical_config(**kw)
    cfg = config.to_abi(cfgd, model, panda.Q_NEUTRAL)
    nx = 21 if ff else 14
    parts = []
    per = [B // len(SCENARIOS) + (1 if i < B % len(SCENARIOS) else 0) for i in range(len(SCENARIOS))]
    for sc, n in zip(SCENARIOS, per):
        if n == 0:
            continue
        seed = synthetic.SCENARIO_SEEDS[sc] + 1000 * 4 + 100000 * rank  # scenario seed + 1000 * config# (+ rank)
        parts.append(synthetic.sample_instances(n, args.horizon, ff, seed=seed, contact_fraction=args.contact_fraction,
                                                scenario=sc, model=model))
    s = {k: np.concatenate([p[k] for p in parts], axis=0) for k in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref")}
    inp = BatchInputs.from_samples(s, nx, args.horizon)
    return cfg, model.to_abi(), inp, nx


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.rows, self.proc, self.device = [], None, device

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except ValueError:
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


def _cpu_rate_probe(args, cfg, model, inp, lib, nthreads: int) -> float:
    """solves/s of the CPU port on a tiny probe (4 instances per thread), used to size the timed sample."""
    from oracle import oracle as orc

    n = min(inp.B, 4 * nthreads)
    sub = inp.slice(np.arange(n))
    t0 = time.perf_counter()
    orc.solve_batch(cfg, model, sub, args.iters, True, nthreads=nthreads, _use=lib)
    return n / max(time.perf_counter() - t0, 1e-6)


def _cpu_sample_size(args, rate: float, seconds: float, B: int, nthreads: int) -> int:
    if args.cpu_sample:
        return min(args.cpu_sample, B)
    n = int(rate * seconds)
    n = max(nthreads, (n // nthreads) * nthreads)
    return min(n, B)


def cpu_baseline(args, cfg, model, inp, nthreads: int, seconds: float = 15.0) -> dict:
    """The oracle port timed on the host cores: a bounded sample (about 15 s of CPU work) of the same workload."""
    from oracle import oracle as orc

    lib, how = orc.timing_lib()
    rate = _cpu_rate_probe(args, cfg, model, inp, lib, nthreads)
    sub = inp.slice(np.arange(_cpu_sample_size(args, rate, seconds, inp.B, nthreads)))
    orc.rollout_stats(reset=True, _use=lib)
    t0 = time.perf_counter()
    orc.solve_batch(cfg, model, sub, args.iters, True, nthreads=nthreads, _use=lib)
    dt = time.perf_counter() - t0
    nroll, niter, hist = orc.rollout_stats(_use=lib)
    return {"value": sub.B / dt, "unit": UNIT, "cores": nthreads, "kind": "port",
            "rollouts_per_iteration": (nroll / niter) if niter else None, "trials_per_step_length": hist[:args.n_alphas_report],
            "sample": f"first {sub.B} instances of the same batch, {args.iters} fixed iterations, oracle/ffmpc_oracle.cpp ({how}), "
                      f"{nthreads} threads over instances, {dt:.1f} s",
            "note": "reference's own CPU path (Crocoddyl/Pinocchio via Python) is not installable on this box"}


def run_reference(args, rank: int, world: int):
    if rank != 0:
        return
    from oracle import oracle as orc

    lib, how = orc.timing_lib()
    nthreads = os.cpu_count() or 1
    cfg, model, probe_inp, nx = make_problem(args, 0, 4 * nthreads)
    rate = _cpu_rate_probe(args, cfg, model, probe_inp, lib, nthreads)
    # each step = a bounded sample (about 8 s of CPU work) of the b200 arm's per-GPU batch
    n = _cpu_sample_size(args, rate, 8.0, args.batch, nthreads)
    cfg, model, inp, nx = make_problem(args, 0, n)
    for _ in range(args.warmup):
        orc.solve_batch(cfg, model, inp.slice(np.arange(min(nthreads, inp.B))), args.iters, True, nthreads=nthreads, _use=lib)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        orc.solve_batch(cfg, model, inp, args.iters, True, nthreads=nthreads, _use=lib)
    dt = time.perf_counter() - t0
    val = inp.B * args.steps / dt
    sample = (f"{inp.B} instances/step of the b200 arm's workload, oracle/ffmpc_oracle.cpp ({how}), {nthreads} threads over instances")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, args.batch, nx),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": nthreads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU arm = from-scratch restatement of the reference algorithm (oracle/); Crocoddyl/Pinocchio themselves are "
                "not installed on this box and cannot be (no network)",
    }
    print(json.dumps(line), flush=True)


def short_run(base_args, local_rank: int, variant: str, contact_model: str, B: int, iters: int, cf: float, steps: int = 2) -> dict:
    """A short device-timed run of another BASELINE configuration (same engine, same timing method), for the `also` key."""
    import argparse as _ap

    import torch

    from franka_force_feedback_mpc_b200 import engine

    a = _ap.Namespace(**vars(base_args))
    a.variant, a.contact_model, a.batch, a.iters, a.contact_fraction = variant, contact_model, B, iters, cf
    cfg, model, inp, nx = make_problem(a, 0, B)
    dev = torch.device("cuda", local_rank)
    N = a.horizon
    eng = engine.FFMPCEngine(cfg, model, device=local_rank, max_batch=B, sub_batches=a.sub_batches)
    t_in = {k: torch.from_numpy(getattr(inp, k)).to(dev) for k in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref")}
    shapes = dict(xs=(B, N + 1, nx), us=(B, N, 7), K=(B, N, 7, nx), k=(B, N, 7), cost=(B,), iters=(B,), ok=(B,))
    t_out = {n: torch.empty(shapes[n], device=dev, dtype=torch.int32 if n in ("iters", "ok") else torch.float64) for n in shapes}
    st = torch.cuda.Stream(device=dev)
    for _ in range(3):
        eng.solve_device(t_in, iters, True, t_out=t_out, stream=st)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(steps):
        eng.solve_device(t_in, iters, True, t_out=t_out, stream=st)
    e1.record(st)
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / steps
    ok = bool((t_out["iters"] == iters).all().item()) and bool(torch.isfinite(t_out["cost"]).all().item())
    eng.close()
    return {"value": B / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "steps": steps, "config": workload_config(a, B, nx), "all_iters_ran": ok}


def workload_config(args, B: int, nx: int) -> dict:
    if args.variant == "ff" and args.contact_model == "normal_1d":
        which = "BASELINE config 4"                                   # the configuration the metric is quoted on (default)
    elif args.variant == "classical" and args.contact_model == "normal_1d":
        which = "BASELINE config 3 shape"
    elif args.variant == "ff":
        which = "BASELINE config 5 shape"
    else:
        which = "non-BASELINE shape"
    return {
        "workload": f"{'ForceFeedback' if args.variant == 'ff' else 'Classical'}CrocoddylMPC OCP sweep ({which}): "
                    f"nx={nx} nu=7 {args.contact_model} contact, horizon {args.horizon}, dt_ocp 0.01, Box-FDDP, "
                    f"{args.iters} fixed iterations, cold start, scenarios {'/'.join(SCENARIOS)}",
        "batch_per_gpu": B, "horizon": args.horizon, "iters": args.iters, "nx": nx, "contact_fraction": args.contact_fraction,
        "parallelism": f"instance-sharded x{args.gpus}, no collective on the solve path",
        "sub_batches": getattr(args, "sub_batches_resolved", None),
        "l2": "per-step working set (node records + trajectories + gains) is GBs, far larger than the 126 MB L2",
    }


def bind_to_gpu_numa_node(local_rank):
    """Pin this process (and with it the first-touch placement of its pinned host buffers) to the CPUs of the NUMA node its GPU
    hangs off, so that eight ranks do not all write their results into one node's memory.  Best effort: returns what it did."""
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        node = -1
        for cand in (bus, bus[4:] if len(bus) > 12 else bus):   # nvml prints an 8-digit domain, sysfs a 4-digit one
            path = f"/sys/bus/pci/devices/{cand}/numa_node"
            if os.path.exists(path):
                node = int(open(path).read().strip())
                break
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()]
        info = {"gpu_numa_node": node, "host_numa_nodes": len(nodes), "bound": False}
        if node >= 0 and len(nodes) > 1:
            cpus = set()
            for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
            cpus &= os.sched_getaffinity(0)
            if cpus:
                os.sched_setaffinity(0, cpus)
                info["bound"] = True
                info["cpus"] = len(cpus)
        return info
    except Exception as e:  # no nvml / sysfs: run unbound
        return {"bound": False, "error": type(e).__name__}


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU path (use --impl reference for the CPU arm)")
    import __graft_entry__ as ge
    from franka_force_feedback_mpc_b200 import _abi, engine
    from franka_force_feedback_mpc_b200.batch import OUTPUT_NAMES

    if not os.path.exists(ge.LIB):
        ge.build_cuda()
    numa = bind_to_gpu_numa_node(local_rank)
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    B, N = args.batch, args.horizon
    total = world * B
    if args.scaling == "strong" and world > 1:
        # the one global sweep (rank-0 seeds), this rank's contiguous slice of it
        from franka_force_feedback_mpc_b200 import sharding as _sh

        lo, hi = _sh.shard_bounds(args.batch, world, rank)
        cfg, model, ginp, nx = make_problem(args, 0, args.batch)
        inp = ginp.slice(np.arange(lo, hi))
        total, B = args.batch, int(_sh.shard_sizes(args.batch, world).max())
        if inp.B != B:   # ragged split: pad this rank's shard with copies of its last instance (timing only; work is max over ranks)
            inp = inp.slice(np.minimum(np.arange(B), inp.B - 1))
    else:
        cfg, model, inp, nx = make_problem(args, rank, B)
    nc = (3 if args.contact_model == "point3d" else 1) if args.contact_fraction > 0 else 0
    eng = engine.FFMPCEngine(cfg, model, device=local_rank, max_batch=B, sub_batches=args.sub_batches)

    args.sub_batches_resolved = eng.info()["sub_batches"]

    # ---- device-resident inputs / outputs (torch tensors are only buffers) ----
    t_in = {k: torch.from_numpy(getattr(inp, k)).to(dev) for k in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref")}
    want = ("xs", "us", "K", "k", "cost", "stop", "iters", "ok", "f_pred")
    shapes = dict(xs=(B, N + 1, nx), us=(B, N, 7), K=(B, N, 7, nx), k=(B, N, 7), cost=(B,), stop=(B,), gaps=(B, N + 1, nx),
                  f_pred=(B, 2, 3), iters=(B,), ok=(B,))
    t_out = {n: torch.empty(shapes[n], device=dev, dtype=torch.int32 if n in ("iters", "ok") else torch.float64) for n in want}
    stream = torch.cuda.Stream(device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def step_device():
        eng.solve_device(t_in, args.iters, True, t_out=t_out, stream=stream)

    eng.set_profiling(False)
    for _ in range(args.warmup):
        step_device()
    barrier()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    fam_ms = {n: 0.0 for n in _abi.FAMILY_NAMES}
    fam_launch = {n: 0 for n in _abi.FAMILY_NAMES}
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with torch.cuda.stream(stream):
        ev0.record(stream)
    for _ in range(args.steps):
        step_device()
    with torch.cuda.stream(stream):
        ev1.record(stream)
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    launches_per_step = int(eng.last_timing()[1])
    clk = clocks.stop() if rank == 0 else None
    # kernel-family split: untimed, on a second handle WITHOUT sub-batches (one stream, an event after every kernel family).
    # The timed steps above overlap their sub-batches, so these family times add up to a little more than ms_per_step;
    # they say where the time goes, `value` says how long it takes.
    eng_p = engine.FFMPCEngine(cfg, model, device=local_rank, max_batch=B, sub_batches=1)
    eng_p.solve_device(t_in, args.iters, True, t_out=t_out, stream=stream)
    eng_p.set_profiling(True)
    eng_p.solve_device(t_in, args.iters, True, t_out=t_out, stream=stream)
    torch.cuda.synchronize(dev)
    prof = eng_p.get_profile()
    for n, (ms, nl) in prof.items():
        fam_ms[n], fam_launch[n] = ms, nl
    eng_p.close()
    t_ms = torch.tensor([ms_total], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms_total = float(t_ms.item())
    iters_out = t_out["iters"].cpu().numpy()
    ok_dev = bool(np.all(iters_out == args.iters)) and bool(torch.isfinite(t_out["cost"]).all().item())

    # ---- end to end through the host-pointer C-ABI call: pinned host buffers, H2D + D2H of every step inside ----
    # ffmpc_solve_batch_host_async + ffmpc_wait: a stream of solves, each copying its own inputs in and its own results out;
    # a finished sub-batch's results cross PCIe while the next step's other sub-batches are solved.  Two sets of pinned
    # output buffers alternate (a call in flight owns its buffers until the wait).
    h_in = {k: torch.from_numpy(getattr(inp, k)).pin_memory() for k in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref")}
    e2e_want = ("xs", "us", "K", "k", "cost", "iters", "ok")
    s_in = _abi.BatchIn()
    s_in.B = B
    for k, t in h_in.items():
        setattr(s_in, k, t.data_ptr())
    h_outs, s_outs = [], []
    for _ in range(2):
        h_out = {n: torch.empty(shapes[n], dtype=torch.int32 if n in ("iters", "ok") else torch.float64).pin_memory() for n in e2e_want}
        s_out = _abi.BatchOut()
        for k, t in h_out.items():
            setattr(s_out, k, t.data_ptr())
        h_outs.append(h_out)
        s_outs.append(s_out)
    h2d = sum(t.numel() * t.element_size() for t in h_in.values())
    d2h = sum(t.numel() * t.element_size() for t in h_outs[0].values())
    for i in range(max(2, min(args.warmup, 3))):
        eng.solve_host_ptrs_async(s_in, s_outs[i % 2], args.iters, True)
    eng.wait()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        eng.solve_host_ptrs_async(s_in, s_outs[i % 2], args.iters, True)
    eng.wait()
    e2e_s = time.perf_counter() - t0
    # the same steps one blocking call at a time (ffmpc_solve_batch_host): what a caller without a pipeline sees
    t0 = time.perf_counter()
    for i in range(args.steps):
        eng.solve_host_ptrs(s_in, s_outs[i % 2], args.iters, True)
    e2e_sync_s = time.perf_counter() - t0
    t_e = torch.tensor([e2e_s, e2e_sync_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
    e2e_s, e2e_sync_s = float(t_e[0].item()), float(t_e[1].item())
    e2e_ok = all(bool(np.isfinite(h["cost"].numpy()).all()) and bool((h["iters"].numpy() == args.iters).all()) for h in h_outs)
    d2h_gbs_rank = d2h * args.steps / e2e_s * 1e-9   # this rank's average device->host rate over the pipelined steps

    # ---- strong scaling: BASELINE.json configs[3] as written, `--batch` instances in total sharded over the GPUs ----
    # (the weak run above solves `--batch` instances on EVERY GPU).  Each rank solves its contiguous slice of the sweep and
    # the per-instance results cost / iters / ok / f_pred are gathered from device memory over NCCL inside the timed region
    # (north_star: "only a final result gather over NVLink"); no collective on the solve path.
    strong = None
    if world > 1 and args.scaling == "weak":
        from franka_force_feedback_mpc_b200 import sharding

        lo, hi = sharding.shard_bounds(B, world, rank)
        _, _, ginp, _ = make_problem(args, 0, B)          # the one global sweep (rank-0 seeds), sliced
        shard = ginp.slice(np.arange(lo, hi))
        Bs = shard.B
        eng_s = engine.FFMPCEngine(cfg, model, device=local_rank, max_batch=Bs, sub_batches=args.sub_batches)
        ts_in = {k: torch.from_numpy(getattr(shard, k)).to(dev) for k in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref")}
        sh = dict(xs=(Bs, N + 1, nx), us=(Bs, N, 7), K=(Bs, N, 7, nx), k=(Bs, N, 7), cost=(Bs,), stop=(Bs,), f_pred=(Bs, 2, 3), iters=(Bs,), ok=(Bs,))
        ts_out = {n: torch.empty(sh[n], device=dev, dtype=torch.int32 if n in ("iters", "ok") else torch.float64) for n in want}
        gathered = None

        def step_strong():
            eng_s.solve_device(ts_in, args.iters, True, t_out=ts_out, stream=stream)
            with torch.cuda.stream(stream):
                return sharding.gather_results({n: ts_out[n] for n in ("cost", "iters", "ok", "f_pred")}, B, world, rank, dist=dist, to_numpy=False)

        for _ in range(max(2, min(args.warmup, 3))):
            step_strong()
        barrier()
        with torch.cuda.stream(stream):
            ev0.record(stream)
        for _ in range(args.steps):
            gathered = step_strong()
        with torch.cuda.stream(stream):
            ev1.record(stream)
        barrier()
        t_s = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
        dist.all_reduce(t_s, op=dist.ReduceOp.MAX)
        ms_strong = float(t_s.item())
        strong = {"total": B, "per_gpu": int(sharding.shard_sizes(B, world).max()), "value": B * args.steps / (ms_strong * 1e-3),
                  "unit": UNIT, "ms_per_step": ms_strong / args.steps, "sub_batches": eng_s.info()["sub_batches"],
                  "gather": "cost / iters / ok / f_pred all-gathered from device memory (NCCL) inside the timed region"}
        if rank == 0:
            strong["gather_ok"] = bool(gathered["cost"].shape[0] == B and torch.isfinite(gathered["cost"]).all().item()
                                       and (gathered["iters"] == args.iters).all().item())
        eng_s.close()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # B=1 latency first: it is launch-bound wall clock, measure it before the CPU baseline loads every host core
    latency = None
    if not args.no_latency:
        one = inp.slice(np.arange(1))
        eng1 = engine.FFMPCEngine(cfg, model, device=local_rank, max_batch=1)
        o1 = None
        lat = []
        for i in range(60):
            t0 = time.perf_counter()
            o1 = eng1.solve_host(one, args.iters, True, want=("xs", "us", "K", "cost", "iters", "ok"), out=o1)
            lat.append(time.perf_counter() - t0)
        lat = np.array(lat[10:]) * 1e3
        latency = {"p50_ms": float(np.percentile(lat, 50)), "p99_ms": float(np.percentile(lat, 99)), "n": int(lat.size),
                   "what": "B=1 solve through ffmpc_solve_batch_host (H2D + 10 fixed iterations + D2H), wall clock"}
        eng1.close()
        if not args.no_cpu_baseline:
            # the same single instance on ONE host core (B=1 GPU latency vs the CPU port on 1 core)
            from oracle import oracle as orc

            lib, how = orc.timing_lib()
            cl = []
            for i in range(7):
                t0 = time.perf_counter()
                orc.solve_batch(cfg, model, one, args.iters, True, nthreads=1, _use=lib)
                cl.append(time.perf_counter() - t0)
            latency["cpu_port_1core_p50_ms"] = float(np.percentile(np.array(cl[2:]) * 1e3, 50))
        # BASELINE configs 1 and 2: one controller, a stream of warm-started receding-horizon ticks (the reference's loop
        # runs one such tick per 5 ms, run_classical.py:203-207); kinematic replay of the flat scenario, see tools/tick_stream.py
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import tick_stream

        latency["tick_stream"] = {}
        for name, variant in (("config1_classical", "classical"), ("config2_ff", "ff")):
            r, _ = tick_stream.gpu_stream(variant, args.tick_stream_ticks, device=local_rank)
            latency["tick_stream"][name] = {"p50_ms": r["gpu_tick_ms"]["p50"], "p99_ms": r["gpu_tick_ms"]["p99"], "ticks": r["ticks"],
                                            "iters_mean": r["iters_mean"], "budget_ms": 5.0, "what": r["gpu_what"], "config": r["config"]}

    # ---- rank 0 only: CPU baseline (also counts the rollouts the line search really runs), roofline, latency ----
    cpu = None
    rollouts = 1.0
    if not args.no_cpu_baseline:
        nthreads = os.cpu_count() or 1
        if world == 1:
            cpu = cpu_baseline(args, cfg, model, inp, nthreads)
            rollouts = cpu.get("rollouts_per_iteration") or 1.0
        else:  # the CPU baseline is an N = 1 figure; here a 1 s sample only counts the rollouts of the line search (FLOP model)
            rollouts = cpu_baseline(args, cfg, model, inp, nthreads, seconds=1.0).get("rollouts_per_iteration") or 1.0
    value = total * args.steps / (ms_total * 1e-3)
    e2e_val = total * args.steps / e2e_s
    peak_tf = engine.dfma_peak_tflops(local_rank)
    knots = B * (N + 1)
    fam_flop = {
        "node_diff": knots * F_DIFF[nc] * args.iters,
        "backward": B * N * f_bwd(nx) * args.iters,
        "forward": knots * (rollouts * F_CALC[nc] + F_EI[nx]) * args.iters,  # rollouts per iteration as counted by the CPU port
    }
    dom = max(("node_diff", "backward", "forward"), key=lambda n: fam_ms[n])
    dom_launches = max(1, fam_launch[dom])
    achieved = fam_flop[dom] / (fam_ms[dom] * 1e-3) * 1e-12 if fam_ms[dom] > 0 else 0.0
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except (OSError, ValueError):
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = frac_exec = None
    counters_note = "profiles/kernel_counters.json not found"
    try:
        with open(os.path.join(ROOT, "profiles", "kernel_counters.json")) as f:
            kc = json.load(f)
        kd = kc["kernels"][dom]
        units_per_launch = {"node_diff": knots, "backward": B * N, "forward": knots}[dom]   # one launch = the whole (sub-)batch family
        traffic = kd["dram_bytes_per_unit"] * units_per_launch * args.iters / dom_launches * (rollouts if dom == "forward" else 1.0)
        flop_exec = kd["flop_executed_per_unit"] * units_per_launch * args.iters * (rollouts if dom == "forward" else 1.0)
        frac_exec = flop_exec / (fam_ms[dom] * 1e-3) * 1e-12 / peak_tf if (fam_ms[dom] > 0 and peak_tf) else None
        counters_note = f"profiles/kernel_counters.json (rev {kc['rev']}, {kc['source']}, B={kc['B_per_launch']} per launch): " \
                        f"{kd['dram_bytes_per_unit']:.0f} B and {kd['flop_executed_per_unit']:.0f} executed FLOP per knot of `{dom}`"
    except (OSError, ValueError, KeyError):
        pass
    solve_ms = sum(fam_ms.values())
    roofline = {
        "bound": "fp64",  # neither "hbm" nor "tensor": the path is bound by the fp64 FMA pipe
        "kernel": dom, "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved / peak_tf if peak_tf else None,
        "peak_source": "fp64 DFMA micro-benchmark run by this bench (ffmpc_dfma_peak); MEASURED_PEAKS.json has no fp64 entry",
        "flop_per_launch": fam_flop[dom] / dom_launches, "ms_per_launch": fam_ms[dom] / dom_launches,
        # DRAM bytes and executed fp64 work of the dominant kernel per launch: per-knot figures of the ncu --set full capture
        # named in `counters` (profiles/kernel_counters.json, written by tools/ncu_summary.py --json from the capture of the
        # revision it names), scaled to this launch's knots
        "traffic": traffic, "frac_executed": frac_exec, "counters": counters_note,
        "kernel_ms_per_step": {n: fam_ms[n] for n in _abi.FAMILY_NAMES},
        "kernel_share": {n: (fam_ms[n] / solve_ms if solve_ms else None) for n in _abi.FAMILY_NAMES},
        "whole_solve": {
            "flop_per_solve": flop_per_solve(nx, nc, N, args.iters, rollouts), "rollouts_per_iteration": rollouts,
            "achieved_tflops": value / world * flop_per_solve(nx, nc, N, args.iters, rollouts) * 1e-12,
            "frac_of_fp64_peak": value / world * flop_per_solve(nx, nc, N, args.iters, rollouts) * 1e-12 / peak_tf if peak_tf else None,
            "io_bytes_per_solve": io_bytes_per_solve(nx, N),
            "hbm_gbs_compulsory": value / world * io_bytes_per_solve(nx, N) * 1e-9,
            "hbm_peak_gbs": hbm_peak, "hbm_peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
            "frac_of_hbm_peak": value / world * io_bytes_per_solve(nx, N) * 1e-9 / hbm_peak,
        },
    }

    also = None
    if world == 1 and not args.no_also and args.variant == "ff" and args.contact_model == "normal_1d":
        # the other throughput configurations of BASELINE.json, short runs (parity for them: tests/test_gpu_parity.py)
        also = {"config3": short_run(args, local_rank, "classical", "normal_1d", 4096, 10, 1.0),
                "config5": short_run(args, local_rank, "ff", "point3d", 16384, 20, 0.75)}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": args.scaling if world > 1 else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_config(args, B, nx),
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ok": e2e_ok,
                "call": "ffmpc_solve_batch_host_async x steps + ffmpc_wait (pinned host buffers; every step copies its inputs in and its "
                        "results out)",
                "blocking_call_value": total * args.steps / e2e_sync_s, "d2h_gbs_rank0": d2h_gbs_rank,
                "numa_rank0": numa},
        "gpu_launches": launches_per_step * args.steps,
        "clocks": clk, "roofline": roofline, "cpu_baseline": cpu, "latency": latency, "all_iters_ran": ok_dev,
    }
    if also is not None:
        line["also"] = also
    if strong is not None:
        # efficiency against the same GPU solving the WHOLE sweep alone = this run's weak per-GPU rate (value / n_gpus)
        strong["efficiency_vs_1gpu"] = strong["value"] / value
        line["strong"] = strong
    print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
