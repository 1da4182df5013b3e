"""CPU CI for the CUDA sources: tests/host_emul compiles the SAME kernel sources (ffmpc_kernels.cuh /
ffmpc_node.cuh / ffmpc_engine.cu) against a fiber-based CUDA emulator (grid/block/warp shuffles/shared
memory) and this file compares that build with the oracle -- so a logic error in a kernel shows up in
`-m "not gpu"` before any GPU time is spent.  The emulator library is test-only and never shipped; the
GPU parity tests proper are tests/test_gpu_parity.py."""
import os
import subprocess

import numpy as np
import pytest

from ffmpc_cases import CONFIG_CASES, make_case, rel_err, run_config_case

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_DIR = os.path.join(HERE, "host_emul")
NODE_KEYS = ("xnext", "cost", "force", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu")


@pytest.fixture(scope="module")
def emu():
    subprocess.check_call(["make", "-C", EMU_DIR, "all"], stdout=subprocess.DEVNULL)
    from franka_force_feedback_mpc_b200 import engine

    lib = engine.load_library(os.path.join(EMU_DIR, "libffmpc_emu.so"))
    assert b"EMULATION" in lib.ffmpc_version()
    return lib


def _engine(emu, cfg, am, B, **opt):
    from franka_force_feedback_mpc_b200 import engine

    return engine.FFMPCEngine(cfg, am, device=0, max_batch=B, _lib=emu, **opt)


@pytest.mark.parametrize("ff", [False, True])
@pytest.mark.parametrize("nc,cf", [(1, 1.0), (1, 0.0), (3, 0.5)])
def test_emulated_node_kernels_match_oracle(emu, orc, ff, nc, cf):
    B, N = 6, 8
    cfg, am, inp = make_case(ff, nc, B, N, 31, cf)
    rng = np.random.default_rng(7)
    inst = np.repeat(np.arange(B), 3).astype(np.int32)
    knot = np.tile([0, 3, N], B).astype(np.int32)
    x = inp.x0[inst] + rng.normal(0, 0.02, (inst.size, inp.nx))
    u = rng.normal(0, 3.0, (inst.size, 7)) + (inp.x0[inst][:, 14:21] if ff else 0.0)
    u[::5] *= 40.0  # some controls beyond the soft limits
    eng = _engine(emu, cfg, am, B)
    out = eng.node_eval(inp, inst, knot, x, u)
    eng.close()
    ref = orc.node_eval(cfg, am, inp, inst, knot, x, u)
    contact = inp.mode[inst] == 1
    for k in NODE_KEYS:
        a, b = out[k], ref[k]
        if k == "force":
            a, b = a[contact], b[contact]
            if a.size == 0:
                continue
        assert rel_err(a, b) < 1e-10, k


# line search (ffmpc_options.line_search): the schedule of step-length groups; "queue" = one step length per pass with
# compaction, "spec" = every step length rolled out at once + k_forward_finalize, "grouped" / "tail" = mixtures
@pytest.mark.parametrize("ff,nc,cf,box,mode", [(False, 1, 1.0, 1, "queue"), (False, 1, 0.0, 1, "spec"), (True, 1, 0.5, 1, "queue"),
                                               (True, 1, 0.5, 1, "spec"), (True, 3, 1.0, 1, "spec"), (True, 1, 1.0, 0, "queue"),
                                               (True, 1, 0.5, 1, "grouped"), (False, 3, 0.5, 1, "tail")])
def test_emulated_solver_matches_oracle(emu, orc, ff, nc, cf, box, mode):
    B, N, iters = 5, 12, 6
    cfg, am, inp = make_case(ff, nc, B, N, 5, cf)
    cfg.use_box_fddp = box
    eng = _engine(emu, cfg, am, B, line_search=mode)
    out = eng.solve_host(inp, iters, True)
    eng.close()
    ref = orc.solve_batch(cfg, am, inp, iters, True)
    assert np.array_equal(out["iters"], ref["iters"]) and np.array_equal(out["ok"], ref["ok"])
    for k in ("xs", "us", "K", "k", "cost", "stop"):
        assert rel_err(out[k], ref[k]) < 1e-6, k
    assert np.abs(out["gaps"] - ref["gaps"]).max() < 1e-9
    c = inp.mode == 1
    if c.any():
        assert rel_err(out["f_pred"][c], ref["f_pred"][c]) < 1e-6


def test_emulated_line_search_schedules_agree_bitwise(emu):
    """Every schedule takes the same decisions from the same rollouts: bit-identical results (a hard start, so that late
    step lengths and rejected iterations occur)."""
    B, N, iters = 6, 10, 5
    cfg, am, inp = make_case(True, 1, B, N, 23, 0.5)
    inp.x0[:, :7] += np.random.default_rng(3).normal(0, 0.4, (B, 7))   # far from the references
    res = {}
    for ls in ("queue", "spec", "grouped", "tail", "fill", "auto"):
        eng = _engine(emu, cfg, am, B, line_search=ls)
        res[ls] = eng.solve_host(inp, iters, True)
        info = eng.info()
        assert info["trial_slots"] == {"queue": 1, "spec": 10, "grouped": 7, "tail": 5, "fill": 10, "auto": 10}[ls]
        eng.close()
    for ls in ("spec", "grouped", "tail", "fill", "auto"):
        for k in ("xs", "us", "K", "k", "cost", "stop", "gaps", "iters", "ok", "f_pred"):
            assert np.array_equal(res["queue"][k], res[ls][k], equal_nan=True), (ls, k)


def test_emulated_early_exit_and_warm_start(emu, orc):
    from franka_force_feedback_mpc_b200.batch import BatchInputs

    B, N = 4, 10
    cfg, am, inp = make_case(True, 1, B, N, 17, 1.0)
    eng = _engine(emu, cfg, am, B)
    a = eng.solve_host(inp, 50, False)
    ra = orc.solve_batch(cfg, am, inp, 50, False)
    assert np.array_equal(a["iters"], ra["iters"]) and np.array_equal(a["ok"], ra["ok"])
    assert np.all(a["ok"] == 1) and np.all(a["stop"] < cfg.th_stop)
    warm = BatchInputs(inp.nx, N, inp.mode, inp.x0, inp.p_ref, inp.v_ref, inp.x_reg_ref, xs_init=a["xs"], us_init=a["us"],
                       tau_ref=np.zeros((B, 7)))
    b = eng.solve_host(warm, 3, True)
    rb = orc.solve_batch(cfg, am, warm, 3, True)
    eng.close()
    for k in ("xs", "us", "cost"):
        assert rel_err(b[k], rb[k]) < 1e-7, k


def test_emulated_capacity_and_argument_errors(emu):
    from franka_force_feedback_mpc_b200 import engine

    cfg, am, inp = make_case(False, 1, 4, 6, 1, 1.0)
    eng = _engine(emu, cfg, am, 2)
    with pytest.raises(engine.FFMPCError) as e:
        eng.solve_host(inp, 2, True)
    assert "max_batch" in str(e.value)
    eng.close()


@pytest.mark.parametrize("ff", [False, True])
def test_emulated_friction_cone_matches_oracle(emu, orc, ff):
    """point3d with w_friction_cone > 0 (mu small enough that facets are violated): knot quantities
    against the oracle and a short solve."""
    from franka_force_feedback_mpc_b200 import engine

    B, N = 6, 8
    cfg, am, inp = make_case(ff, 3, B, N, 31, 0.8, w_friction_cone=2.0e2, mu=0.3)
    rng = np.random.default_rng(7)
    inst = np.repeat(np.arange(B), 3).astype(np.int32)
    knot = np.tile([0, 3, N], B).astype(np.int32)
    x = inp.x0[inst] + rng.normal(0, 0.02, (inst.size, inp.nx))
    u = rng.normal(0, 3.0, (inst.size, 7)) + (inp.x0[inst][:, 14:21] if ff else 0.0)
    eng = _engine(emu, cfg, am, B)
    out = eng.node_eval(inp, inst, knot, x, u)
    ref = orc.node_eval(cfg, am, inp, inst, knot, x, u)
    cfg0, _, _ = make_case(ff, 3, B, N, 31, 0.8)
    assert np.abs(ref["Lxx"] - orc.node_eval(cfg0, am, inp, inst, knot, x, u)["Lxx"]).max() > 1.0   # the term is live
    for k in NODE_KEYS:
        assert rel_err(out[k], ref[k]) < 1e-10, k
    if ff:   # (the classical point3d solve amplifies 1-ulp differences to 1e-5 within a few iterations, with or without the cone)
        got, want = eng.solve_host(inp, 5, True), orc.solve_batch(cfg, am, inp, 5, True)
        assert np.array_equal(got["iters"], want["iters"]) and np.array_equal(got["ok"], want["ok"])
        for k in ("xs", "us", "K", "k", "cost"):
            assert rel_err(got[k], want[k]) < 1e-6, k
    eng.close()


def test_emulated_host_call_split_is_bitwise_neutral(emu):
    """ffmpc_solve_batch_host solves large batches in chunks so that a chunk's device->host copies overlap the next
    chunk's solve (ffmpc_options.sub_batches): the outputs must not depend on the split (mixed modes, so that the mode
    buckets of the chunks differ from those of the whole batch)."""
    B, N = 20, 8
    cfg, am, inp = make_case(True, 1, B, N, 9, 0.6)
    eng = _engine(emu, cfg, am, B, sub_batches=1)
    whole = eng.solve_host(inp, 4, True)
    eng.close()
    eng = _engine(emu, cfg, am, B, sub_batches=3)      # 20 instances -> 7 + 7 + 6
    split = eng.solve_host(inp, 4, True)
    eng.close()
    for k in whole:
        assert np.array_equal(whole[k], split[k], equal_nan=True), k


@pytest.mark.parametrize("case", CONFIG_CASES, ids=[c[0] for c in CONFIG_CASES])
def test_emulated_configuration_matches_oracle(emu, orc, case):
    """tests/test_gpu_configs.py at a small size on the emulator build: inner-regularisation flags, reference modes,
    development presets, dataclass defaults, benchmark horizons."""
    name, ff, nc, cf, horizon, kw = case
    small = (name, ff, nc, cf, 6 if horizon else None, kw)   # the horizon cases only differ by N: keep them cheap here
    run_config_case(lambda cfg, am, B: _engine(emu, cfg, am, B), orc, small, B=3, N=6, iters=2)


def test_golden_consumer_on_an_oracle_made_file(emu, orc):
    """tests/test_gpu_parity.py::test_against_crocoddyl_golden consumes what tools/export_crocoddyl_golden.py writes on a
    machine that has Crocoddyl; that file does not exist on this box.  Here the consumer itself is exercised: a file of the
    SAME format made by the oracle (knot quantities at the cold start, solves), with a model block that differs from the
    default model in the link-7 inertia, so that the report-and-adopt path of the model comparison runs too."""
    import warnings

    from ffmpc_cases import GOLDEN_KNOT_FIELDS, check_against_golden
    from franka_force_feedback_mpc_b200 import config, panda
    from franka_force_feedback_mpc_b200.batch import BatchInputs
    from franka_force_feedback_mpc_b200.synthetic import batched_gravity_torque, sample_instances

    B, N, iters = 3, 5, 2
    m = panda.default_model()
    m.mass = m.mass.copy(); m.mass[6] += 0.73           # "hand merged into link 7"
    g = {f"pin_{k}": np.asarray(getattr(m, k), float) for k in ("joint_pos", "joint_rot", "mass", "com", "inertia", "ee_pos", "ee_rot", "q_lower", "q_upper", "gravity")}
    for tag, ff in (("classical", False), ("ff", True)):
        preset = config.benchmark_ff_config if ff else config.benchmark_classical_config
        cfg = config.to_abi(preset(horizon=N, max_iters=iters), m, panda.Q_NEUTRAL)
        s = sample_instances(B, N, ff, seed=5, contact_fraction=0.75)
        nx = 21 if ff else 14
        inp = BatchInputs.from_samples(s, nx, N)
        inst = np.repeat(np.arange(B), N + 1).astype(np.int32)
        knot = np.tile(np.arange(N + 1), B).astype(np.int32)
        u0 = inp.x0[:, 14:21] if ff else batched_gravity_torque(m, inp.x0[:, :7])
        kn = orc.node_eval(cfg, m.to_abi(), inp, inst, knot, np.repeat(inp.x0, N + 1, axis=0), np.repeat(u0, N + 1, axis=0))
        for f in GOLDEN_KNOT_FIELDS:
            g[f"{tag}_knot_{f}"] = kn[f].reshape((B, N + 1) + kn[f].shape[1:])
        sol = orc.solve_batch(cfg, m.to_abi(), inp, iters, True)
        for k in ("xs", "us", "K", "k", "cost"):
            g[f"{tag}_{k}"] = sol[k]
        g[f"{tag}_iter"] = sol["iters"]
        for k in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref"):
            g[f"{tag}_{k}"] = s[k]
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        differing = check_against_golden(g, lambda cfg, am, B: _engine(emu, cfg, am, B), tol=1e-7)
    assert differing == ["mass"] and any("mass" in str(x.message) for x in w)


def test_emulated_async_host_pipeline_equals_blocking_calls(emu):
    """ffmpc_solve_batch_host_async + ffmpc_wait with several sub-batches (copy streams, export staging reuse, the 1/S start
    offset of a new pipeline): same bits as the blocking call.  (The emulator runs streams in order; this covers the host
    logic of the entry point, the GPU test covers the concurrency.)"""
    from franka_force_feedback_mpc_b200.batch import alloc_outputs, outputs_abi

    B, N, iters = 6, 8, 3
    cfg, am, inp = make_case(True, 1, B, N, 31, 0.5)
    eng = _engine(emu, cfg, am, B, sub_batches=3)
    want = ("xs", "us", "K", "k", "cost", "iters", "ok", "f_pred")
    ref = eng.solve_host(inp, iters, True, want=want)
    outs = [alloc_outputs(B, N, 21, want) for _ in range(2)]
    s_in = inp.to_abi()
    for o in outs:
        eng.solve_host_ptrs_async(s_in, outputs_abi(o), iters, True)
    eng.wait()
    eng.solve_host_ptrs_async(s_in, outputs_abi(outs[0]), iters, True)   # a new pipeline after the wait
    eng.wait()
    eng.close()
    for o in outs:
        for k in want:
            assert np.array_equal(o[k], ref[k], equal_nan=True), k
