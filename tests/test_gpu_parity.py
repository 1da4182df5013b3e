"""GPU parity tests proper (`-m gpu`): the CUDA engine, called through the C ABI of include/ffmpc.h
(ffmpc_solve_batch_host / ffmpc_solve_batch / ffmpc_node_eval_host), against the CPU oracle on the same
seeded inputs, plus size-independent properties at BASELINE.json's full sizes.

Tolerances (fp64 everywhere; north_star: gains K/k, cost, gaps within 1e-8 relative at a fixed iteration count):
  * one action-model evaluation (calc + calcDiff):  1e-10 relative to the largest entry of each block
  * every solve (1 iteration, 10 iterations, plain FDDP, warm start, ragged batches, BASELINE sizes): tests/parity.py --
    1e-8 per instance against the fp64 oracle on K, k, cost, xs, us, gaps and the predicted contact force; an instance
    that misses it goes to the ARBITER, the same oracle source built in IEEE binary128, and must satisfy
    err(engine vs binary128) <= max(1e-8, 3 err(fp64 oracle vs binary128)).  `iters` / `ok` agree exactly.
    The ten-iteration test sends EVERY instance to the arbiter.  Measured (profiles/r2c_arbiter.txt): the engine is as
    close to binary128 as the fp64 oracle is, or closer, on all six variants; classical point3d is ill-conditioned for
    both (K: engine 1.0e-6, oracle 6.6e-7 after ten iterations).
"""
import numpy as np
import pytest

from ffmpc_cases import make_case, rel_err
from parity import check_vs_oracle, inst_rel

pytestmark = pytest.mark.gpu
NODE_KEYS = ("xnext", "cost", "force", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu")
VARIANTS = [(False, 1, 1.0), (False, 1, 0.0), (False, 3, 0.5), (True, 1, 1.0), (True, 1, 0.0), (True, 3, 0.5)]


def _eng(cfg, am, B, **opt):
    from franka_force_feedback_mpc_b200 import engine

    return engine.FFMPCEngine(cfg, am, device=0, max_batch=B, **opt)


_inst_rel = inst_rel


@pytest.mark.parametrize("ff,nc,cf", VARIANTS)
def test_node_calc_and_calcdiff_match_oracle(built, orc, ff, nc, cf):
    B, N = 64, 50
    cfg, am, inp = make_case(ff, nc, B, N, 31, cf)
    rng = np.random.default_rng(7)
    inst = np.repeat(np.arange(B), 4).astype(np.int32)
    knot = np.tile([0, 7, N - 1, N], B).astype(np.int32)
    x = inp.x0[inst] + rng.normal(0, 0.03, (inst.size, inp.nx))
    u = rng.normal(0, 3.0, (inst.size, 7)) + (inp.x0[inst][:, 14:21] if ff else 0.0)
    u[::5] *= 40.0  # beyond the soft torque limits
    eng = _eng(cfg, am, B)
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


@pytest.mark.parametrize("ff,nc,cf", VARIANTS)
def test_one_iteration_gains_cost_gaps(built, orc, ff, nc, cf):
    B, N = 48, 50
    cfg, am, inp = make_case(ff, nc, B, N, 5, cf)
    eng = _eng(cfg, am, B)
    out = eng.solve_host(inp, 1, True)
    eng.close()
    ref = orc.solve_batch(cfg, am, inp, 1, True, nthreads=8)
    check_vs_oracle(out, cfg, am, inp, 1, True, orc, ref=ref, label=(ff, nc, cf))
    assert _inst_rel(out["stop"], ref["stop"]).max() < 1e-7


@pytest.mark.parametrize("ff,nc,cf", VARIANTS)
def test_ten_fixed_iterations_match_oracle(built, orc, ff, nc, cf):
    """The headline bar: ten fixed Box-FDDP iterations, every instance judged by the binary128 arbiter."""
    B, N, iters = 64, 50, 10
    cfg, am, inp = make_case(ff, nc, B, N, 5, cf)
    eng = _eng(cfg, am, B)
    out = eng.solve_host(inp, iters, True)
    eng.close()
    rep = check_vs_oracle(out, cfg, am, inp, iters, True, orc, label=(ff, nc, cf), arbiter="always")
    if ff and nc == 1:
        # the headline path (BASELINE config 4) is well conditioned: 1e-8 against the fp64 oracle itself, every instance
        if not rep["post_convergence_instances"]:
            for k in ("K", "k", "cost", "xs", "us", "gaps"):
                assert rep[k][0] < 1e-8, (k, rep[k])


def test_line_search_modes_agree_bitwise(built):
    """Every line-search schedule (queue / speculative / grouped / tail, and what auto resolves to) takes the same decisions
    from the same rollouts: bit-identical results.  Then a batch large enough for auto to pick the tail schedule, with
    sub-batches, against the plain queue on one stream."""
    B, N, iters = 96, 50, 8
    cfg, am, inp = make_case(True, 1, B, N, 9, 0.8)
    inp.x0[: B // 2, :7] += np.random.default_rng(5).normal(0, 0.3, (B // 2, 7))   # hard starts: late step lengths occur
    res = {}
    for ls in ("queue", "spec", "grouped", "tail", "fill", "auto"):
        eng = _eng(cfg, am, B, line_search=ls)
        res[ls] = eng.solve_host(inp, iters, True)
        eng.close()
    for ls in ("spec", "grouped", "tail", "fill", "auto"):
        for k in ("xs", "us", "K", "k", "cost", "stop", "gaps", "iters", "ok", "f_pred"):
            assert np.array_equal(res["queue"][k], res[ls][k], equal_nan=True), (ls, k)
    B = 9000
    cfg, am, inp = make_case(True, 1, B, 20, 10, 0.7)
    res = {}
    for ls, sb in (("queue", 1), ("auto", 0), ("grouped", 3), ("fill", 2)):
        eng = _eng(cfg, am, B, line_search=ls, sub_batches=sb)
        res[ls] = eng.solve_host(inp, 5, True)
        eng.close()
    for ls in ("auto", "grouped", "fill"):
        for k in ("xs", "us", "K", "k", "cost", "stop", "gaps", "iters", "ok", "f_pred"):
            assert np.array_equal(res["queue"][k], res[ls][k], equal_nan=True), (ls, k)


def test_plain_fddp_and_early_exit(built, orc):
    """use_box_fddp = 0 (`solver="fddp"`, crocoddyl_classical.py:442-445) and the converged early exit."""
    B, N = 32, 30
    cfg, am, inp = make_case(True, 1, B, N, 17, 0.75)
    cfg.use_box_fddp = 0
    eng = _eng(cfg, am, B)
    a = eng.solve_host(inp, 6, True)
    check_vs_oracle(a, cfg, am, inp, 6, True, orc, label="plain FDDP")
    eng.close()
    cfg.use_box_fddp = 1
    eng = _eng(cfg, am, B)
    b = eng.solve_host(inp, 100, False)
    rb = orc.solve_batch(cfg, am, inp, 100, False, nthreads=8)
    eng.close()
    assert np.array_equal(b["ok"], rb["ok"]) and np.all(b["ok"] == 1)
    assert np.all(b["stop"] < cfg.th_stop)
    same = b["iters"] == rb["iters"]
    assert np.mean(same) >= 0.9  # the stop measure sits at the threshold for a few instances: one iteration more or less
    assert np.abs(b["iters"] - rb["iters"]).max() <= 1
    assert _inst_rel(b["cost"][same], rb["cost"][same]).max() < 1e-8
    assert _inst_rel(b["cost"], rb["cost"]).max() < 1e-6


def test_warm_start_inputs_and_given_tau_ref(built, orc):
    from franka_force_feedback_mpc_b200.batch import BatchInputs

    B, N = 32, 40
    cfg, am, inp = make_case(False, 1, B, N, 23, 1.0)
    eng = _eng(cfg, am, B)
    a = eng.solve_host(inp, 4, True)
    # `_shift_guess` (crocoddyl_classical.py:733-757): xs = [x0] + xs_prev[1:], us = us_prev[1:] + [us_prev[-1]]
    xs_init = a["xs"].copy(); xs_init[:, 0] = inp.x0
    us_init = np.concatenate([a["us"][:, 1:], a["us"][:, -1:]], axis=1)
    tau_ref = np.random.default_rng(0).normal(0, 1.0, (B, 7))
    warm = BatchInputs(14, N, inp.mode, inp.x0, inp.p_ref, inp.v_ref, inp.x_reg_ref, tau_ref=tau_ref, xs_init=xs_init, us_init=us_init)
    b = eng.solve_host(warm, 3, True)
    eng.close()
    check_vs_oracle(b, cfg, am, warm, 3, True, orc, label="warm start + given tau_ref")


def test_device_pointer_api_equals_host_api(built):
    """ffmpc_solve_batch (device pointers from torch tensors, caller's stream) == ffmpc_solve_batch_host, bit for bit."""
    import torch

    B, N = 128, 50
    cfg, am, inp = make_case(True, 1, B, N, 3, 0.75)
    eng = _eng(cfg, am, B)
    h = eng.solve_host(inp, 5, True)
    dev = torch.device("cuda", 0)
    t_in = {k: torch.from_numpy(getattr(inp, k)).to(dev) for k in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref")}
    st = torch.cuda.Stream(device=dev)
    d = eng.solve_device(t_in, 5, True, stream=st)
    st.synchronize()
    eng.close()
    for k in ("xs", "us", "K", "k", "cost", "stop", "gaps", "iters", "ok"):
        assert np.array_equal(d[k].cpu().numpy(), h[k]), k


def test_ragged_and_edge_batches(built, orc):
    """B = 1, a batch that is not a multiple of any block size, all-free / all-contact / mixed modes."""
    N = 20
    for B, cf in ((1, 1.0), (1, 0.0), (37, 0.5), (129, 1.0)):
        cfg, am, inp = make_case(True, 1, B, N, 11 + B, cf)
        eng = _eng(cfg, am, max(B, 200))  # handle larger than the batch
        out = eng.solve_host(inp, 3, True)
        eng.close()
        check_vs_oracle(out, cfg, am, inp, 3, True, orc, label=("ragged", B, cf))


def test_errors_through_the_abi(built):
    from franka_force_feedback_mpc_b200 import engine

    cfg, am, inp = make_case(False, 1, 8, 10, 1, 1.0)
    eng = _eng(cfg, am, 4)
    with pytest.raises(engine.FFMPCError) as e:
        eng.solve_host(inp, 2, True)
    assert "max_batch" in str(e.value)
    eng.close()
    with pytest.raises(engine.FFMPCError):
        engine.FFMPCEngine(cfg, am, device=99, max_batch=4)


# ---------------------------------------------------------------------------------------------------
# BASELINE.json full sizes: properties that need no oracle run
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("ff,nc,B,iters,cf", [(False, 1, 4096, 10, 1.0),      # config 3
                                               (True, 1, 16384, 10, 1.0),      # one scenario of config 4 (4 x 16384)
                                               (True, 3, 16384, 20, 0.75)])    # config 5: point3d, 25 % free / 75 % contact, 20 iterations
def test_full_size_properties(built, orc, ff, nc, B, iters, cf):
    N = 50
    cfg, am, inp = make_case(ff, nc, B, N, 1000 * 3 + 11, cf, scenario="actuation_uncertainty" if ff else "flat")
    eng = _eng(cfg, am, B)
    want = ("xs", "us", "K", "k", "cost", "gaps", "iters", "ok", "stop", "f_pred")
    a = eng.solve_host(inp, iters, True, want=want)
    # determinism: the same call twice is bit-identical
    a2 = eng.solve_host(inp, iters, True, want=want)
    for k in want:
        assert np.array_equal(a[k], a2[k], equal_nan=True), k
    # permutation equivariance: instances are independent, results do not depend on batch position or neighbours
    perm = np.random.default_rng(0).permutation(B)
    p = eng.solve_host(inp.slice(perm), iters, True, want=want)
    for k in ("xs", "us", "cost"):
        assert np.array_equal(p[k], a[k][perm]), k
    # fewer iterations never give a lower cost (monotone descent of accepted steps)
    h = eng.solve_host(inp, iters // 2, True, want=want)
    eng.close()
    # (while the trajectory still has gaps FDDP may accept steps that raise the cost, th_acceptnegstep; from the
    #  first gap-free candidate on every accepted step must lower it)
    feas_half = np.abs(h["gaps"]).reshape(B, -1).max(axis=1) < 1e-9
    assert feas_half.mean() > 0.9
    assert np.all((a["cost"] <= h["cost"] + 1e-9 * np.abs(h["cost"]))[feas_half])
    assert np.all(np.isfinite(a["cost"])) and np.array_equal(a["iters"], np.full(B, iters))
    assert np.all(np.abs(a["us"]) <= np.array(cfg.tau_limits) + 1e-12)
    assert np.allclose(a["xs"][:, 0], inp.x0, atol=0, rtol=0) or np.abs(a["gaps"][:, 0]).max() > 0
    closed = np.abs(a["gaps"]).reshape(B, -1).max(axis=1) < 1e-9
    assert closed.mean() > 0.9  # the rest has not accepted a full step yet (oracle agrees, see the sample below)
    # 256 instances drawn from the full batch: gains K / k, cost, gaps, trajectories against the oracle (north_star's bar)
    idx = np.sort(np.random.default_rng(1).choice(B, 256, replace=False))
    check_vs_oracle({k: a[k][idx] for k in want}, cfg, am, inp.slice(idx), iters, True, orc, label=("full size", ff, nc, B))


def test_against_crocoddyl_golden(built):
    """Parity with REAL Crocoddyl, when someone with the reference's conda environment has produced
    tests/golden/crocoddyl_golden.npz with tools/export_crocoddyl_golden.py.  Neither crocoddyl nor pinocchio can be
    installed in this image or on the GPU box, so here this reports: Crocoddyl golden unavailable on this box.
    (The consumer itself is exercised on the CPU with an oracle-made file of the same format: tests/test_emulator.py.)"""
    import os

    from ffmpc_cases import check_against_golden

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "crocoddyl_golden.npz")
    if not os.path.exists(path):
        pytest.skip("Crocoddyl golden: unavailable on this box (needs the reference's conda env; see tools/export_crocoddyl_golden.py)")
    check_against_golden(np.load(path, allow_pickle=False), _eng)


@pytest.mark.parametrize("ff", [False, True])
def test_friction_cone_point3d_matches_oracle(built, orc, ff):
    """SURVEY.md 8f4 (ResidualModelContactFrictionCone, crocoddyl_classical.py:678-686): knot quantities and one FDDP
    iteration with the cone on and violated (mu = 0.3), against the oracle; relative tolerances as for the other
    variants (1e-10 per knot, 1e-8 after one iteration unless the oracle's own rounding sensitivity is larger)."""
    B, N = 64, 50
    cfg, am, inp = make_case(ff, 3, B, N, 31, 0.75, w_friction_cone=2.0e2, mu=0.3)
    rng = np.random.default_rng(7)
    inst = np.repeat(np.arange(B), 4).astype(np.int32)
    knot = np.tile([0, 7, N - 1, N], B).astype(np.int32)
    x = inp.x0[inst] + rng.normal(0, 0.03, (inst.size, inp.nx))
    u = rng.normal(0, 3.0, (inst.size, 7)) + (inp.x0[inst][:, 14:21] if ff else 0.0)
    eng = _eng(cfg, am, B)
    out = eng.node_eval(inp, inst, knot, x, u)
    ref = orc.node_eval(cfg, am, inp, inst, knot, x, u)
    cfg0, _, _ = make_case(ff, 3, B, N, 31, 0.75)
    assert np.abs(ref["Lxx"] - orc.node_eval(cfg0, am, inp, inst, knot, x, u)["Lxx"]).max() > 1.0   # the term is live
    contact = inp.mode[inst] == 1
    for k in NODE_KEYS:
        a, b = (out[k][contact], ref[k][contact]) if k == "force" else (out[k], ref[k])
        assert rel_err(a, b) < 1e-10, k
    got = eng.solve_host(inp, 1, True)
    eng.close()
    check_vs_oracle(got, cfg, am, inp, 1, True, orc, label=("friction cone", ff))


def test_host_call_split_is_bitwise_neutral(built):
    """ffmpc_solve_batch_host overlaps the device->host copies of a chunk with the solve of the next one
    (ffmpc_options.sub_batches): outputs must be bit-identical to the unsplit call."""
    B, N = 2048, 50
    cfg, am, inp = make_case(True, 1, B, N, 9, 0.75)
    eng = _eng(cfg, am, B, sub_batches=1)
    whole = eng.solve_host(inp, 5, True)
    eng.close()
    eng = _eng(cfg, am, B, sub_batches=5)   # 2048 -> 410 x 4 + 408
    split = eng.solve_host(inp, 5, True)
    eng.close()
    for k in whole:
        assert np.array_equal(whole[k], split[k], equal_nan=True), k


def test_pipelined_async_host_calls_equal_blocking_calls(built):
    """ffmpc_solve_batch_host_async: three calls in flight on the sub-batch streams (results leave on the copy streams while
    the next call is being solved, export staging reused only after the previous copy): each call's outputs are bit-identical
    to the blocking call on the same inputs.  Pinned buffers (the copies are then truly asynchronous)."""
    import torch
    from franka_force_feedback_mpc_b200.batch import alloc_outputs, outputs_abi

    B, N, iters = 4096, 30, 4
    cases = [make_case(True, 1, B, N, 40 + i, 0.6)[2] for i in range(3)]
    cfg, am, _ = make_case(True, 1, B, N, 40, 0.6)
    eng = _eng(cfg, am, B, sub_batches=4)
    want = ("xs", "us", "K", "k", "cost", "iters", "ok", "f_pred")
    ref = [eng.solve_host(c, iters, True, want=want) for c in cases]
    pinned_in, s_ins, outs, s_outs = [], [], [], []
    for c in cases:
        t = {k: torch.from_numpy(getattr(c, k)).pin_memory() for k in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref")}
        s = _abi_in(B, t)
        o = {k: torch.from_numpy(v).pin_memory() for k, v in alloc_outputs(B, N, 21, want).items()}
        pinned_in.append(t); s_ins.append(s); outs.append(o)
        s_outs.append(outputs_abi({k: v.numpy() for k, v in o.items()}))
    for rep in range(2):   # the second round reuses every staging range and every host buffer
        for o in outs:
            for v in o.values():
                v.zero_()
        for i in range(3):
            eng.solve_host_ptrs_async(s_ins[i], s_outs[i], iters, True)
        eng.wait()
        for i in range(3):
            for k in want:
                assert np.array_equal(outs[i][k].numpy(), ref[i][k], equal_nan=True), (rep, i, k)
    eng.close()


def _abi_in(B, tensors):
    from franka_force_feedback_mpc_b200 import _abi

    s = _abi.BatchIn()
    s.B = B
    for k, t in tensors.items():
        setattr(s, k, t.data_ptr())
    return s
