#!/usr/bin/env python
"""Dump REAL Crocoddyl / Pinocchio results for the engine's parity tests ("how parity with real
Crocoddyl gets closed").  Runs ONLY where the reference's conda environment exists (crocoddyl, pinocchio,
example_robot_data: `environment.yml:15-17`); none of them is installable in the build container or on the GPU box,
so this script has never been executed there ("Crocoddyl golden: unavailable on this box").

    python tools/export_crocoddyl_golden.py --reference /path/to/Franka-Force-Feedback-MPC --out tests/golden/crocoddyl_golden.npz

What it writes, for the seeded synthetic instances of `synthetic.sample_instances` (same seeds as the tests):
  * the reduced Panda model constants as Pinocchio sees them (joint placements, inertias, frame `panda_link8`),
    to be compared with `panda.default_model()` (the MuJoCo-XML-derived constants the engine uses);
  * per instance and knot, BEFORE any solve: `problem.calc` + `problem.calcDiff` at the cold start (xs_init, us_init):
    xnext, cost, Fx, Fu, Lx, Lu, Lxx, Lxu, Luu -- a mismatch in a solve can then be localised to a knot
    quantity (the consumer checks the knots first);
  * per instance: xs, us, K, k, cost, iter after `--iters` iterations of the reference's own
    `_build_problem` + `SolverBoxFDDP.solve(xs_init, us_init, iters, False)` with the convergence test disabled
    (th_stop = 0), classical and force-feedback, contact and free mode.
`tests/test_gpu_parity.py::test_against_crocoddyl_golden` consumes the file when it is present.
"""
from __future__ import annotations

import argparse
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def bare_controller(ref_mod, cls_name, cfg, q_nom, R_des, dt=0.005):
    """Reference controller without MuJoCo: the constructor's Pinocchio part is replayed, the sim calibration is
    replaced by identity site offsets (the engine's synthetic references are already in the Pinocchio world)."""
    import crocoddyl
    import pinocchio as pin
    from example_robot_data import load as load_erd

    cls = getattr(ref_mod, cls_name)
    c = object.__new__(cls)
    c.cfg, c.sim, c._k, c._warned_keys = cfg, types.SimpleNamespace(dt=dt), 0, set()
    full = load_erd("panda").model
    c.model = cls._make_arm_only_model(c, full)
    c.ee_fid = c.model.getFrameId("panda_link8")
    c.data = c.model.createData()
    c.state = crocoddyl.StateMultibody(c.model)
    c.actuation = crocoddyl.ActuationModelFull(c.state)
    c.nx_mb, c.ndx_mb, c.nx_aug = int(c.state.nx), int(c.state.ndx), int(c.state.nx + c.actuation.nu)
    c.R_mj_from_pin = np.eye(3)            # references are passed in the Pinocchio world already
    c.R_site_from_pin_ee = np.eye(3)
    c.p_site_minus_frame_pin = np.zeros(3)
    c.q_nom, c.R_des = np.asarray(q_nom, float), np.asarray(R_des, float)
    c.xs = c.us = c.Ks = c.ks = None
    c._tau_prev = np.zeros(7)
    return c, pin


def model_constants(model, pin):
    out = {"joint_pos": [], "joint_rot": [], "mass": [], "com": [], "inertia": []}
    for j in range(1, model.njoints):
        M, Y = model.jointPlacements[j], model.inertias[j]
        out["joint_pos"].append(np.array(M.translation))
        out["joint_rot"].append(np.array(M.rotation))
        out["mass"].append(Y.mass)
        out["com"].append(np.array(Y.lever))
        I = np.array(Y.inertia)
        out["inertia"].append(np.array([I[0, 0], I[1, 1], I[2, 2], I[0, 1], I[0, 2], I[1, 2]]))   # PandaModel order: xx yy zz xy xz yz
    f = model.frames[model.getFrameId("panda_link8")]
    out["ee_pos"], out["ee_rot"] = np.array(f.placement.translation), np.array(f.placement.rotation)
    out["q_lower"], out["q_upper"] = np.array(model.lowerPositionLimit), np.array(model.upperPositionLimit)
    out["gravity"] = np.array(model.gravity.linear)
    return {k: np.array(v, dtype=float) for k, v in out.items()}


KNOT_FIELDS = ("xnext", "cost", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu")


def knot_quantities(problem, xs0, us0):
    """ActionData of every knot after problem.calc + calcDiff at (xs0, us0); the terminal knot has no control blocks."""
    problem.calc(xs0, us0)
    problem.calcDiff(xs0, us0)
    datas = list(problem.runningDatas) + [problem.terminalData]
    nx, nu = len(xs0[0]), len(us0[0])
    res = {}
    for f in KNOT_FIELDS:
        rows = []
        for t, d in enumerate(datas):
            v = np.array(getattr(d, f), dtype=float)
            if t == len(datas) - 1 and f in ("Fu", "Lu", "Lxu", "Luu"):   # terminal model: nu-dependent blocks may be empty
                shape = {"Fu": (nx, nu), "Lu": (nu,), "Lxu": (nx, nu), "Luu": (nu, nu)}[f]
                v = np.zeros(shape) if v.size != int(np.prod(shape)) else v.reshape(shape)
            rows.append(v)
        res[f] = np.array(rows)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", required=True, help="checkout of omar-a-aman/Franka-Force-Feedback-MPC")
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "crocoddyl_golden.npz"))
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--horizon", type=int, default=50)
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    sys.path.insert(0, args.reference)
    from franka_force_feedback_mpc_b200 import config, panda, synthetic
    from src.mpc import crocoddyl_classical as rc, crocoddyl_force_feedback as rf

    mdl = panda.default_model()
    R_des = panda.reference_R_des(mdl)
    out = {}
    for tag, ff, mod, cls_name, preset in (("classical", False, rc, "ClassicalCrocoddylMPC", config.benchmark_classical_config),
                                           ("ff", True, rf, "ForceFeedbackCrocoddylMPC", config.benchmark_ff_config)):
        ours = preset(horizon=args.horizon, max_iters=args.iters)
        ref_cfg_cls = getattr(mod, "ClassicalMPCConfig" if not ff else "ForceFeedbackMPCConfig")
        ref_cfg = ref_cfg_cls(**{k: v for k, v in vars(ours).items() if k in ref_cfg_cls.__dataclass_fields__})
        ctl, pin = bare_controller(mod, cls_name, ref_cfg, panda.Q_NEUTRAL, R_des)
        if not ff:
            for k, v in model_constants(ctl.model, pin).items():
                out[f"pin_{k}"] = v
        s = synthetic.sample_instances(args.batch, args.horizon, ff, seed=5, contact_fraction=0.75)
        res = {k: [] for k in ("xs", "us", "K", "k", "cost", "iter")}
        knots = {f: [] for f in KNOT_FIELDS}
        for b in range(args.batch):
            refs = [(s["p_ref"][b, k], s["v_ref"][b, k], bool(s["mode"][b])) for k in range(args.horizon + 1)]
            ctl.traj_fn = lambda t, refs=refs: refs[min(int(round(t / 0.01)), args.horizon)]
            x0 = s["x0"][b]
            problem = ctl._build_problem(0.0, x0, bool(s["mode"][b]))
            solver = ctl._make_solver(problem)
            solver.th_stop = 0.0   # fixed iteration count
            ctl.xs = ctl.us = None
            ctl._tau_prev = np.asarray(pin.rnea(ctl.model, ctl.data, x0[:7], np.zeros(7), np.zeros(7)))
            xs0, us0 = ctl._shift_guess(x0, args.horizon)
            for f, v in knot_quantities(problem, xs0, us0).items():
                knots[f].append(v)
            solver.solve(xs0, us0, args.iters, False)
            res["xs"].append(np.array(solver.xs)); res["us"].append(np.array(solver.us))
            res["K"].append(np.array(solver.K)); res["k"].append(np.array(solver.k))
            res["cost"].append(solver.cost); res["iter"].append(solver.iter)
        for k, v in res.items():
            out[f"{tag}_{k}"] = np.array(v)
        for f, v in knots.items():
            out[f"{tag}_knot_{f}"] = np.array(v)        # [B, N+1, ...] at the cold start
        for k in ("mode", "x0", "p_ref", "v_ref", "x_reg_ref"):
            out[f"{tag}_{k}"] = s[k]
    np.savez_compressed(args.out, **out)
    print("wrote", args.out)


if __name__ == "__main__":
    main()
