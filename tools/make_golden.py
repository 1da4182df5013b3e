#!/usr/bin/env python
"""Generates tests/golden/*.npz by running the REFERENCE's own Python (imported from /root/reference,
never copied) for the pieces of the hot path that are reference source rather than Crocoddyl:

  lpf_golden.npz   _AugmentedLPFActionModel.calc / calcDiff  (src/mpc/crocoddyl_force_feedback.py:149-290)
                   wrapped around an inner action model whose outputs come from the CPU oracle's
                   classical knot (the inner Crocoddyl model itself is not installable here)
  traj_golden.npz  make_approach_then_circle + the runners' 0.2 s hold wrapper
                   (src/tasks/trajectories.py:8-93, src/run/run_classical.py:235-255)
  misc_golden.npz  _ff_alpha_ocp / _policy_epsilon (crocoddyl_force_feedback.py:493-510)

crocoddyl / pinocchio / example_robot_data are not installed: tiny stub modules are injected into
sys.modules so that the reference files import; only pure-numpy reference code is executed.

Run in the build container (needs /root/reference):   python tools/make_golden.py
The fixtures are committed; the GPU box never runs this script.
"""
from __future__ import annotations

import ctypes as C
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("FFMPC_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLD = os.path.join(ROOT, "tests", "golden")


def install_stubs():
    """Fake third-party modules: just enough surface for the reference files to import."""
    croc = types.ModuleType("crocoddyl")

    class StateVector:
        def __init__(self, nx):
            self.nx = int(nx)
            self.ndx = int(nx)

    class ActionModelAbstract:
        def __init__(self, state, nu, nr=0):
            self.state, self.nu, self.nr = state, int(nu), int(nr)

    class ActionDataAbstract:
        def __init__(self, model):
            self.cost = 0.0
            self.xnext = None

    croc.StateVector = StateVector
    croc.ActionModelAbstract = ActionModelAbstract
    croc.ActionDataAbstract = ActionDataAbstract
    pin = types.ModuleType("pinocchio")
    erd = types.ModuleType("example_robot_data")
    erd.load = lambda *a, **k: (_ for _ in ()).throw(RuntimeError("example_robot_data stub"))
    mj = types.ModuleType("mujoco")
    mj.viewer = types.ModuleType("mujoco.viewer")
    for name, mod in (("crocoddyl", croc), ("pinocchio", pin), ("example_robot_data", erd), ("mujoco", mj),
                      ("mujoco.viewer", mj.viewer)):
        sys.modules.setdefault(name, mod)


class OracleInner:
    """Inner IntegratedActionModelEuler stand-in: outputs of the oracle's classical knot."""

    def __init__(self, orc, cfg_inner, am, inp_inner, inst, knot):
        self.orc, self.cfg, self.am, self.inp, self.inst, self.knot = orc, cfg_inner, am, inp_inner, inst, knot

    def createData(self):
        return types.SimpleNamespace()

    def _eval(self, data, x, u):
        o = self.orc.node_eval(self.cfg, self.am, self.inp, [self.inst], [self.knot], x[None], u[None])
        for k in ("xnext", "cost", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"):
            setattr(data, k, o[k][0])
        data.cost = float(o["cost"][0])

    def calc(self, data, x, u):
        self._eval(data, np.asarray(x, float), np.asarray(u, float))

    def calcDiff(self, data, x, u):
        self._eval(data, np.asarray(x, float), np.asarray(u, float))


def gen_lpf():
    from ffmpc_cases import make_case
    from franka_force_feedback_mpc_b200 import _abi
    from franka_force_feedback_mpc_b200.batch import BatchInputs
    from oracle import oracle as orc
    from src.mpc.crocoddyl_force_feedback import _AugmentedLPFActionModel

    N, B, seed = 12, 6, 101
    out = {}
    for tag, nc in (("nc1", 1), ("nc3", 3)):
        cfg, am, inp = make_case(True, nc, B, N, seed, 0.67)
        cfg_in = _abi.Config.from_buffer_copy(bytes(cfg))
        cfg_in.variant = _abi.VARIANT_CLASSICAL
        inp_in = BatchInputs(14, N, inp.mode, inp.x0[:, :14], inp.p_ref, inp.v_ref, inp.x_reg_ref)
        rng = np.random.default_rng(seed + nc)
        M = 24
        inst = rng.integers(0, B, M).astype(np.int32)
        knot = rng.integers(0, N, M).astype(np.int32)  # running knots (terminal quirk a16 is tested separately)
        x = inp.x0[inst] + rng.normal(0.0, 0.05, (M, 21))
        w = inp.x0[inst][:, 14:] + rng.normal(0.0, 3.0, (M, 7))
        w[::3] *= 4.0  # push some controls over the soft limit so that branch is exercised
        lim = np.array([cfg.tau_limits[i] for i in range(7)])
        ywts = np.array([cfg.y_reg_weights[i] for i in range(21)])
        res = {k: [] for k in ("xnext", "cost", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu")}
        for m in range(M):
            inner = OracleInner(orc, cfg_in, am, inp_in, int(inst[m]), int(knot[m]))
            model = _AugmentedLPFActionModel(inner, 14, 7, cfg.ff_alpha, cfg.w_w, cfg.w_w_soft_limits, lim,
                                             cfg.tau_soft_limit_margin, cfg.w_y, inp.x0[inst[m]], ywts)
            d = types.SimpleNamespace(inner=inner.createData())
            model.calc(d, x[m], w[m])
            model.calcDiff(d, x[m], w[m])
            for k in res:
                res[k].append(np.array(getattr(d, k), dtype=float))
        for k, v in res.items():
            out[f"{tag}_{k}"] = np.stack(v)
        out[f"{tag}_x"], out[f"{tag}_w"], out[f"{tag}_inst"], out[f"{tag}_knot"] = x, w, inst, knot
    out["N"], out["B"], out["seed"], out["contact_fraction"] = N, B, seed, 0.67
    np.savez_compressed(os.path.join(GOLD, "lpf_golden.npz"), **out)
    print("lpf_golden.npz", {k: np.shape(v) for k, v in out.items() if k.startswith("nc1")})


def gen_traj():
    from franka_force_feedback_mpc_b200 import panda
    from franka_force_feedback_mpc_b200.panda import R_MJ_FROM_PIN
    from src.tasks.trajectories import make_approach_then_circle

    model = panda.default_model()
    p_pin, _ = model.fk(panda.Q_NEUTRAL)
    ee_start = R_MJ_FROM_PIN @ p_pin
    base = make_approach_then_circle(center=np.array([-0.5, 0.0, 0.342]), radius=0.10, omega=1.5, z_pre=0.392,
                                     z_contact=0.342, t_approach=0.55, ee_start=ee_start.copy(), t_pre=0.25)
    t_c, t_hold = 0.25 + 0.55, 0.2

    def traj(t):  # the runners' hold wrapper, run_classical.py:249-255 (closure in the runner: restated in 5 lines)
        p, v, s = base(t)
        if s and float(t) < t_c + t_hold:
            ph, _, _ = base(t_c)
            return np.asarray(ph, float), np.zeros(3), True
        return p, v, s

    ts = np.concatenate([np.linspace(0.0, 1.2, 241), np.array([0.25, 0.8, 0.8 - 1e-9, 1.0, 1.0 - 1e-12]), np.linspace(1.2, 20.0, 95)])
    P, V, S = zip(*(traj(float(t)) for t in ts))
    np.savez_compressed(os.path.join(GOLD, "traj_golden.npz"), t=ts, p=np.stack(P), v=np.stack(V), surface=np.array(S),
                        ee_start=ee_start)
    print("traj_golden.npz", len(ts))


def gen_misc():
    from src.mpc.crocoddyl_force_feedback import ForceFeedbackCrocoddylMPC as FF

    rows = []
    for fc in (5.0, 18.0, 25.0, 60.0):
        for dt_ocp, dt in ((0.01, 0.005), (0.02, 0.005), (0.01, 0.01)):
            fake = types.SimpleNamespace(cfg=types.SimpleNamespace(ff_cutoff_hz=fc, dt_ocp=dt_ocp, dt=dt, ff_alpha_override=None),
                                         sim=types.SimpleNamespace(dt=dt), _dt_ocp=dt_ocp)
            rows.append((fc, dt_ocp, dt, FF._ff_alpha_ocp(fake), FF._ff_alpha_ctrl(fake), FF._policy_epsilon(fake)))
    np.savez_compressed(os.path.join(GOLD, "misc_golden.npz"), alpha_rows=np.array(rows))
    print("misc_golden.npz", len(rows))


def gen_ctrl():
    """Host-side per-tick logic of the two controller classes, executed from the reference source on bare instances
    (object.__new__: the constructors need pinocchio / example_robot_data) -> ctrl_golden.npz."""
    from src.mpc.crocoddyl_classical import ClassicalCrocoddylMPC as RC, ClassicalMPCConfig as RCC
    from src.mpc.crocoddyl_force_feedback import ForceFeedbackCrocoddylMPC as RF, ForceFeedbackMPCConfig as RFC

    rng = np.random.default_rng(2024)
    out = {}
    N = 6

    def bare(cls, cfg, dt=0.005):
        o = object.__new__(cls)
        o.cfg = cfg
        o.sim = types.SimpleNamespace(dt=dt)
        o._warned_keys = set()
        o._tau_prev = rng.normal(0, 2.0, 7)
        o.actuation = types.SimpleNamespace(nu=7)
        o.nx_mb, o.ndx_mb, o.nx_aug = 14, 14, 21
        o.xs = o.us = o.Ks = o.ks = None
        o._surface_latched, o._contact_loss_count = False, 0
        return o

    # ---- policy + shift, classical ----
    c = bare(RC, RCC(horizon=N, feedback_gain_scale=0.7))
    out["c_tau_prev"] = c._tau_prev.copy()
    x_now = rng.normal(0, 1, 14)
    out["c_x_now"] = x_now
    out["c_policy_cold"] = c._policy_control(x_now)[0]
    xs0, us0 = c._shift_guess(x_now, N)
    out["c_shift_cold_xs"], out["c_shift_cold_us"] = np.stack(xs0), np.stack(us0)
    for tag, L in (("full", N), ("short", N - 2), ("long", N + 3)):
        c.xs = [rng.normal(0, 1, 14) for _ in range(L + 1)]
        c.us = [rng.normal(0, 5, 7) for _ in range(L)]
        c.Ks = [rng.normal(0, 3, (7, 14)) for _ in range(L)]
        out[f"c_{tag}_xs"], out[f"c_{tag}_us"], out[f"c_{tag}_Ks"] = np.stack(c.xs), np.stack(c.us), np.stack(c.Ks)
        out[f"c_{tag}_policy"] = c._policy_control(x_now)[0]
        xs1, us1 = c._shift_guess(x_now, N)
        out[f"c_{tag}_shift_xs"], out[f"c_{tag}_shift_us"] = np.stack(xs1), np.stack(us1)

    # ---- command filter ----
    for tag, filt in (("raw", False), ("filt", True)):
        c = bare(RC, RCC(apply_command_filter=filt))
        out[f"safe_{tag}_prev0"] = c._tau_prev.copy()
        targets = rng.normal(0, 60.0, (12, 7))
        targets[3, 2] = np.nan
        targets[7] *= 5.0
        out[f"safe_{tag}_in"] = targets
        out[f"safe_{tag}_out"] = np.stack([c._safe_tau(tg) for tg in targets])

    # ---- surface latch ----
    c = bare(RC, RCC(phase_source="force_latch", contact_release_steps=4))
    fn = np.concatenate([np.zeros(5), [0.2, 2.5, 3.0, 0.1, 0.1, 0.1, 6.0, 0.1, 0.1, 0.1, 0.1, 0.1, 0.0], np.zeros(4), [0.0, 0.0]])
    ez = np.concatenate([np.full(5, 0.40), np.full(13, 0.352), np.full(4, 0.45), [0.355, 0.45]])
    hint = np.concatenate([np.zeros(3, bool), np.ones(fn.size - 3, bool)])
    hint[-5:-2] = False
    res = []
    for a, b, h in zip(fn, ez, hint):
        obs = types.SimpleNamespace(f_contact_normal=float(a), ee_pos=np.array([0.0, 0.0, float(b)]))
        res.append(c._detect_surface(obs, 0.0, bool(h)))
    out["latch_fn"], out["latch_ez"], out["latch_hint"], out["latch_out"] = fn, ez, hint, np.array(res)

    # ---- FF policy (Eq. 14-18), tau state source, force-prediction alignment ----
    for tag, interp, dt in (("interp", True, 0.005), ("nointerp", False, 0.005), ("fullstep", True, 0.01)):
        f = bare(RF, RFC(horizon=N, ff_use_tau_interpolation=interp, ff_cutoff_hz=25.0, dt_ocp=0.01), dt=dt)
        f.__class__._dt_ocp.fget  # property exists on the reference class
        y_now = rng.normal(0, 1, 21)
        out[f"f_{tag}_y_now"] = y_now
        out[f"f_{tag}_cold"] = f._policy_control(y_now)[0]
        f.xs = [rng.normal(0, 1, 21) for _ in range(N + 1)]
        f.us = [rng.normal(0, 5, 7) for _ in range(N)]
        f.Ks = [rng.normal(0, 3, (7, 21)) for _ in range(N)]
        out[f"f_{tag}_xs"], out[f"f_{tag}_us"], out[f"f_{tag}_Ks"] = np.stack(f.xs), np.stack(f.us), np.stack(f.Ks)
        out[f"f_{tag}_policy"] = f._policy_control(y_now)[0]
        xs1, us1 = f._shift_guess(y_now, N)
        out[f"f_{tag}_shift_xs"], out[f"f_{tag}_shift_us"] = np.stack(xs1), np.stack(us1)
        f.xs = f.us = None
        xs1, us1 = f._shift_guess(y_now, N)
        out[f"f_{tag}_shift_cold_us"] = np.stack(us1)
    f = bare(RF, RFC(ff_align_min_samples=20, ff_align_window=64, ff_align_max_lag=4))
    f._fn_pred_corr, f._fn_pred_hist_raw, f._fn_pred_hist_meas = np.nan, [], []
    T = 90
    raw = 8.0 + 2.0 * np.sin(np.arange(T) * 0.2) + rng.normal(0, 0.1, T)
    meas = 1.1 * np.roll(raw, 2) + 0.5 + rng.normal(0, 0.1, T)
    meas[10] = np.nan
    surf = np.ones(T, bool)
    surf[:3] = False
    res, corr = [], []
    for a, b, sfc in zip(raw, meas, surf):
        res.append(f._align_logged_force_prediction(float(a), float(b), bool(sfc)))
        corr.append(f._fn_pred_corr)
    out["align_raw"], out["align_meas"], out["align_surf"], out["align_out"], out["align_corr"] = raw, meas, surf, np.array(res), np.array(corr)
    np.savez_compressed(os.path.join(GOLD, "ctrl_golden.npz"), **out)
    print("ctrl_golden.npz", len(out), "arrays")


def gen_scen():
    """The reference's ScenarioUncertaintyInjector (src/run/uncertainty_profiles.py) and RunLogger
    (src/utils/logging.py) executed as they are on seeded inputs -> scen_golden.npz."""
    import json
    import tempfile

    from src.run.uncertainty_profiles import ScenarioUncertaintyInjector, UncertaintyProfileConfig, config_for_scenario
    from src.sim.franka_sim import Observation
    from src.utils.logging import RunLogger

    out = {}
    T = 14
    rng = np.random.default_rng(77)
    q = rng.normal(0, 1, (T, 7)); dq = rng.normal(0, 1, (T, 7)); bias = rng.normal(0, 5, (T, 7)); cmd = rng.normal(0, 10, (T, 7))
    out["in_q"], out["in_dq"], out["in_bias"], out["in_cmd"] = q, dq, bias, cmd
    z7, z3 = np.zeros(7), np.zeros(3)
    cases = [(0.005, config_for_scenario("actuation_uncertainty", seed=15), 0.2),
             (0.001, config_for_scenario("actuation_uncertainty", seed=15), 0.2),
             (0.002, UncertaintyProfileConfig(seed=3, delta_obs_cycles=5, delta_cmd_s=4.0e-3, sigma_tau=0.2, a_min=0.8, a_max=1.2), 0.35),
             (0.005, UncertaintyProfileConfig(seed=99, delta_obs_cycles=0, delta_cmd_s=0.0), 1.0)]
    out["case_dt"] = np.array([c[0] for c in cases]); out["case_alpha"] = np.array([c[2] for c in cases])
    out["case_cfg"] = np.array([[c[1].a_min, c[1].a_max, c[1].b_min, c[1].b_max, c[1].sigma_q, c[1].sigma_dq, c[1].sigma_tau,
                                 c[1].delta_obs_cycles, c[1].delta_cmd_s, c[1].seed] for c in cases])
    for i, (dt, cfg, alpha) in enumerate(cases):
        inj = ScenarioUncertaintyInjector(dt=dt, nu=7, config=cfg, tau_lpf_alpha=alpha)
        res = {k: [] for k in ("q", "dq", "tau_meas", "tau_meas_filt", "tau_meas_act", "tau_meas_act_filt", "tau_bias", "plant")}
        for k in range(T):
            o = Observation(q=q[k], dq=dq[k], tau_meas=z7, tau_meas_filt=z7, tau_meas_act=z7, tau_meas_act_filt=z7, tau_cmd=z7,
                            tau_act=z7, tau_constraint=z7, tau_total=z7, tau_bias=bias[k], f_contact_world=z3, f_contact_normal=0.0,
                            f_contact_normal_world_z=0.0, f_contact_tangent=0.0, contact_count_ee=0, contact_count_table=0,
                            table_normal_world=np.array([0.0, 0.0, 1.0]))
            d = inj.observation_for_controller(o)
            for name in res:
                if name != "plant":
                    res[name].append(np.array(getattr(d, name)))
            res["plant"].append(inj.command_for_plant(cmd[k]))
        for name, v in res.items():
            out[f"inj{i}_{name}"] = np.array(v)
        out[f"inj{i}_meta"] = np.array(json.dumps(inj.meta()))

    # RunLogger: a short run with the field kinds the runners log (run_classical.py:452-481)
    with tempfile.TemporaryDirectory() as tmp:
        lg = RunLogger(run_name="classical_flat", results_dir=tmp, notes={"scenario": "flat", "scene": "assets/scenes/x.xml"})
        for k in range(5):
            lg.log(t=0.005 * (k + 1), ee_pos=q[k, :3].copy(), tau_cmd=cmd[k].copy(), err_tan=float(abs(dq[k, 0])), fn_meas=float(bias[k, 0]),
                   fn_pred=float("nan") if k == 1 else float(bias[k, 1]), contact=int(k % 2), solver_iters=int(3 + k), big=np.arange(12.0) + k)
        lg.set_meta(total_time=0.025, dt=0.005, scenario_label="Flat table", benchmark_mode=True, uncertainty_profile=None,
                    torque_scale=np.ones(7), rms_tangential_error=float(np.sqrt(0.5)), cfg_summary={"horizon": 36, "w": [1.0, 2.0]})
        lg.save()
        out["log_csv"] = np.array(open(lg.path_csv, newline="").read())
        out["log_meta"] = np.array(open(lg.path_meta).read())
        z = np.load(lg.path_npz, allow_pickle=True)
        out["log_npz_keys"] = np.array(sorted(z.files))
        for k in z.files:
            out[f"log_npz_{k}"] = z[k]
    np.savez_compressed(os.path.join(GOLD, "scen_golden.npz"), **out)
    print("scen_golden.npz", len(out), "arrays")


def gen_defaults():
    """Defaults of the reference's two config dataclasses (crocoddyl_classical.py:12-110, crocoddyl_force_feedback.py:12-146),
    field by field, as JSON: config.py and ffmpc_default_config are tested against them (tests/test_config_defaults.py)."""
    import dataclasses
    import json

    from src.mpc.crocoddyl_classical import ClassicalMPCConfig
    from src.mpc.crocoddyl_force_feedback import ForceFeedbackMPCConfig

    def dump(c):
        out = {}
        for f in dataclasses.fields(c):
            v = getattr(c, f.name)
            out[f.name] = v.tolist() if isinstance(v, np.ndarray) else v
        return out

    res = {"ClassicalMPCConfig": dump(ClassicalMPCConfig()), "ForceFeedbackMPCConfig": dump(ForceFeedbackMPCConfig())}
    with open(os.path.join(GOLD, "config_defaults.json"), "w") as f:
        json.dump(res, f, indent=1, sort_keys=True)
    print("config_defaults.json", {k: len(v) for k, v in res.items()})


def main():
    if not os.path.isdir(REF):
        raise SystemExit(f"{REF} not found: the golden fixtures can only be regenerated where the reference is mounted")
    os.makedirs(GOLD, exist_ok=True)
    install_stubs()
    sys.path.insert(0, REF)
    gen_lpf()
    gen_traj()
    gen_misc()
    gen_ctrl()
    gen_scen()
    gen_defaults()


if __name__ == "__main__":
    main()
