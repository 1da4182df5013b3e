"""Host-side mirror of the reference's controller classes (franka_force_feedback_mpc_b200/controllers.py).

* Per-tick logic (`_policy_control`, `_shift_guess`, `_safe_tau`, `_detect_surface`, `_align_logged_force_prediction`)
  is replayed against tests/golden/ctrl_golden.npz, produced by tools/make_golden.py by running the REFERENCE's own
  methods (crocoddyl_classical.py:260-303, 733-779; crocoddyl_force_feedback.py:1014-1093, 1301-1371) on bare instances.
* `compute_control` is driven for a few ticks with the CPU oracle injected as the solver (test infrastructure) to check the
  seam: what is sent to the batched solve (cold start, shifted warm start, references in the Pinocchio world) and what is
  read back.  The GPU test runs the same ticks through the real engine.
"""
from __future__ import annotations

import os
import types

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
G = np.load(os.path.join(HERE, "golden", "ctrl_golden.npz"))
N = 6


class FakeSim:
    """Kinematic stand-in for FrankaMujocoSim: just enough for the controller constructor and compute_control."""

    dt = 0.005

    def __init__(self, model, q):
        from franka_force_feedback_mpc_b200.panda import R_MJ_FROM_PIN
        from franka_force_feedback_mpc_b200.synthetic import batched_gravity_torque

        self.model, self.q, self.dq = model, np.array(q, float), np.zeros(7)
        self.R, self.grav = R_MJ_FROM_PIN, batched_gravity_torque
        self.tau_cmd = self.grav(model, self.q[None])[0]

    def get_observation(self, with_ee=True, with_jacobian=False):
        p, _ = self.model.fk(self.q)
        g = self.grav(self.model, self.q[None])[0]
        return types.SimpleNamespace(q=self.q.copy(), dq=self.dq.copy(), tau_bias=g, ee_pos=self.R @ p, ee_quat=np.array([0.0, 1.0, 0.0, 0.0]),
                                     f_contact_normal=0.0, tau_cmd=self.tau_cmd.copy(), tau_meas_act_filt=self.tau_cmd.copy())


def _make(cls_name, cfg, solver=None, dt=0.005):
    from franka_force_feedback_mpc_b200 import controllers, panda, synthetic

    model = panda.default_model()
    sim = FakeSim(model, panda.Q_NEUTRAL)
    sim.dt = dt
    traj = synthetic.benchmark_trajectory(model)
    ctl = getattr(controllers, cls_name)(sim, lambda t: traj(t), config=cfg, solver=solver)
    return ctl, sim


def _set_plan(ctl, tag, prefix):
    ctl.xs = list(G[f"{prefix}_{tag}_xs"])
    ctl.us = list(G[f"{prefix}_{tag}_us"])
    ctl.Ks = list(G[f"{prefix}_{tag}_Ks"])


def test_classical_policy_and_shift_match_reference_source():
    from franka_force_feedback_mpc_b200.config import ClassicalMPCConfig

    ctl, _ = _make("ClassicalCrocoddylMPC", ClassicalMPCConfig(horizon=N, feedback_gain_scale=0.7))
    ctl._tau_prev = G["c_tau_prev"].copy()
    x_now = G["c_x_now"]
    np.testing.assert_allclose(ctl._policy_control(x_now)[0], G["c_policy_cold"], rtol=0, atol=0)
    xs, us = ctl._shift_guess(x_now, N)
    np.testing.assert_array_equal(np.stack(xs), G["c_shift_cold_xs"])
    np.testing.assert_array_equal(np.stack(us), G["c_shift_cold_us"])
    for tag in ("full", "short", "long"):
        _set_plan(ctl, tag, "c")
        u, idx = ctl._policy_control(x_now)
        assert idx == 0
        np.testing.assert_allclose(u, G[f"c_{tag}_policy"], rtol=1e-14, atol=1e-14)   # PLUS-sign feedback of the reference
        xs, us = ctl._shift_guess(x_now, N)
        np.testing.assert_array_equal(np.stack(xs), G[f"c_{tag}_shift_xs"])
        np.testing.assert_array_equal(np.stack(us), G[f"c_{tag}_shift_us"])


@pytest.mark.parametrize("tag,filt", [("raw", False), ("filt", True)])
def test_safe_tau_matches_reference_source(tag, filt):
    from franka_force_feedback_mpc_b200.config import ClassicalMPCConfig

    ctl, _ = _make("ClassicalCrocoddylMPC", ClassicalMPCConfig(apply_command_filter=filt))
    ctl._tau_prev = G[f"safe_{tag}_prev0"].copy()
    got = np.stack([ctl._safe_tau(t) for t in G[f"safe_{tag}_in"]])
    np.testing.assert_allclose(got, G[f"safe_{tag}_out"], rtol=1e-15, atol=1e-15)


def test_surface_latch_matches_reference_source():
    from franka_force_feedback_mpc_b200.config import ClassicalMPCConfig

    ctl, _ = _make("ClassicalCrocoddylMPC", ClassicalMPCConfig(phase_source="force_latch", contact_release_steps=4))
    got = []
    for fn, ez, hint in zip(G["latch_fn"], G["latch_ez"], G["latch_hint"]):
        obs = types.SimpleNamespace(f_contact_normal=float(fn), ee_pos=np.array([0.0, 0.0, float(ez)]))
        got.append(ctl._detect_surface(obs, 0.0, bool(hint)))
    assert got == list(G["latch_out"])
    assert any(got) and not all(got)


@pytest.mark.parametrize("tag,interp,dt", [("interp", True, 0.005), ("nointerp", False, 0.005), ("fullstep", True, 0.01)])
def test_ff_policy_and_shift_match_reference_source(tag, interp, dt):
    from franka_force_feedback_mpc_b200.config import ForceFeedbackMPCConfig

    cfg = ForceFeedbackMPCConfig(horizon=N, ff_use_tau_interpolation=interp, ff_cutoff_hz=25.0, dt_ocp=0.01)
    ctl, _ = _make("ForceFeedbackCrocoddylMPC", cfg, dt=dt)
    y_now = G[f"f_{tag}_y_now"]
    np.testing.assert_array_equal(ctl._policy_control(y_now)[0], G[f"f_{tag}_cold"])
    _set_plan(ctl, tag, "f")
    u, idx = ctl._policy_control(y_now)
    assert idx == 0
    np.testing.assert_allclose(u, G[f"f_{tag}_policy"], rtol=1e-13, atol=1e-13)   # Eq. 14-18
    xs, us = ctl._shift_guess(y_now, N)
    np.testing.assert_array_equal(np.stack(xs), G[f"f_{tag}_shift_xs"])
    np.testing.assert_array_equal(np.stack(us), G[f"f_{tag}_shift_us"])
    ctl.xs = ctl.us = None
    _, us = ctl._shift_guess(y_now, N)
    np.testing.assert_array_equal(np.stack(us), G[f"f_{tag}_shift_cold_us"])   # cold start holds tau_hat


def test_force_prediction_alignment_matches_reference_source():
    from franka_force_feedback_mpc_b200.config import ForceFeedbackMPCConfig

    ctl, _ = _make("ForceFeedbackCrocoddylMPC", ForceFeedbackMPCConfig(ff_align_min_samples=20, ff_align_window=64, ff_align_max_lag=4))
    got, corr = [], []
    for a, b, s in zip(G["align_raw"], G["align_meas"], G["align_surf"]):
        got.append(ctl._align_logged_force_prediction(float(a), float(b), bool(s)))
        corr.append(ctl._fn_pred_corr)
    np.testing.assert_allclose(np.array(got), G["align_out"], rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(np.array(corr), G["align_corr"], rtol=1e-10, atol=1e-10, equal_nan=True)


# ---- compute_control through the seam -----------------------------------------------------------------
class OracleSolver:
    """Test stand-in for the engine: same call, answered by the CPU oracle; records what the controller sent."""

    def __init__(self, ctl_getter):
        self.calls, self._ctl = [], ctl_getter

    def solve(self, inp, max_iters):
        from oracle import oracle as orc

        ctl = self._ctl()
        cfg_abi, model_abi = ctl._abi()
        self.calls.append(inp)
        return orc.solve_batch(cfg_abi, model_abi, inp, max_iters, False, nthreads=1)


REF_INFO_KEYS_CLASSICAL = {"ok", "cost", "iters", "tau_raw_inf", "tau_cmd_inf", "surface_mode", "unstable", "fn_pred", "solved_now", "policy_idx"}
REF_INFO_KEYS_FF = REF_INFO_KEYS_CLASSICAL | {"tau_des_inf", "tau_meas_state_inf", "fn_pred_raw", "fn_pred_corr"}


def _run_ticks(cls_name, cfg, solver_factory, ticks=3, t0=1.5):
    holder = {}
    solver = solver_factory(lambda: holder["ctl"])
    ctl, sim = _make(cls_name, cfg, solver=solver)
    holder["ctl"] = ctl
    taus, infos = [], []
    for k in range(ticks):
        obs = sim.get_observation()
        obs.f_contact_normal = 7.5
        tau = ctl.compute_control(obs, t0 + k * sim.dt)
        taus.append(tau)
        infos.append(dict(ctl.last_info))
        sim.q = sim.q + 1e-3 * np.sin(np.arange(7) + k)   # small kinematic drift so that x0 changes between ticks
        sim.tau_cmd = tau.copy()
    return ctl, solver, np.stack(taus), infos


@pytest.mark.parametrize("cls_name,ff", [("ClassicalCrocoddylMPC", False), ("ForceFeedbackCrocoddylMPC", True)])
def test_compute_control_seam_with_oracle_solver(orc, cls_name, ff):
    from franka_force_feedback_mpc_b200 import config

    cfg = (config.benchmark_ff_config if ff else config.benchmark_classical_config)(horizon=8, max_iters=4)
    ctl, solver, taus, infos = _run_ticks(cls_name, cfg, OracleSolver)
    nx = 21 if ff else 14
    assert np.isfinite(taus).all() and taus.shape == (3, 7)
    assert set(infos[-1]) == (REF_INFO_KEYS_FF if ff else REF_INFO_KEYS_CLASSICAL)
    assert all(i["solved_now"] and i["surface_mode"] and not i["unstable"] for i in infos)   # t = 1.5 s: contact phase
    assert np.isfinite(infos[-1]["fn_pred"]) and infos[-1]["iters"] >= 1
    first, second = solver.calls[0], solver.calls[1]
    assert first.B == 1 and first.nx == nx and first.mode[0] == 1
    # tick 1: cold start = x0 repeated, controls = gravity torque / tau_hat
    np.testing.assert_array_equal(first.xs_init[0], np.repeat(first.x0, 9, axis=0))
    # tick 2: shifted warm start (crocoddyl_classical.py:733-757): [x0] + xs_prev[1:], us_prev[1:] + [us_prev[-1]]
    np.testing.assert_array_equal(second.xs_init[0, 0], second.x0[0])
    assert not np.array_equal(second.xs_init[0, 1], second.x0[0])
    np.testing.assert_array_equal(second.us_init[0, -1], second.us_init[0, -2])
    # references are the trajectory mapped MuJoCo -> Pinocchio world: p_pin = R^T p_mj - offset
    p_mj, v_mj, surf = ctl.traj_fn(1.5 + 3 * 0.01)
    np.testing.assert_allclose(first.p_ref[0, 3], ctl._pos_mj_to_pin(p_mj), atol=1e-15)
    np.testing.assert_allclose(first.v_ref[0, 3], ctl._vel_mj_to_pin(v_mj), atol=1e-15)
    assert abs(np.linalg.det(ctl.R_des) - 1.0) < 1e-9


def test_controller_without_gpu_fails_loudly():
    """No solver injected -> the product path constructs the CUDA engine, which must refuse to run without a device."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from franka_force_feedback_mpc_b200 import config, engine

    ctl, sim = _make("ClassicalCrocoddylMPC", config.benchmark_classical_config(horizon=4, max_iters=1))
    with pytest.raises((engine.FFMPCError, OSError)):
        ctl.compute_control(sim.get_observation(), 1.5)


@pytest.mark.gpu
@pytest.mark.parametrize("cls_name,ff", [("ClassicalCrocoddylMPC", False), ("ForceFeedbackCrocoddylMPC", True)])
def test_compute_control_on_gpu_matches_oracle_solver(built, orc, cls_name, ff):
    from franka_force_feedback_mpc_b200 import config

    cfg = (config.benchmark_ff_config if ff else config.benchmark_classical_config)(horizon=20, max_iters=6)
    _, _, tau_ref, info_ref = _run_ticks(cls_name, cfg, OracleSolver, ticks=4)
    _, _, tau_gpu, info_gpu = _run_ticks(cls_name, cfg, lambda getter: None, ticks=4)
    np.testing.assert_allclose(tau_gpu, tau_ref, rtol=1e-6, atol=1e-6)
    for a, b in zip(info_gpu, info_ref):
        assert a["iters"] == b["iters"] and a["ok"] == b["ok"] and a["surface_mode"] == b["surface_mode"]
        assert abs(a["cost"] - b["cost"]) <= 1e-8 * max(1.0, abs(b["cost"]))
        assert abs(a["fn_pred"] - b["fn_pred"]) <= 1e-6 * max(1.0, abs(b["fn_pred"]))
