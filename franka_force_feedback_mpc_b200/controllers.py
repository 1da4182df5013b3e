"""Host-side mirror of the reference's controller interface (SURVEY.md 8b "Level-1 API", 8f1):

    ClassicalCrocoddylMPC(sim, traj_fn, config=..., lock_fingers=True, pin_ee_frame="panda_link8")
    ForceFeedbackCrocoddylMPC(...same...)
    .compute_control(obs, t) -> np.ndarray(7);  .last_info;  .xs / .us / .Ks

so that the reference's runners (`src/run/run_*.py`) and its CPU MuJoCo loop can drive the B200 engine with B = 1.
Everything Crocoddyl/Pinocchio did at the seam `crocoddyl_classical.py:344-388` / `crocoddyl_force_feedback.py:584-628`
(problem build, solve, read-back) is ONE call into the C ABI (`ffmpc_solve_batch_host`); what stays on the host is the
O(7) per-tick logic of the reference, restated here:

    surface latch            crocoddyl_classical.py:286-303
    warm-start shift         crocoddyl_classical.py:733-757   (FF :1014-1039)
    feedback policy          crocoddyl_classical.py:759-779   (FF Eq. 14-18, :1041-1093)
    unstable guard           crocoddyl_classical.py:393-403
    command filter           crocoddyl_classical.py:260-284
    force-prediction logging crocoddyl_force_feedback.py:1219-1243, 1301-1371

The rigid-body model is this package's `PandaModel` (the reference loads it from example_robot_data, which is not
installable here).  `solver=` lets tests inject a stand-in for the engine (the CPU oracle or the kernel emulator); the
product default constructs `FFMPCEngine`, which raises without a GPU -- there is no CPU fallback.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import numpy as np

from . import config as _config
from . import panda as _panda
from .batch import BatchInputs
from .config import ClassicalMPCConfig, ForceFeedbackMPCConfig

_NEVER = -1_000_000_000
MODE_FREE, MODE_CONTACT = 0, 1


def quat_wxyz_to_R(q) -> np.ndarray:
    w, x, y, z = np.asarray(q, dtype=float).reshape(4) / (np.linalg.norm(q) + 1e-12)
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def vertical_down_rotation_mj() -> np.ndarray:
    """EE z axis pointing down, x along world x (MuJoCo world), crocoddyl_classical.py:240-248."""
    z = np.array([0.0, 0.0, -1.0])
    y = np.cross(z, np.array([1.0, 0.0, 0.0]))
    y /= np.linalg.norm(y) + 1e-12
    x = np.cross(y, z)
    x /= np.linalg.norm(x) + 1e-12
    return np.column_stack([x, y, z])


class _EngineSolver:
    """Default solver: the CUDA engine through the C ABI, B = 1."""

    def __init__(self, cfg_abi, model_abi, device: int = 0, _lib=None):
        from .engine import FFMPCEngine

        self.eng = FFMPCEngine(cfg_abi, model_abi, device=device, max_batch=1, _lib=_lib)

    def solve(self, inp: BatchInputs, max_iters: int) -> dict:
        # fixed_iters = False: keep Crocoddyl's converged early exit (solver.solve(xs, us, maxiter, False))
        return self.eng.solve_host(inp, max_iters, fixed_iters=False, want=("xs", "us", "K", "cost", "iters", "ok", "f_pred"))


class _PandaMPC:
    """Logic shared by the two controllers; subclasses define the state layout and the policy."""

    variant_ff = False

    def __init__(self, sim, traj_fn: Callable[[float], Tuple[np.ndarray, np.ndarray, bool]], config=None, lock_fingers: bool = True,
                 pin_ee_frame: str = "panda_link8", model: Optional[_panda.PandaModel] = None, solver=None, device: int = 0):
        if pin_ee_frame != "panda_link8":
            raise ValueError(f"Pinocchio frame '{pin_ee_frame}' does not exist in panda model.")
        self.sim, self.traj_fn = sim, traj_fn
        self.cfg = config if config is not None else (ForceFeedbackMPCConfig() if self.variant_ff else ClassicalMPCConfig())
        self.model = model if model is not None else _panda.default_model()
        self.data = None          # the reference exposes a pinocchio Data here; the engine keeps its own workspace
        self.ee_fid = 0           # single operational frame of PandaModel
        self._k = 0
        self._warned_keys = set()
        self.R_mj_from_pin = np.diag([-1.0, -1.0, 1.0])
        obs0 = sim.get_observation(with_ee=True, with_jacobian=False)
        self.q_nom = np.asarray(obs0.q, dtype=float).copy()
        p_pin, R_pin = self.model.fk(self.q_nom)
        if getattr(obs0, "ee_quat", None) is not None:
            R_site_mj = quat_wxyz_to_R(obs0.ee_quat)
        else:
            R_site_mj = np.asarray(sim.data.site_xmat[sim.ee_site_id], dtype=float).reshape(3, 3).copy()
        self.R_site_from_pin_ee = R_pin.T @ self.R_mj_from_pin.T @ R_site_mj
        if getattr(obs0, "ee_pos", None) is not None:
            p_site_mj = np.asarray(obs0.ee_pos, dtype=float).reshape(3)
        else:
            p_site_mj = np.asarray(sim.data.site_xpos[sim.ee_site_id], dtype=float).reshape(3)
        self.p_site_minus_frame_pin = self.R_mj_from_pin.T @ p_site_mj - p_pin
        self.R_des = self._rot_mj_to_pin(vertical_down_rotation_mj())
        self.xs: Optional[List[np.ndarray]] = None
        self.us: Optional[List[np.ndarray]] = None
        self.Ks: Optional[List[np.ndarray]] = None
        self.ks = None
        self._tau_prev = np.asarray(obs0.tau_bias, dtype=float).copy()
        self._last_solve_step = _NEVER
        self._last_solve_ok, self._last_solve_cost, self._last_solve_iters = False, np.nan, -1
        self._surface_latched, self._contact_loss_count = False, 0
        self._prev_surface_mode: Optional[bool] = None
        self.last_info = {"ok": False, "cost": np.nan, "iters": -1, "tau_raw_inf": np.nan, "tau_cmd_inf": np.nan,
                          "surface_mode": False, "unstable": False, "fn_pred": np.nan}
        self._solver = solver
        self._device = device
        self._abi_cache = None

    # ---- frames ---------------------------------------------------------------------------------
    @property
    def _dt_ocp(self) -> float:
        return float(max(_config.dt_ocp_of(self.cfg), 1.0e-6))

    def _pos_mj_to_pin(self, p_mj) -> np.ndarray:
        return self.R_mj_from_pin.T @ np.asarray(p_mj, dtype=float) - self.p_site_minus_frame_pin

    def _vel_mj_to_pin(self, v_mj) -> np.ndarray:
        return self.R_mj_from_pin.T @ np.asarray(v_mj, dtype=float)

    def _rot_mj_to_pin(self, R_mj_site) -> np.ndarray:
        return self.R_mj_from_pin.T @ np.asarray(R_mj_site, dtype=float) @ self.R_site_from_pin_ee.T

    # ---- per-tick host logic ------------------------------------------------------------------------
    def _safe_tau(self, tau_target) -> np.ndarray:
        c = self.cfg
        tau = np.array(tau_target, dtype=float)
        if not np.isfinite(tau).all():
            tau = self._tau_prev.copy()
        tau = np.clip(tau, -c.tau_limits, c.tau_limits)
        if bool(c.apply_command_filter):
            step = np.clip(tau - self._tau_prev, -c.tau_trust_inf, c.tau_trust_inf)
            slew = np.asarray(c.tau_rate_limit, dtype=float) * float(getattr(self.sim, "dt", c.dt))
            limited = self._tau_prev + np.clip(step, -slew, slew)
            a = float(np.clip(c.tau_smoothing_alpha, 0.0, 1.0))
            tau = np.clip((1.0 - a) * self._tau_prev + a * limited, -c.tau_limits, c.tau_limits)
        self._tau_prev = tau.copy()
        return tau

    def _detect_surface(self, obs, t: float, surf_hint: bool) -> bool:
        c = self.cfg
        fn = float(getattr(obs, "f_contact_normal", 0.0))
        ee_pos = getattr(obs, "ee_pos", None)
        ee_z = float(ee_pos[2]) if ee_pos is not None else float("inf")
        near = np.isfinite(ee_z) and ee_z <= float(c.z_contact) + float(c.z_contact_band)
        if self._surface_latched:
            self._contact_loss_count = self._contact_loss_count + 1 if fn < c.fn_contact_off else 0
            if self._contact_loss_count >= int(c.contact_release_steps):
                self._surface_latched, self._contact_loss_count = False, 0
        elif fn > c.fn_contact_on or (surf_hint and near):
            self._surface_latched, self._contact_loss_count = True, 0
        return self._surface_latched

    def _cold_control(self, x0: np.ndarray) -> np.ndarray:
        raise NotImplementedError

    def _shift_guess(self, x0: np.ndarray, N: int):
        """Receding-horizon warm start: previous solution shifted by one knot, x0 in front; cold start otherwise."""
        if self.xs is None or self.us is None or len(self.us) < N:
            u = self._cold_control(x0)
            return [x0.copy() for _ in range(N + 1)], [u.copy() for _ in range(N)]
        xs = [x0.copy()] + [a.copy() for a in self.xs[1:N + 1]]
        xs += [self.xs[-1].copy() for _ in range(N + 1 - len(xs))]
        us = [a.copy() for a in self.us[1:N]]
        us += [self.us[-1].copy() for _ in range(N - len(us))]
        return xs, us

    def _drop_plan(self):
        self.xs = self.us = self.Ks = self.ks = None
        self._last_solve_step = _NEVER

    def _surface_mode(self, obs, t: float) -> bool:
        _, _, hint = self.traj_fn(t)
        if str(self.cfg.phase_source).strip().lower() == "force_latch":
            now = bool(self._detect_surface(obs, t, hint))
        else:
            now = bool(hint)
        if self._prev_surface_mode is None:
            self._prev_surface_mode = now
        elif now != self._prev_surface_mode:
            self._drop_plan()      # a mode switch invalidates the shifted warm start
            self._prev_surface_mode = now
        return now

    # ---- the seam: one batched-solve call (B = 1) ---------------------------------------------------
    def _references(self, t0: float):
        N, dt = int(self.cfg.horizon), self._dt_ocp
        p = np.empty((1, N + 1, 3))
        v = np.empty((1, N + 1, 3))
        for k in range(N + 1):
            p_mj, v_mj, _ = self.traj_fn(t0 + k * dt)
            p[0, k] = self._pos_mj_to_pin(p_mj)
            v[0, k] = self._vel_mj_to_pin(v_mj)
        return p, v

    def _posture_reference(self, x0_mb: np.ndarray) -> np.ndarray:
        if str(self.cfg.posture_ref_mode).strip().lower() == "q_nom":
            return np.concatenate([self.q_nom, np.zeros(7)])
        return np.asarray(x0_mb, dtype=float).copy()

    def _tau_reference(self, q_now: np.ndarray) -> Optional[np.ndarray]:
        mode = str(self.cfg.torque_ref_mode).strip().lower()
        if mode == "zero":
            return np.zeros((1, 7))
        if mode == "gravity_qnom":
            from .synthetic import batched_gravity_torque

            return batched_gravity_torque(self.model, self.q_nom[None])
        return None  # "gravity_x0": RNEA(q0, 0, 0), evaluated inside the engine

    def _abi(self):
        if self._abi_cache is None:
            self._abi_cache = (_config.to_abi(self.cfg, self.model, self.q_nom, self.R_des), self.model.to_abi())
        return self._abi_cache

    def _solve(self, t: float, x0: np.ndarray, surface_now: bool) -> dict:
        N, nx = int(self.cfg.horizon), x0.size
        p_ref, v_ref = self._references(t)
        xs_init, us_init = self._shift_guess(x0, N)
        inp = BatchInputs(nx, N, [MODE_CONTACT if surface_now else MODE_FREE], x0[None], p_ref, v_ref,
                          self._posture_reference(x0[:14])[None], tau_ref=self._tau_reference(x0[:7]),
                          xs_init=np.stack(xs_init)[None], us_init=np.stack(us_init)[None])
        if self._solver is None:
            cfg_abi, model_abi = self._abi()
            self._solver = _EngineSolver(cfg_abi, model_abi, self._device)
        out = self._solver.solve(inp, int(self.cfg.max_iters))
        self._last_solve_step = self._k
        self._last_solve_ok = bool(out["ok"][0])
        self._last_solve_cost = float(out["cost"][0])
        self._last_solve_iters = int(out["iters"][0])
        if N > 0 and np.isfinite(out["us"][0, 0]).all():
            self.xs = [out["xs"][0, k].copy() for k in range(N + 1)]
            self.us = [out["us"][0, k].copy() for k in range(N)]
            self.Ks = [out["K"][0, k].copy() for k in range(N)]
            self.ks = None
        return out

    def _shift_plan_between_solves(self):
        if self.us is None or self.xs is None:
            return
        for name in ("us", "xs", "Ks"):
            seq = getattr(self, name)
            if seq is not None and len(seq) > 1:
                setattr(self, name, seq[1:] + [seq[-1]])
        self.ks = None

    def _need_solve(self) -> bool:
        period = max(1, int(self.cfg.mpc_update_steps))
        return self.us is None or self.xs is None or (self._k - self._last_solve_step) >= period

    def _is_unstable(self, cost: float, tau_raw_inf: float) -> bool:
        return (not np.isfinite(cost)) or cost > float(self.cfg.max_solver_cost) or tau_raw_inf > float(self.cfg.max_tau_raw_inf)


class ClassicalCrocoddylMPC(_PandaMPC):
    """State (q, v), control tau (nx = 14, nu = 7).  crocoddyl_classical.py:113-440."""

    variant_ff = False

    def _cold_control(self, x0):
        return self._tau_prev

    def _policy_control(self, x_now: np.ndarray):
        """us[0] + s K0 (x_now - xs[0]) -- the PLUS sign is the reference's (crocoddyl_classical.py:775-777)."""
        if self.us is None or len(self.us) == 0:
            return self._tau_prev.copy(), -1
        u = np.array(self.us[0], dtype=float)
        if self.cfg.use_feedback_policy and self.Ks is not None and self.xs is not None and len(self.Ks) > 0 and len(self.xs) > 0:
            u = u + float(self.cfg.feedback_gain_scale) * (np.asarray(self.Ks[0], dtype=float) @ (x_now - self.xs[0]))
        return u, 0

    def compute_control(self, obs, t: float) -> np.ndarray:
        self._k += 1
        v = np.asarray(obs.dq, dtype=float)
        x0 = np.concatenate([np.asarray(obs.q, dtype=float), v])
        surface_now = self._surface_mode(obs, t)
        ok, cost, iters = self._last_solve_ok, float(self._last_solve_cost), int(self._last_solve_iters)
        fn_pred = float(self.last_info.get("fn_pred", np.nan))
        solved_now = False
        if self._need_solve():
            out = self._solve(t, x0, surface_now)
            ok, cost, iters = self._last_solve_ok, self._last_solve_cost, self._last_solve_iters
            fn_pred = float(out["f_pred"][0, 0, 2]) if surface_now else np.nan   # contact force of knot 0, world z (:905-942)
            solved_now = True
        tau_raw, policy_idx = self._policy_control(x0)
        tau_raw_inf = float(np.max(np.abs(tau_raw)))
        unstable = self._is_unstable(cost, tau_raw_inf)
        if unstable:
            tau_raw = np.asarray(obs.tau_bias, dtype=float) - float(self.cfg.fallback_dq_damping) * v
            self._drop_plan()
        tau_cmd = self._safe_tau(tau_raw)
        self.last_info = {"ok": bool(ok), "cost": float(cost), "iters": iters, "tau_raw_inf": tau_raw_inf,
                          "tau_cmd_inf": float(np.max(np.abs(tau_cmd))), "surface_mode": bool(surface_now), "unstable": bool(unstable),
                          "fn_pred": float(fn_pred) if np.isfinite(fn_pred) else np.nan, "solved_now": bool(solved_now),
                          "policy_idx": int(policy_idx)}
        if not solved_now:
            self._shift_plan_between_solves()
        return tau_cmd


class ForceFeedbackCrocoddylMPC(_PandaMPC):
    """Augmented state (q, v, tau_hat), control w (nx = 21, nu = 7).  crocoddyl_force_feedback.py:293-695."""

    variant_ff = True
    nx_mb, nx_aug = 14, 21

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self._fn_pred_corr = np.nan
        self._fn_pred_hist_raw: List[float] = []
        self._fn_pred_hist_meas: List[float] = []

    def _cold_control(self, y0):
        return np.asarray(y0[14:21], dtype=float)

    def _ff_alpha_ocp(self) -> float:
        return _config.ff_alpha_ocp(self.cfg)

    def _ff_alpha_ctrl(self) -> float:
        if self.cfg.ff_alpha_override is not None:
            return float(np.clip(float(self.cfg.ff_alpha_override), 0.0, 0.999999))
        wc = 2.0 * np.pi * float(max(self.cfg.ff_cutoff_hz, 0.0))
        return float(np.clip(np.exp(-wc * float(getattr(self.sim, "dt", self.cfg.dt))), 0.0, 0.999999))

    def _policy_epsilon(self) -> float:
        return float(np.clip(float(getattr(self.sim, "dt", self.cfg.dt)) / self._dt_ocp, 0.0, 1.0))

    _TAU_SOURCES = {
        "tau_meas_act_filt": ("tau_meas_act_filt", "tau_meas_act", "tau_cmd"),
        "tau_meas_act": ("tau_meas_act", "tau_cmd"),
        "tau_cmd": ("tau_cmd",),
        "tau_meas_filt": ("tau_meas_filt", "tau_meas"),
        "tau_meas": ("tau_meas",),
        "tau_total": ("tau_total", "tau_meas"),
    }

    def _tau_state_from_obs(self, obs) -> np.ndarray:
        src = str(self.cfg.ff_tau_state_source).strip().lower()
        if src == "auto":
            src = "tau_meas_filt" if bool(self.cfg.ff_use_tau_meas_filt) else "tau_meas"
        for key in self._TAU_SOURCES.get(src, ("tau_meas_act_filt", "tau_meas_act", "tau_cmd", "tau_meas")) + ("tau_cmd",):
            val = getattr(obs, key, None)
            if val is None:
                continue
            tau = np.asarray(val, dtype=float).reshape(7)
            if np.isfinite(tau).all():
                return tau
        return np.zeros(7)

    def _policy_control(self, y_now: np.ndarray):
        """Interpolated filtered-torque policy with Riccati feedback (Eq. 14-18 of the reference, :1041-1093)."""
        if self.us is None or self.xs is None or len(self.us) == 0 or len(self.xs) == 0:
            return np.asarray(y_now[14:21], dtype=float).copy(), -1
        alpha = self._ff_alpha_ocp()
        eps = self._policy_epsilon() if bool(self.cfg.ff_use_tau_interpolation) else 0.0
        y_nom = np.asarray(self.xs[0], dtype=float)
        tau0 = y_nom[14:21]
        tau1 = np.asarray(self.xs[1], dtype=float)[14:21] if len(self.xs) > 1 else alpha * tau0 + (1.0 - alpha) * np.asarray(self.us[0], float)
        tau = tau0 + eps * (tau1 - tau0)
        if self.cfg.use_feedback_policy and self.Ks is not None and len(self.Ks) > 0:
            K0 = np.asarray(self.Ks[0], dtype=float)
            if K0.ndim == 2 and K0.shape[1] >= 21:
                gain = eps * (1.0 - alpha)
                x_err = y_nom[:14] - y_now[:14]
                tau_err = tau0 - y_now[14:21]
                fb = gain * (K0[:, :14] @ x_err) + (np.eye(7) + gain * (K0[:, 14:21] - np.eye(7))) @ tau_err
                tau = tau + float(self.cfg.feedback_gain_scale) * fb
        return np.asarray(tau, dtype=float).reshape(7), 0

    def _force_next_step(self, f_pred: np.ndarray) -> float:
        """|f_z| of knots 0 / 1 blended to t + dt_mpc (:1219-1243); f_pred = [2, 3] contact forces from the engine."""
        f0, f1 = abs(float(f_pred[0, 2])), abs(float(f_pred[1, 2]))
        if int(self.cfg.horizon) == 1 or not np.isfinite(f1):
            return f0
        if not np.isfinite(f0):
            return f1
        if float(getattr(self.sim, "dt", self.cfg.dt)) >= self._dt_ocp - 1.0e-9:
            return f0
        eps = self._policy_epsilon()
        return float((1.0 - eps) * f0 + eps * f1)

    def _align_logged_force_prediction(self, fn_pred_raw: float, fn_meas: float, surface_now: bool) -> float:
        """Logging only: affine fit + lag of the predicted force against the measured one over a sliding window (:1301-1371)."""
        c = self.cfg
        if not np.isfinite(fn_pred_raw):
            return np.nan
        if not surface_now or not bool(c.ff_align_force_prediction):
            self._fn_pred_corr = np.nan
            return float(fn_pred_raw)
        if np.isfinite(fn_meas):
            win = int(max(c.ff_align_window, 16))
            self._fn_pred_hist_raw = (self._fn_pred_hist_raw + [float(fn_pred_raw)])[-win:]
            self._fn_pred_hist_meas = (self._fn_pred_hist_meas + [float(fn_meas)])[-win:]
        min_n = int(max(c.ff_align_min_samples, 8))
        raw, meas = np.asarray(self._fn_pred_hist_raw, float), np.asarray(self._fn_pred_hist_meas, float)
        n = min(raw.size, meas.size)
        if n < min_n:
            self._fn_pred_corr = np.nan
            return float(fn_pred_raw)
        corr_min = float(max(c.ff_align_corr_threshold, 0.0))
        best = None
        for lag in range(min(int(max(c.ff_align_max_lag, 0)), n - min_n) + 1):
            x, y = (raw[:-lag], meas[lag:]) if lag > 0 else (raw, meas)
            if min(x.size, y.size) < min_n:
                continue
            xc, yc = x - x.mean(), y - y.mean()
            den = float(np.linalg.norm(xc) * np.linalg.norm(yc))
            if den < 1.0e-9:
                continue
            corr = float(xc @ yc / den)
            if abs(corr) < corr_min:
                continue
            a, b = np.linalg.lstsq(np.column_stack([x, np.ones_like(x)]), y, rcond=None)[0]
            rmse = float(np.sqrt(np.mean((a * x + b - y) ** 2)))
            if best is None or rmse < best[0]:
                best = (rmse, lag, corr, float(a), float(b))
        if best is None:
            self._fn_pred_corr = np.nan
            return float(fn_pred_raw)
        _, lag, corr, a, b = best
        self._fn_pred_corr = corr
        return float(max(a * float(raw[max(0, raw.size - 1 - lag)]) + b, 0.0))

    def compute_control(self, obs, t: float) -> np.ndarray:
        self._k += 1
        v = np.asarray(obs.dq, dtype=float)
        tau_hat = self._tau_state_from_obs(obs)
        y0 = np.concatenate([np.asarray(obs.q, dtype=float), v, tau_hat])
        surface_now = self._surface_mode(obs, t)
        ok, cost, iters = self._last_solve_ok, float(self._last_solve_cost), int(self._last_solve_iters)
        fn_pred_raw = float(self.last_info.get("fn_pred_raw", self.last_info.get("fn_pred", np.nan)))
        solved_now = False
        if self._need_solve():
            out = self._solve(t, y0, surface_now)
            ok, cost, iters = self._last_solve_ok, self._last_solve_cost, self._last_solve_iters
            fn_pred_raw = self._force_next_step(out["f_pred"][0]) if surface_now else np.nan
            solved_now = True
        tau_des, policy_idx = self._policy_control(y0)
        tau_raw = np.array(tau_des, dtype=float)
        if bool(self.cfg.ff_inverse_actuation_model):
            a_ctrl = self._ff_alpha_ctrl()
            tau_raw = (tau_raw - a_ctrl * tau_hat) / max(1.0e-6, 1.0 - a_ctrl)
        tau_raw_inf = float(np.max(np.abs(tau_raw)))
        unstable = self._is_unstable(cost, tau_raw_inf)
        if unstable:
            tau_raw = np.asarray(obs.tau_bias, dtype=float) - float(self.cfg.fallback_dq_damping) * v
            self._drop_plan()
        tau_cmd = self._safe_tau(tau_raw)
        fn_pred = self._align_logged_force_prediction(fn_pred_raw, float(getattr(obs, "f_contact_normal", np.nan)), surface_now)
        nan_if = lambda x: float(x) if np.isfinite(x) else np.nan
        self.last_info = {"ok": bool(ok), "cost": float(cost), "iters": iters, "tau_des_inf": float(np.max(np.abs(tau_des))),
                          "tau_meas_state_inf": float(np.max(np.abs(tau_hat))), "tau_raw_inf": tau_raw_inf,
                          "tau_cmd_inf": float(np.max(np.abs(tau_cmd))), "surface_mode": bool(surface_now), "unstable": bool(unstable),
                          "fn_pred": nan_if(fn_pred), "fn_pred_raw": nan_if(fn_pred_raw), "fn_pred_corr": nan_if(self._fn_pred_corr),
                          "solved_now": bool(solved_now), "policy_idx": int(policy_idx)}
        if not solved_now:
            self._shift_plan_between_solves()
        return tau_cmd
