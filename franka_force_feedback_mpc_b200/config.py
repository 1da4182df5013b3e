"""Controller configuration: host-side mirror of the reference's two config dataclasses.

``ClassicalMPCConfig``      <- src/mpc/crocoddyl_classical.py:12-110
``ForceFeedbackMPCConfig``  <- src/mpc/crocoddyl_force_feedback.py:12-146

Field names, defaults and meaning are the reference's, so ``src/run/run_*.py`` can construct them
unchanged (run_classical.py:269-361, run_force_feedback.py:272-387).  :func:`to_abi` lowers a
config to the plain C struct ``ffmpc_config`` of ``include/ffmpc.h``; everything that is a pure
function of static config (soft joint-limit bounds, LPF alpha, R_des) is resolved here once, not
per tick.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

from . import _abi
from .panda import PandaModel, TAU_LIMITS, reference_R_des


def _arr(*v):
    return field(default_factory=lambda: np.array(v, dtype=float))


@dataclass
class _CommonMPCConfig:
    # timing
    horizon: int = 20
    dt: float = 0.01
    dt_ocp: Optional[float] = None
    # free-space tracking
    w_ee_pos: float = 2.0e2
    w_ee_ori: float = 1.0e1
    ori_weights: np.ndarray = _arr(2.0, 2.0, 0.15)
    # regularisation
    w_posture: float = 5.0e-1
    w_v: float = 2.5e-1
    w_tau: float = 1.0e-3
    w_tau_smooth: float = 5.0e-2
    posture_ref_mode: str = "x0"
    torque_ref_mode: str = "gravity_x0"
    w_tau_soft_limits: float = 0.0
    tau_soft_limit_margin: float = 0.2
    w_q_soft_limits: float = 0.0
    q_soft_limit_margin: float = 0.05
    # contact phase objectives
    z_contact: float = 0.35
    z_press: float = 0.0020
    w_plane_z: float = 0.0
    w_vz: float = 0.0
    w_tangent_pos: float = 2.0e2
    w_tangent_vel: float = 1.0e2
    # contact modelling
    contact_name: str = "ee_contact"
    contact_model: str = "normal_1d"
    mu: float = 0.6
    friction_margin: float = 1e-3
    w_friction_cone: float = 2.0e2
    w_unilateral: float = 5.0e1
    contact_gains: np.ndarray = _arr(0.0, 60.0)
    contact_inv_damping: float = 1.0e-8
    strict_force_residual_dim: bool = True
    fn_des: float = 8.0
    w_fn: float = 2.0e1
    # orientation damping
    w_wdamp: float = 2.0e1
    w_wdamp_weights: np.ndarray = _arr(1.5, 1.5, 0.2)
    # surface detection
    phase_source: str = "trajectory"
    fn_contact_on: float = 2.0
    fn_contact_off: float = 0.5
    z_contact_band: float = 0.01
    # command safety filtering
    tau_limits: np.ndarray = _arr(87, 87, 87, 87, 12, 12, 12)
    tau_rate_limit: np.ndarray = _arr(450, 450, 450, 450, 180, 180, 180)
    tau_trust_inf: float = 40.0
    tau_smoothing_alpha: float = 0.35
    apply_command_filter: bool = False
    v_damp_weights: np.ndarray = _arr(1.0, 1.0, 1.0, 1.0, 0.4, 0.4, 0.4)
    # solver
    max_iters: int = 20
    use_box_fddp: bool = True
    mpc_update_steps: int = 1
    use_feedback_policy: bool = True
    feedback_gain_scale: float = 1.0
    verbose: bool = False
    debug_every: int = 25
    max_solver_cost: float = 1.0e8
    max_tau_raw_inf: float = 3.0e2
    fallback_dq_damping: float = 5.0
    contact_release_steps: int = 25


@dataclass
class ClassicalMPCConfig(_CommonMPCConfig):
    """crocoddyl_classical.py:12-110 (same defaults)."""


@dataclass
class ForceFeedbackMPCConfig(_CommonMPCConfig):
    """crocoddyl_force_feedback.py:12-146 (defaults that differ from the classical config are
    overridden below)."""

    w_w: float = 8.0e-4
    w_y: float = 3.0e-3
    y_q_weights: np.ndarray = _arr(0.2, 0.2, 0.2, 0.2, 0.1, 0.1, 0.1)
    y_v_weights: np.ndarray = _arr(0.08, 0.08, 0.08, 0.08, 0.05, 0.05, 0.05)
    y_tau_weights: np.ndarray = _arr(0.35, 0.35, 0.35, 0.35, 0.2, 0.2, 0.2)
    use_inner_state_reg: bool = True
    use_inner_tau_reg: bool = True
    w_tau_soft_limits: float = 1.5
    w_w_soft_limits: float = 2.0
    w_q_soft_limits: float = 8.0
    feedback_gain_scale: float = 0.35
    ff_cutoff_hz: float = 18.0
    ff_alpha_override: Optional[float] = None
    ff_use_tau_meas_filt: bool = True
    ff_tau_state_source: str = "tau_meas_act_filt"
    ff_use_tau_interpolation: bool = True
    ff_align_force_prediction: bool = True
    ff_align_window: int = 240
    ff_align_min_samples: int = 80
    ff_align_corr_threshold: float = 0.05
    ff_align_max_lag: int = 8
    ff_inverse_actuation_model: bool = False
    ff_tau_feedback_gain: float = 1.0


# ------------------------------------------------------------------------------------------------
# runner presets (values of SURVEY.md App. B)
# ------------------------------------------------------------------------------------------------
def benchmark_classical_config(**over) -> ClassicalMPCConfig:
    """The `benchmark_mode` preset of src/run/run_classical.py:269-315."""
    kw = dict(
        horizon=36, dt=0.005, dt_ocp=0.01, z_contact=0.342, z_press=0.0065, w_ee_pos=1.2e3, w_ee_ori=5.0e1,
        ori_weights=np.array([2.4, 2.4, 0.3]), w_posture=1.5e-1, w_v=8.0e-2, posture_ref_mode="q_nom",
        w_tau=8.0e-4, torque_ref_mode="gravity_x0", w_tau_soft_limits=2.0, w_q_soft_limits=8.0,
        q_soft_limit_margin=0.05, w_tau_smooth=0.0, w_tangent_pos=2.6e3, w_tangent_vel=7.0e2, w_plane_z=1.2e3,
        w_vz=5.0e2, w_friction_cone=0.0, w_unilateral=3.0e1, mu=1.0, contact_gains=np.array([140.0, 80.0]),
        fn_des=22.0, w_fn=2.8e1, w_wdamp=6.0e1, w_wdamp_weights=np.array([1.8, 1.8, 0.3]), fn_contact_on=1.0,
        fn_contact_off=0.1, z_contact_band=0.012, max_iters=10, mpc_update_steps=1, use_feedback_policy=True,
        feedback_gain_scale=0.55, max_solver_cost=1.0e8, max_tau_raw_inf=3.0e2, contact_release_steps=60,
        contact_model="normal_1d", phase_source="trajectory", apply_command_filter=False,
        strict_force_residual_dim=True, debug_every=100,
    )
    kw.update(over)
    return ClassicalMPCConfig(**kw)


def benchmark_ff_config(**over) -> ForceFeedbackMPCConfig:
    """The `benchmark_mode` preset of src/run/run_force_feedback.py:272-330."""
    kw = dict(
        horizon=40, dt=0.005, dt_ocp=0.01, z_contact=0.342, z_press=0.0065, w_ee_pos=1.2e3, w_ee_ori=4.5e1,
        ori_weights=np.array([2.2, 2.2, 0.3]), w_posture=1.0e-1, w_v=5.0e-2, posture_ref_mode="q_nom", w_tau=8.0e-4,
        w_w=6.0e-4, w_w_soft_limits=2.0, w_y=8.0e-4, y_q_weights=np.array([0.15] * 4 + [0.08] * 3),
        y_v_weights=np.array([0.05] * 4 + [0.03] * 3), y_tau_weights=np.array([0.12] * 4 + [0.08] * 3),
        use_inner_state_reg=True, use_inner_tau_reg=True, torque_ref_mode="gravity_x0", w_tau_soft_limits=1.5,
        w_q_soft_limits=8.0, q_soft_limit_margin=0.05, w_tau_smooth=0.0, w_tangent_pos=3.6e3, w_tangent_vel=1.2e3,
        w_plane_z=9.0e2, w_vz=3.0e2, w_friction_cone=0.0, w_unilateral=3.0e1, mu=1.0,
        contact_gains=np.array([145.0, 85.0]), fn_des=22.0, w_fn=3.0e1, w_wdamp=7.0e1,
        w_wdamp_weights=np.array([1.8, 1.8, 0.3]), fn_contact_on=1.0, fn_contact_off=0.1, z_contact_band=0.012,
        max_iters=10, mpc_update_steps=1, use_feedback_policy=True, feedback_gain_scale=0.55, max_solver_cost=1.0e8,
        max_tau_raw_inf=3.0e2, contact_release_steps=80, contact_model="normal_1d", phase_source="trajectory",
        apply_command_filter=False, strict_force_residual_dim=True, ff_tau_state_source="tau_meas_act_filt",
        ff_cutoff_hz=25.0, ff_inverse_actuation_model=True, ff_tau_feedback_gain=1.0, debug_every=500,
    )
    kw.update(over)
    return ForceFeedbackMPCConfig(**kw)


def dev_classical_config(**over) -> ClassicalMPCConfig:
    """The development (non-benchmark) preset of src/run/run_classical.py:316-361."""
    kw = dict(
        horizon=50, dt=0.005, dt_ocp=0.01, z_contact=0.342, z_press=0.0075, w_ee_pos=1.3e3, w_ee_ori=6.0e1,
        ori_weights=np.array([2.4, 2.4, 0.2]), w_posture=1.5e-1, w_v=8.0e-2, posture_ref_mode="q_nom", w_tau=2.0e-3,
        torque_ref_mode="gravity_x0", w_tau_soft_limits=4.0, w_q_soft_limits=10.0, q_soft_limit_margin=0.10,
        w_tau_smooth=5.0e-2, w_tangent_pos=3.5e3, w_tangent_vel=1.3e3, w_plane_z=6.0e2, w_vz=2.5e2, w_friction_cone=0.0,
        w_unilateral=4.0e1, mu=1.0, contact_gains=np.array([150.0, 85.0]), fn_des=30.0, w_fn=5.5e1, w_wdamp=8.0e1,
        w_wdamp_weights=np.array([2.0, 2.0, 0.3]), fn_contact_on=1.0, fn_contact_off=0.05, z_contact_band=0.012,
        max_iters=10, mpc_update_steps=1, use_feedback_policy=True, feedback_gain_scale=0.60, max_tau_raw_inf=2.0e2,
        contact_release_steps=80, contact_model="normal_1d", phase_source="trajectory", apply_command_filter=False,
        strict_force_residual_dim=True, debug_every=100,
    )
    kw.update(over)
    return ClassicalMPCConfig(**kw)


def dev_ff_config(**over) -> ForceFeedbackMPCConfig:
    """The development (non-benchmark) preset of src/run/run_force_feedback.py:331-387."""
    kw = dict(
        horizon=50, dt=0.005, dt_ocp=0.01, z_contact=0.342, z_press=0.0080, w_ee_pos=1.8e3, w_ee_ori=5.5e1,
        ori_weights=np.array([2.4, 2.4, 0.2]), w_posture=6.0e-2, w_v=2.0e-2, posture_ref_mode="q_nom", w_tau=2.0e-3,
        w_w=3.0e-4, w_w_soft_limits=3.0, w_y=5.0e-4, y_q_weights=np.array([0.2] * 4 + [0.1] * 3),
        y_v_weights=np.array([0.08] * 4 + [0.05] * 3), y_tau_weights=np.array([0.10] * 4 + [0.06] * 3),
        use_inner_state_reg=True, use_inner_tau_reg=True, torque_ref_mode="gravity_x0", w_tau_soft_limits=1.5,
        w_q_soft_limits=10.0, q_soft_limit_margin=0.10, w_tau_smooth=5.0e-2, w_tangent_pos=4.8e3, w_tangent_vel=1.9e3,
        w_plane_z=5.0e2, w_vz=2.0e2, w_friction_cone=0.0, w_unilateral=3.0e1, mu=1.0,
        contact_gains=np.array([150.0, 90.0]), fn_des=26.0, w_fn=4.0e1, w_wdamp=8.0e1,
        w_wdamp_weights=np.array([2.0, 2.0, 0.3]), fn_contact_on=1.0, fn_contact_off=0.05, z_contact_band=0.012,
        max_iters=10, mpc_update_steps=1, use_feedback_policy=True, feedback_gain_scale=0.60, max_tau_raw_inf=2.2e2,
        contact_release_steps=80, contact_model="normal_1d", phase_source="trajectory", apply_command_filter=False,
        strict_force_residual_dim=True, ff_tau_state_source="tau_meas_act_filt", ff_cutoff_hz=90.0,
        ff_inverse_actuation_model=True, debug_every=500,
    )
    kw.update(over)
    return ForceFeedbackMPCConfig(**kw)


PRESETS = {
    ("benchmark", False): benchmark_classical_config, ("benchmark", True): benchmark_ff_config,
    ("dev", False): dev_classical_config, ("dev", True): dev_ff_config,
    ("defaults", False): ClassicalMPCConfig, ("defaults", True): ForceFeedbackMPCConfig,   # the dataclass defaults
}


# ------------------------------------------------------------------------------------------------
# lowering to the C struct
# ------------------------------------------------------------------------------------------------
def dt_ocp_of(cfg) -> float:
    """`_dt_ocp` property (crocoddyl_classical.py:184-187)."""
    dt = cfg.dt if cfg.dt_ocp is None else float(cfg.dt_ocp)
    return float(max(dt, 1.0e-6))


def ff_alpha_ocp(cfg) -> float:
    """`_ff_alpha_ocp` (crocoddyl_force_feedback.py:493-497)."""
    if cfg.ff_alpha_override is not None:
        return float(np.clip(float(cfg.ff_alpha_override), 0.0, 0.999999))
    wc = 2.0 * math.pi * float(max(cfg.ff_cutoff_hz, 0.0))
    return float(np.clip(math.exp(-wc * dt_ocp_of(cfg)), 0.0, 0.999999))


def contact_nc_of(cfg) -> int:
    """Contact-model switch of `_make_dam` (crocoddyl_classical.py:637-643)."""
    mode = str(cfg.contact_model).strip().lower()
    return 3 if mode in ("point3d", "3d", "rigid3d", "route_a_3d") else 1


def q_soft_limit_terms(cfg, model: PandaModel, q_nom: np.ndarray):
    """Residual reference and barrier bounds of `_make_q_soft_limit_cost`
    (crocoddyl_classical.py:487-519): residual = q - q_ref, bounds on the residual."""
    q_lb = np.asarray(model.lowerPositionLimit, dtype=float)
    q_ub = np.asarray(model.upperPositionLimit, dtype=float)
    fin_lb, fin_ub = np.isfinite(q_lb), np.isfinite(q_ub)
    both = fin_lb & fin_ub
    q_ref = np.asarray(q_nom, dtype=float).copy()
    q_ref[both] = 0.5 * (q_lb[both] + q_ub[both])
    margin = float(max(cfg.q_soft_limit_margin, 0.0))
    lo, hi = q_lb.copy(), q_ub.copy()
    lo[fin_lb] += margin
    hi[fin_ub] -= margin
    bad = both & (lo > hi)
    if np.any(bad):
        mid = 0.5 * (q_lb[bad] + q_ub[bad])
        lo[bad], hi[bad] = mid - 1.0e-3, mid + 1.0e-3
    lb = np.full(7, -np.inf)
    ub = np.full(7, np.inf)
    lb[fin_lb] = lo[fin_lb] - q_ref[fin_lb]
    ub[fin_ub] = hi[fin_ub] - q_ref[fin_ub]
    return q_ref, lb, ub


def to_abi(cfg, model: PandaModel, q_nom: np.ndarray, R_des: Optional[np.ndarray] = None) -> _abi.Config:
    """Lower a Classical/ForceFeedback config to ``ffmpc_config``.

    Raises ``ValueError`` for configurations the engine does not implement, mirroring the
    reference's construction-time errors (crocoddyl_classical.py:141-142, 869-870)."""
    ff = isinstance(cfg, ForceFeedbackMPCConfig)
    nc = contact_nc_of(cfg)
    if int(cfg.horizon) < 1:
        raise ValueError("horizon must be >= 1")
    c = _abi.Config()
    c.variant = _abi.VARIANT_FF if ff else _abi.VARIANT_CLASSICAL
    c.horizon = int(cfg.horizon)
    c.contact_nc = nc
    c.use_box_fddp = 1 if cfg.use_box_fddp else 0
    c.use_inner_state_reg = int(bool(getattr(cfg, "use_inner_state_reg", True)))
    c.use_inner_tau_reg = int(bool(getattr(cfg, "use_inner_tau_reg", True)))
    c.dt_ocp = dt_ocp_of(cfg)
    c.w_ee_pos = float(cfg.w_ee_pos)
    c.w_ee_ori = float(cfg.w_ee_ori)
    _abi.set_array(c.ori_weights, cfg.ori_weights)
    _abi.set_array(c.R_des, reference_R_des(model) if R_des is None else R_des)
    c.w_posture = float(cfg.w_posture)
    c.w_v = float(cfg.w_v)
    _abi.set_array(c.v_damp_weights, cfg.v_damp_weights)
    c.w_q_soft_limits = float(max(cfg.w_q_soft_limits, 0.0))
    q_ref, lb, ub = q_soft_limit_terms(cfg, model, q_nom)
    _abi.set_array(c.q_soft_ref, q_ref)
    _abi.set_array(c.q_soft_lb, lb)
    _abi.set_array(c.q_soft_ub, ub)
    c.w_wdamp = float(cfg.w_wdamp)
    _abi.set_array(c.w_wdamp_weights, cfg.w_wdamp_weights)
    c.w_tau = float(cfg.w_tau)
    c.w_tau_soft_limits = float(max(cfg.w_tau_soft_limits, 0.0))
    c.tau_soft_limit_margin = float(cfg.tau_soft_limit_margin)
    _abi.set_array(c.tau_limits, cfg.tau_limits)
    c.z_press = float(cfg.z_press)
    c.w_tangent_pos = float(cfg.w_tangent_pos)
    c.w_tangent_vel = float(cfg.w_tangent_vel)
    c.w_plane_z = float(max(cfg.w_plane_z, 0.0))
    c.w_vz = float(max(cfg.w_vz, 0.0))
    c.w_unilateral = float(max(cfg.w_unilateral, 0.0))
    c.friction_margin = float(cfg.friction_margin)
    c.w_fn = float(max(cfg.w_fn, 0.0))
    c.fn_des = float(cfg.fn_des)
    _abi.set_array(c.contact_gains, cfg.contact_gains)
    c.contact_inv_damping = float(cfg.contact_inv_damping)
    if ff:
        c.ff_alpha = ff_alpha_ocp(cfg)
        c.w_w = float(cfg.w_w)
        c.w_w_soft_limits = float(cfg.w_w_soft_limits)
        c.w_y = float(cfg.w_y)
        _abi.set_array(c.y_reg_weights, np.concatenate([cfg.y_q_weights, cfg.y_v_weights, cfg.y_tau_weights]))
    # crocoddyl solver defaults
    c.reg_min, c.reg_max, c.reg_incfactor, c.reg_decfactor = 1e-9, 1e9, 10.0, 10.0
    c.th_stepdec, c.th_stepinc, c.th_acceptstep, c.th_acceptnegstep, c.th_grad = 0.5, 0.01, 0.1, 2.0, 1e-12
    c.th_stop = 5e-5 if cfg.use_box_fddp else 1e-9
    c.n_alphas = 10
    c.boxqp_maxiter, c.boxqp_th_acceptstep, c.boxqp_th_grad, c.boxqp_reg = 100, 0.1, 1e-5, 0.0
    # friction cone: only built by the reference for point3d (crocoddyl_classical.py:678)
    c.w_friction_cone = float(cfg.w_friction_cone) if nc == 3 else 0.0
    c.mu = float(cfg.mu)
    return c


def tick_policy(cfg, dt_mpc: Optional[float] = None) -> _abi.TickPolicy:
    """`ffmpc_tick_policy` of a controller config: the per-tick host logic of `compute_control` the batched tick runs on the
    device (policy scale, Eq. 14 interpolation, unstable guard, `_safe_tau`).  dt_mpc = sim.dt (default cfg.dt)."""
    p = _abi.TickPolicy()
    dt = float(cfg.dt if dt_mpc is None else dt_mpc)
    p.use_feedback_policy = int(bool(cfg.use_feedback_policy))
    p.ff_use_tau_interpolation = int(bool(getattr(cfg, "ff_use_tau_interpolation", True)))
    p.ff_inverse_actuation_model = int(bool(getattr(cfg, "ff_inverse_actuation_model", False)))
    p.apply_command_filter = int(bool(cfg.apply_command_filter))
    p.feedback_gain_scale = float(cfg.feedback_gain_scale)
    p.dt_mpc = dt
    if getattr(cfg, "ff_alpha_override", None) is not None:
        p.ff_alpha_ctrl = float(np.clip(float(cfg.ff_alpha_override), 0.0, 0.999999))
    else:
        p.ff_alpha_ctrl = float(np.clip(np.exp(-2.0 * np.pi * float(max(getattr(cfg, "ff_cutoff_hz", 0.0), 0.0)) * dt), 0.0, 0.999999))
    p.max_solver_cost, p.max_tau_raw_inf = float(cfg.max_solver_cost), float(cfg.max_tau_raw_inf)
    p.fallback_dq_damping = float(cfg.fallback_dq_damping)
    p.tau_trust_inf, p.tau_smoothing_alpha = float(cfg.tau_trust_inf), float(cfg.tau_smoothing_alpha)
    _abi.set_array(p.tau_rate_limit, np.asarray(cfg.tau_rate_limit, dtype=float))
    return p
