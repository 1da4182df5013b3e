"""Synthetic problem instances of the shapes BASELINE.json names.

Everything here is host-side numpy input generation: the reference trajectory of the runners
(src/run/run_classical.py:221-255 -> src/tasks/trajectories.py:8-93, restated, and checked
against the reference's own function in tests/test_reference_pieces.py), batched Panda
kinematics for sampling consistent initial states, and the per-config instance distributions.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

from .panda import PandaModel, Q_NEUTRAL, R_MJ_FROM_PIN, default_model, reference_R_des

SCENARIO_SEEDS = {  # src/run/run_classical.py:30-39
    "flat": 11, "tilted_5": 12, "tilted_10": 13, "tilted_15": 14, "actuation_uncertainty": 15,
}
SCENARIO_TILT_DEG = {"flat": 0.0, "tilted_5": 5.0, "tilted_10": 10.0, "tilted_15": 15.0, "actuation_uncertainty": 0.0}


# ------------------------------------------------------------------------------------------------
# reference trajectory (MuJoCo world), vectorised over query times
# ------------------------------------------------------------------------------------------------
@dataclass
class ApproachThenCircle:
    """Pre-approach (0..t_pre) -> approach (..t_pre+t_approach) -> circle at z_contact, with the
    0.2 s "hold" of the benchmark runners (run_classical.py:247-255).  Restates
    `make_approach_then_circle` (src/tasks/trajectories.py:8-93)."""

    center: np.ndarray
    radius: float = 0.10
    omega: float = 1.5
    z_contact: float = 0.342
    z_pre: float = 0.392
    t_approach: float = 0.55
    t_pre: float = 0.25
    ee_start: Optional[np.ndarray] = None
    t_hold: float = 0.2

    def _base(self, t: np.ndarray) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        t = np.asarray(t, dtype=float)
        c = np.asarray(self.center, dtype=float)
        T_a = max(float(self.t_approach), 1e-6)
        T_p = max(float(self.t_pre), 0.0)
        p_cs = np.array([c[0] + self.radius, c[1], self.z_contact])
        p_start = p_cs + np.array([0.0, 0.0, 0.08]) if self.ee_start is None else np.asarray(self.ee_start, float)
        p_pre = np.array([p_cs[0], p_cs[1], self.z_pre])

        def blend(p0, p1, tau, T):
            s_lin = np.clip(tau / T, 0.0, 1.0)
            s = s_lin * s_lin * (3.0 - 2.0 * s_lin)
            ds = 6.0 * s_lin * (1.0 - s_lin) / T
            return (1.0 - s)[..., None] * p0 + s[..., None] * p1, ds[..., None] * (p1 - p0)

        p = np.zeros(t.shape + (3,))
        v = np.zeros(t.shape + (3,))
        surf = np.zeros(t.shape, dtype=bool)
        in_pre = (T_p > 0.0) & (t < T_p)
        in_app = (~in_pre) & (t < T_p + T_a)
        in_circ = ~(in_pre | in_app)
        pp, vp = blend(p_start, p_pre, t, T_p if T_p > 0 else 1.0)
        pa, va = blend(p_pre if T_p > 0 else p_start, p_cs, t - T_p, T_a)
        th = self.omega * (t - (T_p + T_a))
        pc = np.stack([c[0] + self.radius * np.cos(th), c[1] + self.radius * np.sin(th), np.full(t.shape, self.z_contact)], -1)
        vc = np.stack([-self.radius * self.omega * np.sin(th), self.radius * self.omega * np.cos(th), np.zeros(t.shape)], -1)
        for mask, pm, vm in ((in_pre, pp, vp), (in_app, pa, va), (in_circ, pc, vc)):
            p[mask] = pm[mask]
            v[mask] = vm[mask]
        surf[in_circ] = True
        return p, v, surf

    def __call__(self, t) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        """(p_ref, v_ref, surface) in the MuJoCo world, for scalar or array `t`."""
        t = np.asarray(t, dtype=float)
        p, v, surf = self._base(t)
        t_c = float(max(self.t_pre, 0.0) + max(self.t_approach, 1e-6))
        hold = surf & (t < t_c + self.t_hold)
        if np.any(hold):
            p_hold, _, _ = self._base(np.asarray(t_c))
            p[hold] = p_hold
            v[hold] = 0.0
        return p, v, surf


def benchmark_trajectory(model: Optional[PandaModel] = None) -> ApproachThenCircle:
    """Trajectory of the benchmark runners: table top z=0.32 + tool radius 0.03 - 8 mm = 0.342,
    circle centre = table centre (-0.5, 0) in the MuJoCo world (run_classical.py:221-246)."""
    model = model or default_model()
    p_pin, _ = model.fk(Q_NEUTRAL)
    return ApproachThenCircle(center=np.array([-0.5, 0.0, 0.342]), radius=0.10, omega=1.5, z_contact=0.342,
                              z_pre=0.392, t_approach=0.55, t_pre=0.25, ee_start=R_MJ_FROM_PIN @ p_pin)


def refs_pin(traj: ApproachThenCircle, t0: np.ndarray, N: int, dt_ocp: float,
             p_site_minus_frame_pin: Optional[np.ndarray] = None):
    """p_ref/v_ref[B, N+1, 3] in the Pinocchio world at t0 + k dt_ocp (crocoddyl_classical.py:526-530,
    250-255: p_pin = R^T p_mj - offset, v_pin = R^T v_mj)."""
    t0 = np.asarray(t0, dtype=float).reshape(-1)
    tk = t0[:, None] + dt_ocp * np.arange(N + 1)[None, :]
    p_mj, v_mj, _ = traj(tk)
    off = np.zeros(3) if p_site_minus_frame_pin is None else np.asarray(p_site_minus_frame_pin, float)
    return p_mj @ R_MJ_FROM_PIN - off, v_mj @ R_MJ_FROM_PIN  # R is symmetric/diagonal: p @ R == (R^T p)^T


# ------------------------------------------------------------------------------------------------
# batched kinematics (numpy)
# ------------------------------------------------------------------------------------------------
def batched_frames(model: PandaModel, q: np.ndarray):
    """World joint rotations [B,7,3,3], positions [B,7,3], EE position [B,3], EE rotation [B,3,3]."""
    q = np.asarray(q, dtype=float)
    B = q.shape[0]
    R = np.broadcast_to(np.eye(3), (B, 3, 3)).copy()
    p = np.zeros((B, 3))
    Rs = np.zeros((B, 7, 3, 3))
    ps = np.zeros((B, 7, 3))
    for i in range(7):
        c, s = np.cos(q[:, i]), np.sin(q[:, i])
        Rz = np.zeros((B, 3, 3))
        Rz[:, 0, 0], Rz[:, 0, 1], Rz[:, 1, 0], Rz[:, 1, 1], Rz[:, 2, 2] = c, -s, s, c, 1.0
        p = p + R @ model.joint_pos[i]
        R = R @ model.joint_rot[i] @ Rz
        Rs[:, i], ps[:, i] = R, p
    pf = p + R @ model.ee_pos
    return Rs, ps, pf, R @ model.ee_rot


def batched_jacobian(model: PandaModel, q: np.ndarray):
    Rs, ps, pf, Rf = batched_frames(model, q)
    z = Rs[:, :, :, 2]  # [B,7,3]
    Jl = np.cross(z, pf[:, None, :] - ps)  # [B,7,3]
    J = np.concatenate([Jl, z], axis=2).transpose(0, 2, 1)  # [B,6,7]
    return J, pf, Rf, Rs, ps


def batched_gravity_torque(model: PandaModel, q: np.ndarray) -> np.ndarray:
    """g(q) = RNEA(q,0,0): tau_j = -sum_{i>=j} m_i g . (z_j x (c_i - p_j))."""
    Rs, ps, _, _ = batched_frames(model, q)
    cw = ps + np.einsum("bijk,ik->bij", Rs, model.com)  # [B,7,3]
    z = Rs[:, :, :, 2]
    tau = np.zeros((q.shape[0], 7))
    for j in range(7):
        for i in range(j, 7):
            tau[:, j] -= model.mass[i] * np.einsum("bk,k->b", np.cross(z[:, j], cw[:, i] - ps[:, j]), model.gravity)
    return tau


def _log3_batched(R: np.ndarray) -> np.ndarray:
    tr = np.clip((np.trace(R, axis1=1, axis2=2) - 1.0) * 0.5, -1.0, 1.0)
    th = np.arccos(tr)
    w = np.stack([R[:, 2, 1] - R[:, 1, 2], R[:, 0, 2] - R[:, 2, 0], R[:, 1, 0] - R[:, 0, 1]], -1)
    sc = np.where(th < 1e-9, 0.5, th / (2.0 * np.sin(np.where(th < 1e-9, 1.0, th))))
    return sc[:, None] * w


def batched_ik(model: PandaModel, p_des: np.ndarray, R_des: np.ndarray, q_init: np.ndarray, iters: int = 40,
               damping: float = 1e-4) -> np.ndarray:
    """Damped least-squares IK (`IK_oracle` of SURVEY.md 8d), vectorised over instances."""
    q = np.array(q_init, dtype=float)
    if q.ndim == 1:
        q = np.broadcast_to(q, (p_des.shape[0], 7)).copy()
    for _ in range(iters):
        J, pf, Rf, _, _ = batched_jacobian(model, q)
        e = np.concatenate([p_des - pf, _log3_batched(R_des[None] @ Rf.transpose(0, 2, 1))], axis=1)
        A = J @ J.transpose(0, 2, 1) + damping * np.eye(6)
        dq = np.einsum("bij,bi->bj", J, np.linalg.solve(A, e[..., None])[..., 0])
        nrm = np.linalg.norm(dq, axis=1, keepdims=True)
        dq = dq * np.minimum(1.0, 0.5 / np.maximum(nrm, 1e-12))
        q = np.clip(q + dq, model.q_lower + 0.05, model.q_upper - 0.05)
    return q


# ------------------------------------------------------------------------------------------------
# instance sampling
# ------------------------------------------------------------------------------------------------
def sample_instances(B: int, N: int, variant_ff: bool, seed: int, contact_fraction: float = 1.0,
                     scenario: str = "flat", dt_ocp: float = 0.01, model: Optional[PandaModel] = None,
                     fn_bias: float = 22.0) -> dict:
    """Synthetic batch per SURVEY.md 8d.  Returns numpy arrays
    mode[B], x0[B,nx], p_ref/v_ref[B,N+1,3], x_reg_ref[B,14] (posture_ref_mode="q_nom"), t0[B].

    * contact instances: t ~ U(1.0, 20.0); free instances: t ~ U(0, 0.8)
    * q0 = IK(p_ref(t), R_des) + N(0, 0.02^2) clipped to limits -/+ 0.05; v0 = J^+ v_ref + N(0, 0.05^2)
    * FF: tau_hat = g(q0) + J^T [0,0,-fn_bias] (contact) + N(0, 0.5^2)
    * tilted_*: EE z-offset dz = -x_rel tan(tilt) about world y through the table centre
      (run_classical.py:94-106); the tilt is hidden from the controller so the OCP is unchanged
    * actuation_uncertainty: obs noise sigma_q=5e-4, sigma_dq=2e-3 and
      tau_hat = a tau + b + N(0,0.05^2), a~U(0.95,1.05), b~U(-0.1,0.1) (uncertainty_profiles.py:21-30)
    """
    model = model or default_model()
    rng = np.random.default_rng(seed)
    traj = benchmark_trajectory(model)
    R_des = reference_R_des(model)
    n_contact = int(round(B * contact_fraction))
    mode = np.zeros(B, dtype=np.int32)
    mode[:n_contact] = 1
    rng.shuffle(mode)
    t0 = np.where(mode == 1, rng.uniform(1.0, 20.0, B), rng.uniform(0.0, 0.8, B))
    p_ref, v_ref = refs_pin(traj, t0, N, dt_ocp)

    p0 = p_ref[:, 0].copy()
    tilt = np.deg2rad(SCENARIO_TILT_DEG.get(scenario, 0.0))
    if tilt != 0.0:
        x_rel_mj = (R_MJ_FROM_PIN @ p0.T).T[:, 0] - (-0.5)
        p0[:, 2] += np.where(mode == 1, -x_rel_mj * np.tan(tilt), 0.0)
    q_seed = Q_NEUTRAL.copy()
    q0 = batched_ik(model, p0, R_des, q_seed)
    q0 = np.clip(q0 + rng.normal(0.0, 0.02, (B, 7)), model.q_lower + 0.05, model.q_upper - 0.05)
    J, _, _, _, _ = batched_jacobian(model, q0)
    Jl = J[:, :3]
    v0 = np.einsum("bij,bi->bj", Jl, np.linalg.solve(Jl @ Jl.transpose(0, 2, 1) + 1e-6 * np.eye(3), v_ref[:, 0][..., None])[..., 0])
    v0 = v0 + rng.normal(0.0, 0.05, (B, 7))
    if scenario == "actuation_uncertainty":
        q0 = q0 + rng.normal(0.0, 5e-4, (B, 7))
        v0 = v0 + rng.normal(0.0, 2e-3, (B, 7))
    x_mb = np.concatenate([q0, v0], axis=1)
    out = dict(mode=mode, t0=t0, p_ref=p_ref, v_ref=v_ref,
               x_reg_ref=np.broadcast_to(np.concatenate([Q_NEUTRAL, np.zeros(7)]), (B, 14)).copy())
    if variant_ff:
        g = batched_gravity_torque(model, q0)
        f = np.zeros((B, 3))
        f[:, 2] = np.where(mode == 1, -fn_bias, 0.0)
        tau = g + np.einsum("bij,bi->bj", Jl, f)
        if scenario == "actuation_uncertainty":
            a = rng.uniform(0.95, 1.05, (B, 1))
            b = rng.uniform(-0.1, 0.1, (B, 7))
            tau = a * tau + b + rng.normal(0.0, 0.05, (B, 7))
        tau = tau + rng.normal(0.0, 0.5, (B, 7))
        out["x0"] = np.concatenate([x_mb, tau], axis=1)
    else:
        out["x0"] = x_mb
    return out
