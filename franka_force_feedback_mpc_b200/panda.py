"""Panda 7-DoF arm constants and small host-side (numpy) kinematics helpers.

The reference loads its model with ``example_robot_data.load("panda")`` + ``buildReducedModel``
(src/mpc/crocoddyl_classical.py:137-139, 189-197); neither package exists in this image.  The only
Panda kinematic / inertial data under /root/reference is the MuJoCo model
``assets/scenes/panda_robot.xml:113-189`` (a "nohand" Panda), transcribed here.
The MuJoCo base yaw (``quat="0 0 0 1"``, :98) is dropped: these constants live in the
"Pinocchio world" of the reference, ``p_mj = diag(-1,-1,1) p_pin`` (crocoddyl_classical.py:149-151).

A different model (e.g. the reduced Pinocchio model exported by ``tools/export_crocoddyl_golden.py``)
can be loaded from JSON with :func:`load_model_json` without recompiling anything.

The numpy helpers below are HOST-side conveniences (controller calibration at construction time,
synthetic-input generation).  They are not on the solve path: that path is CUDA only.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

from . import _abi

# parent -> joint translation, quaternion (w,x,y,z) as written in the MJCF, mass, CoM,
# fullinertia (Ixx Iyy Izz Ixy Ixz Iyz about the CoM), joint range
_LINKS = [
    # link1
    dict(pos=(0.0, 0.0, 0.333), quat=(1, 0, 0, 0), mass=4.970684, com=(0.003875, 0.002081, -0.04762),
         inertia=(0.70337, 0.70661, 0.009117, -0.000139, 0.006772, 0.019169), range=(-2.8973, 2.8973)),
    # link2
    dict(pos=(0.0, 0.0, 0.0), quat=(1, -1, 0, 0), mass=0.646926, com=(-0.003141, -0.02872, 0.003495),
         inertia=(0.007962, 0.02811, 0.025995, -0.003925, 0.010254, 0.000704), range=(-1.7628, 1.7628)),
    # link3
    dict(pos=(0.0, -0.316, 0.0), quat=(1, 1, 0, 0), mass=3.228604, com=(0.027518, 0.039252, -0.066502),
         inertia=(0.037242, 0.036155, 0.01083, -0.004761, -0.011396, -0.012805), range=(-2.8973, 2.8973)),
    # link4
    dict(pos=(0.0825, 0.0, 0.0), quat=(1, 1, 0, 0), mass=3.587895, com=(-0.05317, 0.104419, 0.027454),
         inertia=(0.025853, 0.019552, 0.028323, 0.007796, -0.001332, 0.008641), range=(-3.0718, -0.0698)),
    # link5
    dict(pos=(-0.0825, 0.384, 0.0), quat=(1, -1, 0, 0), mass=1.225946, com=(-0.011953, 0.041065, -0.038437),
         inertia=(0.035549, 0.029474, 0.008627, -0.002117, -0.004037, 0.000229), range=(-2.8973, 2.8973)),
    # link6
    dict(pos=(0.0, 0.0, 0.0), quat=(1, 1, 0, 0), mass=1.666555, com=(0.060149, -0.014117, -0.010517),
         inertia=(0.001964, 0.004354, 0.005433, 0.000109, -0.001158, 0.000341), range=(-0.0175, 3.7525)),
    # link7
    dict(pos=(0.088, 0.0, 0.0), quat=(1, 1, 0, 0), mass=0.735522, com=(0.010517, -0.004252, 0.061597),
         inertia=(0.012516, 0.010027, 0.004815, -0.000428, -0.001196, -0.000741), range=(-2.8973, 2.8973)),
]
EE_POS = (0.0, 0.0, 0.107)  # panda_link8 / MJCF "tool" body (:189)
# MuJoCo ee_site orientation in the tool body (:189); only used to emulate the site calibration
EE_SITE_QUAT = (0.3826834, 0.0, 0.0, 0.9238795)
Q_NEUTRAL = np.array([0.0, -0.758, 0.0, -2.22, 0.0, 1.43, 0.0])  # keyframe "neutral" (:233)
TAU_LIMITS = np.array([87.0, 87.0, 87.0, 87.0, 12.0, 12.0, 12.0])  # (:10, 218-221)
R_MJ_FROM_PIN = np.diag([-1.0, -1.0, 1.0])


def quat_to_rot(q) -> np.ndarray:
    w, x, y, z = np.asarray(q, dtype=float) / np.linalg.norm(q)
    return np.array(
        [
            [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
            [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
            [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)],
        ]
    )


@dataclass
class PandaModel:
    joint_pos: np.ndarray = field(default_factory=lambda: np.array([l["pos"] for l in _LINKS], dtype=float))
    joint_rot: np.ndarray = field(default_factory=lambda: np.array([quat_to_rot(l["quat"]) for l in _LINKS]))
    mass: np.ndarray = field(default_factory=lambda: np.array([l["mass"] for l in _LINKS], dtype=float))
    com: np.ndarray = field(default_factory=lambda: np.array([l["com"] for l in _LINKS], dtype=float))
    inertia: np.ndarray = field(default_factory=lambda: np.array([l["inertia"] for l in _LINKS], dtype=float))
    ee_pos: np.ndarray = field(default_factory=lambda: np.array(EE_POS, dtype=float))
    ee_rot: np.ndarray = field(default_factory=lambda: np.eye(3))
    gravity: np.ndarray = field(default_factory=lambda: np.array([0.0, 0.0, -9.81]))
    q_lower: np.ndarray = field(default_factory=lambda: np.array([l["range"][0] for l in _LINKS], dtype=float))
    q_upper: np.ndarray = field(default_factory=lambda: np.array([l["range"][1] for l in _LINKS], dtype=float))

    nq = 7
    nv = 7

    # --- pinocchio-like attribute names the reference's runners read -----------------------------
    @property
    def lowerPositionLimit(self) -> np.ndarray:
        return self.q_lower

    @property
    def upperPositionLimit(self) -> np.ndarray:
        return self.q_upper

    def to_abi(self) -> _abi.Model:
        m = _abi.Model()
        _abi.set_array(m.joint_pos, self.joint_pos)
        _abi.set_array(m.joint_rot, self.joint_rot)
        _abi.set_array(m.mass, self.mass)
        _abi.set_array(m.com, self.com)
        _abi.set_array(m.inertia, self.inertia)
        _abi.set_array(m.ee_pos, self.ee_pos)
        _abi.set_array(m.ee_rot, self.ee_rot)
        _abi.set_array(m.gravity, self.gravity)
        return m

    # --- host-side kinematics ---------------------------------------------------------------------
    def joint_frames(self, q):
        """World rotations / positions of the 7 joint frames."""
        q = np.asarray(q, dtype=float).reshape(7)
        R = np.eye(3)
        p = np.zeros(3)
        Rs, ps = [], []
        for i in range(7):
            c, s = np.cos(q[i]), np.sin(q[i])
            Rz = np.array([[c, -s, 0.0], [s, c, 0.0], [0.0, 0.0, 1.0]])
            p = p + R @ self.joint_pos[i]
            R = R @ self.joint_rot[i] @ Rz
            Rs.append(R)
            ps.append(p)
        return Rs, ps

    def fk(self, q):
        """EE frame (``panda_link8``) placement: (p[3], R[3,3])."""
        Rs, ps = self.joint_frames(q)
        return ps[-1] + Rs[-1] @ self.ee_pos, Rs[-1] @ self.ee_rot

    def frame_jacobian(self, q) -> np.ndarray:
        """6x7 LOCAL_WORLD_ALIGNED Jacobian of the EE frame, rows [linear; angular]."""
        Rs, ps = self.joint_frames(q)
        pf = ps[-1] + Rs[-1] @ self.ee_pos
        J = np.zeros((6, 7))
        for j in range(7):
            z = Rs[j][:, 2]
            J[:3, j] = np.cross(z, pf - ps[j])
            J[3:, j] = z
        return J

    def ik(self, p_des, R_des, q_init, iters: int = 60, damping: float = 1e-4) -> np.ndarray:
        """Damped least-squares IK for synthetic x0 sampling (SURVEY.md 8d `IK_oracle`)."""
        q = np.asarray(q_init, dtype=float).copy()
        for _ in range(iters):
            p, R = self.fk(q)
            e = np.concatenate([p_des - p, _log3(R_des @ R.T)])
            if np.linalg.norm(e) < 1e-10:
                break
            J = self.frame_jacobian(q)
            dq = J.T @ np.linalg.solve(J @ J.T + damping * np.eye(6), e)
            q = np.clip(q + dq, self.q_lower + 0.05, self.q_upper - 0.05)
        return q


def _log3(R: np.ndarray) -> np.ndarray:
    tr = np.clip((np.trace(R) - 1.0) * 0.5, -1.0, 1.0)
    th = np.arccos(tr)
    w = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    if th < 1e-9:
        return 0.5 * w
    return th / (2.0 * np.sin(th)) * w


def default_model() -> PandaModel:
    return PandaModel()


def load_model_json(path: str) -> PandaModel:
    """Load a model exported as JSON (same field names as :class:`PandaModel`)."""
    with open(path) as f:
        d = json.load(f)
    m = PandaModel()
    for k, v in d.items():
        if hasattr(m, k):
            setattr(m, k, np.asarray(v, dtype=float))
    return m


def reference_R_des(model: Optional[PandaModel] = None) -> np.ndarray:
    """``R_des`` exactly as the reference calibrates it (crocoddyl_classical.py:155-157, 199-211,
    241-258) when the plant is the MJCF Panda: the MuJoCo ee_site is the tool body rotated by
    ``EE_SITE_QUAT`` and the MuJoCo world is the Pinocchio world yawed by 180 deg, so
    ``R_site_from_pin_ee = Rz(135 deg)`` independent of q, and
    ``R_des = R_mj_from_pin^T @ [x,y,z]_down @ R_site_from_pin_ee^T``."""
    R_site_from_pin_ee = quat_to_rot(EE_SITE_QUAT)
    z = np.array([0.0, 0.0, -1.0])
    x = np.array([1.0, 0.0, 0.0])
    y = np.cross(z, x)
    y /= np.linalg.norm(y) + 1e-12
    x = np.cross(y, z)
    x /= np.linalg.norm(x) + 1e-12
    R_down = np.column_stack([x, y, z])
    return R_MJ_FROM_PIN.T @ R_down @ R_site_from_pin_ee.T
