"""ctypes mirror of ``include/ffmpc.h`` (the C-ABI drop-in boundary).

Field order and types must match the header exactly; ``tests/test_abi.py`` checks the struct
sizes against the compiled library (``ffmpc_abi_sizes``).
"""
from __future__ import annotations

import ctypes as C

NJ = 7
NX_MB = 14
NX_FF = 21

OK = 0
ERR_ARG = 1
ERR_CUDA = 2
ERR_NO_DEVICE = 3
ERR_CAPACITY = 4
STATIC_MAX_POINTS = 512  # FFMPC_STATIC_MAX_POINTS / FFMPC_STATIC_MAX_BATCH (ffmpc_node_eval_static)
STATIC_MAX_BATCH = 512

VARIANT_CLASSICAL = 0
VARIANT_FF = 1
MODE_FREE = 0
MODE_CONTACT = 1

NFAM = 4
FAMILY_NAMES = ("other", "node_diff", "backward", "forward")

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class Model(C.Structure):
    _fields_ = [
        ("joint_pos", (C.c_double * 3) * NJ),
        ("joint_rot", (C.c_double * 9) * NJ),
        ("mass", C.c_double * NJ),
        ("com", (C.c_double * 3) * NJ),
        ("inertia", (C.c_double * 6) * NJ),
        ("ee_pos", C.c_double * 3),
        ("ee_rot", C.c_double * 9),
        ("gravity", C.c_double * 3),
    ]


class Config(C.Structure):
    _fields_ = [
        ("variant", C.c_int),
        ("horizon", C.c_int),
        ("contact_nc", C.c_int),
        ("use_box_fddp", C.c_int),
        ("use_inner_state_reg", C.c_int),
        ("use_inner_tau_reg", C.c_int),
        ("dt_ocp", C.c_double),
        ("w_ee_pos", C.c_double),
        ("w_ee_ori", C.c_double),
        ("ori_weights", C.c_double * 3),
        ("R_des", C.c_double * 9),
        ("w_posture", C.c_double),
        ("w_v", C.c_double),
        ("v_damp_weights", C.c_double * NJ),
        ("w_q_soft_limits", C.c_double),
        ("q_soft_ref", C.c_double * NJ),
        ("q_soft_lb", C.c_double * NJ),
        ("q_soft_ub", C.c_double * NJ),
        ("w_wdamp", C.c_double),
        ("w_wdamp_weights", C.c_double * 3),
        ("w_tau", C.c_double),
        ("w_tau_soft_limits", C.c_double),
        ("tau_soft_limit_margin", C.c_double),
        ("tau_limits", C.c_double * NJ),
        ("z_press", C.c_double),
        ("w_tangent_pos", C.c_double),
        ("w_tangent_vel", C.c_double),
        ("w_plane_z", C.c_double),
        ("w_vz", C.c_double),
        ("w_unilateral", C.c_double),
        ("friction_margin", C.c_double),
        ("w_fn", C.c_double),
        ("fn_des", C.c_double),
        ("contact_gains", C.c_double * 2),
        ("contact_inv_damping", C.c_double),
        ("ff_alpha", C.c_double),
        ("w_w", C.c_double),
        ("w_w_soft_limits", C.c_double),
        ("w_y", C.c_double),
        ("y_reg_weights", C.c_double * NX_FF),
        ("reg_min", C.c_double),
        ("reg_max", C.c_double),
        ("reg_incfactor", C.c_double),
        ("reg_decfactor", C.c_double),
        ("th_stepdec", C.c_double),
        ("th_stepinc", C.c_double),
        ("th_acceptstep", C.c_double),
        ("th_acceptnegstep", C.c_double),
        ("th_grad", C.c_double),
        ("th_stop", C.c_double),
        ("n_alphas", C.c_int),
        ("boxqp_maxiter", C.c_int),
        ("boxqp_th_acceptstep", C.c_double),
        ("boxqp_th_grad", C.c_double),
        ("boxqp_reg", C.c_double),
        ("w_friction_cone", C.c_double),
        ("mu", C.c_double),
    ]


class BatchIn(C.Structure):
    _fields_ = [
        ("B", C.c_int),
        ("mode", C.c_void_p),
        ("x0", C.c_void_p),
        ("p_ref", C.c_void_p),
        ("v_ref", C.c_void_p),
        ("x_reg_ref", C.c_void_p),
        ("tau_ref", C.c_void_p),
        ("xs_init", C.c_void_p),
        ("us_init", C.c_void_p),
    ]


class BatchOut(C.Structure):
    _fields_ = [
        ("xs", C.c_void_p),
        ("us", C.c_void_p),
        ("K", C.c_void_p),
        ("k", C.c_void_p),
        ("cost", C.c_void_p),
        ("stop", C.c_void_p),
        ("gaps", C.c_void_p),
        ("f_pred", C.c_void_p),
        ("iters", C.c_void_p),
        ("ok", C.c_void_p),
    ]


class NodeOut(C.Structure):
    _fields_ = [
        ("xnext", C.c_void_p),
        ("cost", C.c_void_p),
        ("force", C.c_void_p),
        ("Fx", C.c_void_p),
        ("Fu", C.c_void_p),
        ("Lx", C.c_void_p),
        ("Lu", C.c_void_p),
        ("Lxx", C.c_void_p),
        ("Lxu", C.c_void_p),
        ("Luu", C.c_void_p),
    ]


class TickPolicy(C.Structure):
    _fields_ = [
        ("use_feedback_policy", C.c_int),
        ("ff_use_tau_interpolation", C.c_int),
        ("ff_inverse_actuation_model", C.c_int),
        ("apply_command_filter", C.c_int),
        ("feedback_gain_scale", C.c_double),
        ("dt_mpc", C.c_double),
        ("ff_alpha_ctrl", C.c_double),
        ("max_solver_cost", C.c_double),
        ("max_tau_raw_inf", C.c_double),
        ("fallback_dq_damping", C.c_double),
        ("tau_trust_inf", C.c_double),
        ("tau_smoothing_alpha", C.c_double),
        ("tau_rate_limit", C.c_double * 7),
    ]


class TickIn(C.Structure):
    _fields_ = [
        ("B", C.c_int),
        ("reset", C.c_int),
        ("mode", C.c_void_p),
        ("x_now", C.c_void_p),
        ("p_ref", C.c_void_p),
        ("v_ref", C.c_void_p),
        ("x_reg_ref", C.c_void_p),
        ("tau_bias", C.c_void_p),
        ("tau_ref", C.c_void_p),
    ]


class TickOut(C.Structure):
    _fields_ = [
        ("tau_cmd", C.c_void_p),
        ("tau_raw", C.c_void_p),
        ("cost", C.c_void_p),
        ("fn_pred", C.c_void_p),
        ("iters", C.c_void_p),
        ("ok", C.c_void_p),
        ("unstable", C.c_void_p),
        ("xs", C.c_void_p),
        ("us", C.c_void_p),
    ]


LS_AUTO, LS_QUEUE, LS_SPECULATIVE, LS_GROUPED, LS_TAIL, LS_FILL = 0, 1, 2, 3, 4, 5
LINE_SEARCH_MODES = {"auto": LS_AUTO, "queue": LS_QUEUE, "spec": LS_SPECULATIVE, "grouped": LS_GROUPED, "tail": LS_TAIL, "fill": LS_FILL}


class Options(C.Structure):
    _fields_ = [("line_search", C.c_int), ("sub_batches", C.c_int), ("reserved", C.c_int * 6)]


class Info(C.Structure):
    _fields_ = [("sub_batches", C.c_int), ("sub_batch_capacity", C.c_int), ("line_search", C.c_int), ("trial_slots", C.c_int),
                ("reserved", C.c_int * 4)]


# every symbol include/ffmpc.h declares (tests check the product library exports all of them)
EXPORTED_SYMBOLS = [
    "ffmpc_default_config",
    "ffmpc_default_options",
    "ffmpc_create",
    "ffmpc_create_ex",
    "ffmpc_destroy",
    "ffmpc_solve_batch",
    "ffmpc_solve_batch_host",
    "ffmpc_solve_batch_host_async",
    "ffmpc_wait",
    "ffmpc_node_eval_host",
    "ffmpc_node_eval_static",
    "ffmpc_tick_batch",
    "ffmpc_tick_batch_host",
    "ffmpc_last_timing",
    "ffmpc_set_profiling",
    "ffmpc_get_profile",
    "ffmpc_dfma_peak",
    "ffmpc_sync",
    "ffmpc_get_info",
    "ffmpc_last_error",
    "ffmpc_version",
]


def set_array(dst, values):
    """Copy a (nested) python/numpy sequence into a ctypes array field."""
    import numpy as np

    flat = np.asarray(values, dtype=float).reshape(-1)
    buf = (C.c_double * flat.size).from_buffer(dst)
    if flat.size != len(buf):
        raise ValueError("size mismatch")
    for i, v in enumerate(flat):
        buf[i] = float(v)


def get_array(src, shape=None):
    import numpy as np

    n = C.sizeof(src) // C.sizeof(C.c_double)
    arr = np.frombuffer((C.c_double * n).from_buffer(src), dtype=np.float64).copy()
    return arr.reshape(shape) if shape is not None else arr
