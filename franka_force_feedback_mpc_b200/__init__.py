"""B200-native batched FDDP / Box-FDDP MPC engine for the Franka Panda controllers of
omar-a-aman/Franka-Force-Feedback-MPC (classical torque MPC and force-feedback MPC).

The compute path is hand-written CUDA for sm_100a behind the C ABI of ``include/ffmpc.h``;
this package is the host-side mirror of the reference's controller interface.
"""
from .config import (  # noqa: F401
    ClassicalMPCConfig,
    ForceFeedbackMPCConfig,
    benchmark_classical_config,
    benchmark_ff_config,
)
from .controllers import ClassicalCrocoddylMPC, ForceFeedbackCrocoddylMPC  # noqa: F401
from .panda import PandaModel, default_model  # noqa: F401

__all__ = [
    "ClassicalCrocoddylMPC",
    "ForceFeedbackCrocoddylMPC",
    "ClassicalMPCConfig",
    "ForceFeedbackMPCConfig",
    "benchmark_classical_config",
    "benchmark_ff_config",
    "PandaModel",
    "default_model",
]
