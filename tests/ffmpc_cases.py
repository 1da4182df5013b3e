"""Synthetic problem factories shared by the tests (importable as a plain module)."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def make_case(variant_ff: bool, nc: int, B: int, N: int, seed: int, contact_fraction: float, scenario: str = "flat", **over):
    """(cfg_abi, model_abi, BatchInputs) of a synthetic batch with the runner benchmark presets."""
    from franka_force_feedback_mpc_b200 import config, panda, synthetic
    from franka_force_feedback_mpc_b200.batch import BatchInputs

    m = panda.default_model()
    kw = dict(horizon=N, contact_model="point3d" if nc == 3 else "normal_1d")
    kw.update(over)
    cfgd = config.benchmark_ff_config(**kw) if variant_ff else config.benchmark_classical_config(**kw)
    cfg = config.to_abi(cfgd, m, panda.Q_NEUTRAL)
    s = synthetic.sample_instances(B, N, variant_ff, seed=seed, contact_fraction=contact_fraction, scenario=scenario)
    return cfg, m.to_abi(), BatchInputs.from_samples(s, 21 if variant_ff else 14, N)


def rel_err(a, b) -> float:
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
