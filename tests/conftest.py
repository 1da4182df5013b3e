"""Shared fixtures.  `-m "not gpu"` runs on a CPU-only box (oracle, host logic, ABI, emulator);
`-m gpu` needs a B200 and calls the CUDA engine through the C ABI."""
from __future__ import annotations

import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))
if HERE not in sys.path:
    sys.path.insert(0, HERE)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def has_gpu() -> bool:
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built():
    """Compile the CUDA library (nvcc cross-compiles without a GPU) and the oracle once per session."""
    import __graft_entry__ as ge

    ge.build_cuda()
    ge.build_oracle()
    return ge


@pytest.fixture(scope="session")
def orc(built):
    from oracle import oracle

    return oracle


@pytest.fixture(scope="session")
def model():
    from franka_force_feedback_mpc_b200 import panda

    return panda.default_model()


from ffmpc_cases import make_case, rel_err  # noqa: E402,F401
