"""GPU parity over the configuration space of the reference's OCP builder (`-m gpu`, through the C ABI): the switches
the benchmark preset leaves in one position -- inner regularisation flags of the force-feedback model, posture / torque
reference modes, the runners' development presets, the dataclass defaults (w_plane_z = w_vz = 0, default friction cone)
and the benchmark horizons N = 36 / 40.  Knot quantities at 1e-10, one and five FDDP iterations on the bar of
tests/parity.py.  The same cases run at a small size on the CPU emulator build in tests/test_emulator.py."""
import pytest

from ffmpc_cases import CONFIG_CASES, run_config_case

pytestmark = pytest.mark.gpu


def _eng(cfg, am, B):
    from franka_force_feedback_mpc_b200 import engine

    return engine.FFMPCEngine(cfg, am, device=0, max_batch=B)


@pytest.mark.parametrize("case", CONFIG_CASES, ids=[c[0] for c in CONFIG_CASES])
@pytest.mark.parametrize("iters", [1, 5])
def test_configuration_matches_oracle(built, orc, case, iters):
    run_config_case(_eng, orc, case, B=32, N=50, iters=iters)
