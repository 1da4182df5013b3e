"""Pins the CPU oracle (and, under `-m gpu`, the CUDA engine) to outputs of the REFERENCE's own Python:
the fixtures under tests/golden/ were produced by tools/make_golden.py, which imports
/root/reference/src/... under stub third-party modules and executes the reference source itself.
  * _AugmentedLPFActionModel.calc / calcDiff   crocoddyl_force_feedback.py:149-290
  * make_approach_then_circle + hold wrapper   trajectories.py:8-93, run_classical.py:235-255
  * _ff_alpha_ocp                              crocoddyl_force_feedback.py:493-497
Tolerance: 1e-12 relative (same fp64 arithmetic, different summation order in the block assembly)."""
import os

import numpy as np
import pytest

from ffmpc_cases import make_case, rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KEYS = ("xnext", "cost", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu")
TOL = 1e-12


def _lpf_case(nc):
    g = np.load(os.path.join(GOLD, "lpf_golden.npz"))
    cfg, am, inp = make_case(True, nc, int(g["B"]), int(g["N"]), int(g["seed"]), float(g["contact_fraction"]))
    tag = f"nc{nc}"
    return g, tag, cfg, am, inp


@pytest.mark.parametrize("nc", [1, 3])
def test_oracle_lpf_wrapper_matches_reference_source(orc, nc):
    g, tag, cfg, am, inp = _lpf_case(nc)
    out = orc.node_eval(cfg, am, inp, g[f"{tag}_inst"], g[f"{tag}_knot"], g[f"{tag}_x"], g[f"{tag}_w"])
    for k in KEYS:
        assert rel_err(out[k], g[f"{tag}_{k}"]) < TOL, k
    # the fixture exercises the soft-limit branch and both contact modes
    assert np.count_nonzero(np.diagonal(g[f"{tag}_Luu"], axis1=1, axis2=2) > cfg.w_w + 1e-9) > 0
    assert set(np.unique(inp.mode[g[f"{tag}_inst"]])) == {0, 1}


@pytest.mark.gpu
@pytest.mark.parametrize("nc", [1, 3])
def test_gpu_lpf_wrapper_matches_reference_source(built, nc):
    from franka_force_feedback_mpc_b200 import engine

    g, tag, cfg, am, inp = _lpf_case(nc)
    eng = engine.FFMPCEngine(cfg, am, device=0, max_batch=inp.B)
    out = eng.node_eval(inp, g[f"{tag}_inst"], g[f"{tag}_knot"], g[f"{tag}_x"], g[f"{tag}_w"])
    eng.close()
    for k in KEYS:
        # analytic derivatives on the GPU vs forward-mode AD in the oracle that fed the reference wrapper
        assert rel_err(out[k], g[f"{tag}_{k}"]) < 1e-9, k


def test_trajectory_matches_reference_source():
    from franka_force_feedback_mpc_b200 import synthetic

    g = np.load(os.path.join(GOLD, "traj_golden.npz"))
    traj = synthetic.benchmark_trajectory()
    assert np.allclose(traj.ee_start, g["ee_start"], atol=1e-15)
    p, v, s = traj(g["t"])
    assert np.array_equal(s, g["surface"])
    assert np.abs(p - g["p"]).max() < 1e-14
    assert np.abs(v - g["v"]).max() < 1e-13


def test_ff_alpha_matches_reference_source():
    from franka_force_feedback_mpc_b200 import config

    g = np.load(os.path.join(GOLD, "misc_golden.npz"))
    for fc, dt_ocp, dt, a_ocp, a_ctrl, eps in g["alpha_rows"]:
        c = config.benchmark_ff_config(ff_cutoff_hz=float(fc), dt_ocp=float(dt_ocp), dt=float(dt))
        assert abs(config.ff_alpha_ocp(c) - a_ocp) < 1e-15


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="reference tree not mounted (GPU box)")
def test_fixtures_are_reproducible_from_reference(tmp_path, monkeypatch):
    """Where the reference is mounted, regenerate the fixtures and compare with the committed ones."""
    import importlib.util
    import sys

    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(os.path.dirname(GOLD), "..", "tools", "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    monkeypatch.setattr(mg, "GOLD", str(tmp_path))
    saved = {k: sys.modules.get(k) for k in ("crocoddyl", "pinocchio", "example_robot_data", "mujoco", "mujoco.viewer", "src")}
    try:
        mg.install_stubs()
        sys.path.insert(0, mg.REF)
        mg.gen_traj()
        mg.gen_lpf()
    finally:
        sys.path.remove(mg.REF)
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
        for k in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
            sys.modules.pop(k, None)
    for name in ("traj_golden.npz", "lpf_golden.npz"):
        a, b = np.load(os.path.join(GOLD, name)), np.load(os.path.join(str(tmp_path), name))
        assert set(a.files) == set(b.files)
        for k in a.files:
            assert np.allclose(a[k], b[k], rtol=1e-13, atol=1e-13), (name, k)
