"""the reference's runner module imported UNCHANGED from /root/reference with
`franka_force_feedback_mpc_b200.dropin` installed: its `from src.mpc.crocoddyl_classical import ...` binds the engine-backed
controllers, and its own `_check_pin_mj_alignment` (which drives `pin.forwardKinematics / updateFramePlacements / data.oMf`)
runs on the stand-in `pinocchio` module.  MuJoCo (the plant) is unavailable on this box: a stub module and a kinematic fake of
`FrankaMujocoSim` take its place; the solve is answered by the CPU oracle through the controllers' solver seam.  Runs in a
subprocess so that the substituted `sys.modules` entries cannot leak into other tests."""
import os
import subprocess
import sys
import textwrap

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("FFMPC_REFERENCE", "/root/reference")

SCRIPT = textwrap.dedent('''
    import sys, types
    import numpy as np
    sys.path.insert(0, %(root)r); sys.path.insert(0, %(tests)r); sys.path.insert(0, %(ref)r)
    from franka_force_feedback_mpc_b200 import controllers, dropin, panda, synthetic

    mj = types.ModuleType("mujoco"); mj.viewer = types.ModuleType("mujoco.viewer")
    mj.mj_forward = lambda model, data: data.forward()
    sys.modules["mujoco"], sys.modules["mujoco.viewer"] = mj, mj.viewer

    holder = {}
    class Solver:                                  # the controllers' solver seam, answered by the CPU oracle
        def solve(self, inp, max_iters):
            from oracle import oracle as orc
            cfg_abi, model_abi = holder["ctl"]._abi()
            return orc.solve_batch(cfg_abi, model_abi, inp, max_iters, False, nthreads=1)

    done = dropin.install(solver=Solver())
    assert "src.mpc.crocoddyl_classical" in done and "src.mpc.crocoddyl_force_feedback" in done
    import src.run.run_classical as rc            # the reference's file, unmodified
    import src.run.run_force_feedback as rf
    assert issubclass(rc.ClassicalCrocoddylMPC, controllers.ClassicalCrocoddylMPC)
    assert issubclass(rf.ForceFeedbackCrocoddylMPC, controllers.ForceFeedbackCrocoddylMPC)
    assert rc.ClassicalMPCConfig().horizon > 0 and rc.pin is sys.modules["pinocchio"]

    model = panda.default_model()
    R = panda.R_MJ_FROM_PIN
    class Data:                                    # the slices of mujoco.MjData the runner touches
        def __init__(self, q):
            self.qpos, self.qvel, self.qacc = np.array(q, float), np.zeros(7), np.zeros(7)
            self.site_xpos, self.site_xmat = np.zeros((1, 3)), np.zeros((1, 9))
            self.forward()
        def forward(self):
            p, Rp = model.fk(self.qpos)
            self.site_xpos[0], self.site_xmat[0] = R @ p, (R @ Rp).reshape(9)
    class Sim:                                     # kinematic fake of FrankaMujocoSim
        dt = 0.005
        def __init__(self):
            self.data = Data(panda.Q_NEUTRAL)
            self.model = types.SimpleNamespace(jnt_range=np.stack([model.q_lower, model.q_upper], axis=1))
            self.qpos_adr, self.jnt_ids, self.ee_site_id = np.arange(7), np.arange(7), 0
        def get_observation(self, with_ee=True, with_jacobian=False):
            g = synthetic.batched_gravity_torque(model, self.data.qpos[None])[0]
            return types.SimpleNamespace(q=self.data.qpos.copy(), dq=self.data.qvel.copy(), tau_bias=g, ee_pos=self.data.site_xpos[0].copy(),
                                         ee_quat=None, f_contact_normal=0.0, tau_cmd=g.copy(), tau_meas_act_filt=g.copy())
    sim = Sim()
    traj = synthetic.benchmark_trajectory(model)
    cfg = rc.ClassicalMPCConfig(horizon=6, max_iters=2)
    mpc = rc.ClassicalCrocoddylMPC(sim=sim, traj_fn=lambda t: traj(t), config=cfg)
    holder["ctl"] = mpc
    stats = rc._check_pin_mj_alignment(sim, mpc, samples=8, seed=0)          # the runner's own function
    assert stats["samples"] == 8 and stats["max_pos_m"] < 1e-12 and stats["max_rot_deg"] < 1e-5, stats
    assert np.array_equal(sim.data.qpos, panda.Q_NEUTRAL)                       # the check restored the plant state
    tau = mpc.compute_control(sim.get_observation(), 1.5)
    assert tau.shape == (7,) and np.isfinite(tau).all() and mpc.last_info["iters"] >= 1
    assert rc._scenario_seed("actuation_uncertainty") == 15
    print("DROPIN-OK", stats["max_pos_m"], mpc.last_info["cost"])
''')


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src", "run")), reason="reference checkout not mounted")
def test_reference_runner_module_runs_on_the_dropin(orc):
    code = SCRIPT % {"root": ROOT, "tests": HERE, "ref": REF}
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0 and "DROPIN-OK" in r.stdout, r.stdout + r.stderr
