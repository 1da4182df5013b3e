"""Device-side receding-horizon tick (`ffmpc_tick_batch_host`): B controllers ticked at once must produce
what B independent copies of the controller mirror (`controllers.py`, itself pinned to the reference's own methods by
tests/test_controllers.py) produce when each drives its own B = 1 solve: same warm-start shift, same policy
(classical feedback with the reference's sign; FF Eq. 14-18), same unstable guard and command filter, plan dropped on a
mode switch.  CPU: the kernel sources under the fiber emulator (test-only build); GPU: the real engine."""
from __future__ import annotations

import os
import subprocess
import types

import numpy as np
import pytest

from test_controllers import FakeSim

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_DIR = os.path.join(HERE, "host_emul")


class EngineSolver:
    """B = 1 engine behind the controller mirror (same library as the batched engine under test)."""

    def __init__(self, cfg_abi, model_abi, lib):
        from franka_force_feedback_mpc_b200 import engine

        self.eng = engine.FFMPCEngine(cfg_abi, model_abi, device=0, max_batch=1, _lib=lib)

    def solve(self, inp, max_iters):
        return self.eng.solve_host(inp, max_iters, fixed_iters=False, want=("xs", "us", "K", "cost", "iters", "ok", "f_pred"))


class OracleB1Solver:
    """The CPU oracle behind the controller mirror: a reference for the tick that shares NOTHING with the engine under test."""

    def __init__(self, cfg_abi, model_abi, lib=None):
        self.cfg, self.model = cfg_abi, model_abi

    def solve(self, inp, max_iters):
        from oracle import oracle as orc

        return orc.solve_batch(self.cfg, self.model, inp, max_iters, False, nthreads=1)


def _drive(lib, ff: bool, filt: bool, ticks: int = 5, N: int = 8, iters: int = 3, mirror=None, tol_cost=1e-9, tol_fn=1e-8):
    """Returns the largest |tau_cmd(batched tick) - tau_cmd(independent mirrors)| over the run.  `mirror`: solver class behind
    the B independent controller mirrors (default: a B = 1 engine of the same library; OracleB1Solver: the CPU oracle)."""
    Solver = mirror or EngineSolver
    from franka_force_feedback_mpc_b200 import config, controllers, engine, panda, synthetic

    preset = config.benchmark_ff_config if ff else config.benchmark_classical_config
    cfg = preset(horizon=N, max_iters=iters, apply_command_filter=filt)
    model = panda.default_model()
    traj = synthetic.benchmark_trajectory(model)
    t_start = [1.5, 0.10, 0.7925, 2.3]   # contact, free, free -> contact switch after two ticks (t_c = 0.8), contact
    B = len(t_start)
    rng = np.random.default_rng(3)
    ctls, sims = [], []
    for b in range(B):
        sim = FakeSim(model, panda.Q_NEUTRAL + rng.normal(0, 0.02, 7))
        ctl = getattr(controllers, "ForceFeedbackCrocoddylMPC" if ff else "ClassicalCrocoddylMPC")(sim, lambda t: traj(t), config=cfg)
        ctl._solver = Solver(*ctl._abi(), lib)
        ctls.append(ctl)
        sims.append(sim)
    # all controllers share q_nom-dependent config only through _abi(); use controller 0's for the batched engine
    # (q_nom differs per controller -> q soft-limit reference differs: build every controller on the same nominal pose)
    cfg_abi, model_abi = ctls[0]._abi()
    for c in ctls[1:]:
        c._abi_cache = (cfg_abi, model_abi)
        c.q_nom = ctls[0].q_nom.copy()
        c._solver = Solver(cfg_abi, model_abi, lib)
    ctls[0]._solver = Solver(cfg_abi, model_abi, lib)
    eng = engine.FFMPCEngine(cfg_abi, model_abi, device=0, max_batch=B, _lib=lib)
    pol = config.tick_policy(cfg, dt_mpc=sims[0].dt)
    nx = 21 if ff else 14
    worst = 0.0
    modes_seen = set()
    for k in range(ticks):
        obs_all = [s.get_observation() for s in sims]
        ts = [t0 + k * sims[0].dt for t0 in t_start]
        ref_tau, ref_info = [], []
        for c, o, t in zip(ctls, obs_all, ts):
            ref_tau.append(c.compute_control(o, t))
            ref_info.append(dict(c.last_info))
        mode = np.array([int(i["surface_mode"]) for i in ref_info], np.int32)
        modes_seen.update(mode.tolist())
        x_now = np.stack([np.concatenate([o.q, o.dq] + ([c._tau_state_from_obs(o)] if ff else [])) for c, o in zip(ctls, obs_all)])
        refs = [c._references(t) for c, t in zip(ctls, ts)]
        out = eng.tick_host(mode, x_now, np.concatenate([r[0] for r in refs]), np.concatenate([r[1] for r in refs]),
                            np.stack([c._posture_reference(x[:14]) for c, x in zip(ctls, x_now)]),
                            np.stack([o.tau_bias for o in obs_all]), pol, iters, fixed_iters=False, reset=(k == 0))
        assert out["tau_cmd"].shape == (B, 7) and nx == x_now.shape[1]
        worst = max(worst, float(np.abs(out["tau_cmd"] - np.stack(ref_tau)).max()))
        for b in range(B):
            assert out["iters"][b] == ref_info[b]["iters"] and bool(out["ok"][b]) == ref_info[b]["ok"], (k, b)
            assert bool(out["unstable"][b]) == ref_info[b]["unstable"]
            assert abs(out["cost"][b] - ref_info[b]["cost"]) <= tol_cost * max(1.0, abs(ref_info[b]["cost"]))
            fr = ref_info[b]["fn_pred_raw"] if ff else ref_info[b]["fn_pred"]
            assert (np.isnan(out["fn_pred"][b]) and np.isnan(fr)) or abs(out["fn_pred"][b] - fr) <= tol_fn * max(1.0, abs(fr))
        # plant stand-in: small kinematic drift + the commanded torque becomes the next measured torque
        for b, s in enumerate(sims):
            s.q = s.q + 2e-3 * np.sin(np.arange(7) + k + b)
            s.dq = 0.05 * np.cos(np.arange(7) * 0.7 + k)
            s.tau_cmd = ref_tau[b].copy()
    eng.close()
    assert modes_seen == {0, 1}
    return worst


@pytest.fixture(scope="module")
def emu():
    subprocess.check_call(["make", "-C", EMU_DIR, "all"], stdout=subprocess.DEVNULL)
    from franka_force_feedback_mpc_b200 import engine

    lib = engine.load_library(os.path.join(EMU_DIR, "libffmpc_emu.so"))
    assert b"EMULATION" in lib.ffmpc_version()
    return lib


@pytest.mark.parametrize("ff,filt", [(False, False), (True, True)])
def test_emulated_batched_tick_equals_independent_controllers(emu, ff, filt):
    assert _drive(emu, ff, filt, ticks=4, N=6, iters=2) < 1e-9


@pytest.mark.parametrize("ff,filt", [(False, True), (True, False)])
def test_emulated_batched_tick_equals_oracle_backed_controllers(emu, ff, filt):
    """Same run, the mirrors solving with the CPU ORACLE: nothing of the engine on the reference side (a solve regression
    cannot cancel).  Torques agree to 1e-6 Nm (1e-8 relative solver parity through the feedback gains)."""
    assert _drive(emu, ff, filt, ticks=4, N=6, iters=2, mirror=OracleB1Solver, tol_cost=1e-8, tol_fn=1e-7) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("ff,filt", [(False, True), (True, False), (True, True)])
def test_gpu_batched_tick_equals_independent_controllers(built, ff, filt):
    from franka_force_feedback_mpc_b200 import engine

    assert _drive(engine.load_library(), ff, filt, ticks=6, N=30, iters=5) < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("ff,filt", [(False, True), (True, False), (True, True)])
def test_gpu_batched_tick_equals_oracle_backed_controllers(built, ff, filt):
    """SURVEY.md 8f2 against an independent reference: the device tick vs B controller mirrors whose solves are answered by
    the CPU oracle."""
    from franka_force_feedback_mpc_b200 import engine

    assert _drive(engine.load_library(), ff, filt, ticks=6, N=30, iters=5, mirror=OracleB1Solver, tol_cost=1e-8, tol_fn=1e-7) < 1e-6


def test_tick_argument_errors(emu):
    from franka_force_feedback_mpc_b200 import config, engine, panda

    cfg = config.benchmark_classical_config(horizon=4)
    m = panda.default_model()
    eng = engine.FFMPCEngine(config.to_abi(cfg, m, panda.Q_NEUTRAL), m.to_abi(), device=0, max_batch=2, _lib=emu)
    pol = config.tick_policy(cfg)
    z = lambda *s: np.zeros(s)
    with pytest.raises(engine.FFMPCError):   # batch larger than the handle
        eng.tick_host(np.zeros(3, np.int32), z(3, 14), z(3, 5, 3), z(3, 5, 3), z(3, 14), z(3, 7), pol, 1)
    pol.dt_mpc = 0.0
    with pytest.raises(engine.FFMPCError):
        eng.tick_host(np.zeros(2, np.int32), z(2, 14), z(2, 5, 3), z(2, 5, 3), z(2, 14), z(2, 7), pol, 1)
    eng.close()


def test_tick_forgets_plans_after_a_solve_or_a_batch_size_change(emu):
    """The stored plans ARE the solver workspace (instance b of sub-batch c): a tick after ffmpc_solve_batch_host on the
    handle, or with another batch size, must behave like a tick with `reset` set instead of warm-starting from (and
    applying the feedback gains of) whatever now sits in the workspace."""
    from franka_force_feedback_mpc_b200 import config, engine, panda, synthetic
    from franka_force_feedback_mpc_b200.batch import BatchInputs

    N, B = 5, 3
    cfg = config.benchmark_classical_config(horizon=N, max_iters=2)
    m = panda.default_model()
    cfg_abi, am = config.to_abi(cfg, m, panda.Q_NEUTRAL), m.to_abi()
    pol = config.tick_policy(cfg, dt_mpc=0.005)
    s = synthetic.sample_instances(B, N, False, seed=5, contact_fraction=0.5)
    args = lambda n: (s["mode"][:n], s["x0"][:n], s["p_ref"][:n], s["v_ref"][:n], s["x_reg_ref"][:n], np.zeros((n, 7)))
    fresh = engine.FFMPCEngine(cfg_abi, am, device=0, max_batch=B, _lib=emu)
    want3 = fresh.tick_host(*args(3), pol, 2, reset=True)
    want2 = fresh.tick_host(*args(2), pol, 2, reset=True)
    fresh.close()
    eng = engine.FFMPCEngine(cfg_abi, am, device=0, max_batch=B, _lib=emu)
    eng.tick_host(*args(3), pol, 2, reset=True)
    other = synthetic.sample_instances(B, N, False, seed=99, contact_fraction=0.5)
    eng.solve_host(BatchInputs.from_samples(other, 14, N), 2, True)        # overwrites the workspace the plans live in
    got3 = eng.tick_host(*args(3), pol, 2, reset=False)
    got2 = eng.tick_host(*args(2), pol, 2, reset=False)                    # another batch size than the previous tick
    eng.close()
    for k in ("tau_cmd", "tau_raw", "cost", "iters"):
        assert np.array_equal(got3[k], want3[k]), k
        assert np.array_equal(got2[k], want2[k]), k
