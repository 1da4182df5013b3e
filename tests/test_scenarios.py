"""Scenario / uncertainty harness against fixtures produced by running the reference's own
`ScenarioUncertaintyInjector` and `RunLogger` (tools/make_golden.py::gen_scen -> tests/golden/scen_golden.npz).
Bit-exact: the injector is integer/stream work (same generator, same draw order), the log files are byte formats."""
from __future__ import annotations

import json
import os

import numpy as np
import pytest

from franka_force_feedback_mpc_b200 import scenarios as sc

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "scen_golden.npz"), allow_pickle=False)


def _cfg(row):
    return sc.UncertaintyProfileConfig(a_min=row[0], a_max=row[1], b_min=row[2], b_max=row[3], sigma_q=row[4], sigma_dq=row[5],
                                       sigma_tau=row[6], delta_obs_cycles=int(row[7]), delta_cmd_s=row[8], seed=int(row[9]))


NAMES = ("q", "dq", "tau_meas", "tau_meas_filt", "tau_meas_act", "tau_meas_act_filt", "tau_bias")


def _run(inj, q, dq, bias, cmd, rows):
    """Feed the golden input stream to every loop (row b of the batch = golden case rows[b])."""
    B, T = len(rows), q.shape[0]
    got = {n: [] for n in NAMES + ("plant",)}
    for k in range(T):
        rep = lambda a: np.repeat(a[k][None], B, axis=0)
        d = inj.observation_for_controller(q=rep(q), dq=rep(dq), tau_bias=rep(bias))
        for n in NAMES:
            got[n].append(d[n])
        got["plant"].append(inj.command_for_plant(rep(cmd)))
    return {n: np.stack(v) for n, v in got.items()}   # [T, B, 7]


def test_each_injector_case_alone_is_bit_identical(gold):
    for i in range(len(gold["case_dt"])):
        inj = sc.BatchedUncertaintyInjector(float(gold["case_dt"][i]), 7, [_cfg(gold["case_cfg"][i])], float(gold["case_alpha"][i]))
        got = _run(inj, gold["in_q"], gold["in_dq"], gold["in_bias"], gold["in_cmd"], [i])
        for n, v in got.items():
            assert np.array_equal(v[:, 0], gold[f"inj{i}_{n}"]), (i, n)
        assert inj.meta(0) == json.loads(str(gold[f"inj{i}_meta"]))


def test_loops_with_different_delays_share_one_batch(gold):
    """Per-loop delays inside one batch: golden case 0 in rows 0 and 2 with a differently seeded, longer-delayed loop
    between them.  Rows 0 and 2 must stay identical to the single-loop golden; row 1 must see the observation from
    four ticks back and a zero command until three commands are queued."""
    c0 = _cfg(gold["case_cfg"][0])
    other = sc.UncertaintyProfileConfig(seed=1234, delta_obs_cycles=20, delta_cmd_s=0.015)
    inj = sc.BatchedUncertaintyInjector(0.005, 7, [c0, other, c0], 0.2)
    assert inj.obs_delay_steps.tolist() == [0, 4, 0] and inj.cmd_delay_steps.tolist() == [0, 3, 0]
    got = _run(inj, gold["in_q"], gold["in_dq"], gold["in_bias"], gold["in_cmd"], [0, 1, 2])
    for n, v in got.items():
        assert np.array_equal(v[:, 0], gold[f"inj0_{n}"]) and np.array_equal(v[:, 2], gold[f"inj0_{n}"]), n
    # the delayed loop sees observation max(k-4, 0) (before noise: tau_bias passes through untouched) and a zero
    # command until three commands are queued
    assert np.array_equal(got["tau_bias"][:, 1], gold["in_bias"][np.maximum(np.arange(14) - 4, 0)])
    quiet = got["plant"][:3, 1] - inj.b[1]
    assert np.abs(quiet).max() < 6 * 0.05            # a*0 + b + N(0, sigma_tau)
    assert np.abs(got["plant"][5, 1] - inj.a[1] * gold["in_cmd"][2] - inj.b[1]).max() < 6 * 0.05


def test_scenario_tables_and_errors():
    assert sc.SCENARIOS == ("flat", "tilted_5", "tilted_10", "tilted_15", "actuation_uncertainty")
    assert [sc.scenario_seed(s) for s in sc.SCENARIOS] == [11, 12, 13, 14, 15] and sc.scenario_seed("tilted") == 16
    assert sc.scenario_seed("nope") == 99
    assert sc.scenario_settings("tilted_10")["tilt_deg"] == 10.0
    assert sc.scenario_settings("actuation_uncertainty")["torque_scale"].tolist() == [0.90, 1.08, 0.92, 1.05, 0.88, 1.10, 0.86]
    with pytest.raises(ValueError):
        sc.scenario_settings("nope")
    assert sc.config_for_scenario("flat") is None and sc.config_for_scenario(" Actuation_Uncertainty ", 15).seed == 15
    assert sc.injector_for_scenarios(["flat", "tilted_5"], 0.005) is None
    with pytest.raises(ValueError):
        sc.injector_for_scenarios(["flat", "actuation_uncertainty"], 0.005)
    inj = sc.injector_for_scenarios(["actuation_uncertainty"] * 3, 0.005, seeds=[15, 16, 17])
    assert inj.B == 3 and len({float(a) for a in inj.a}) == 3
    tau = np.ones((2, 7))
    assert np.array_equal(sc.plant_command(tau, ["flat", "actuation_uncertainty"], None)[1], sc.scenario_settings("actuation_uncertainty")["torque_scale"])
    with pytest.raises(ValueError):
        inj.observation_for_controller(q=np.zeros((2, 7)), dq=np.zeros((2, 7)))
    with pytest.raises(ValueError):
        inj.observation_for_controller(q=np.zeros((3, 7)))


def test_injector_statistics_over_many_seeds():
    """Config-4 statement of SURVEY.md 8d: a ~ U(0.95, 1.05), b ~ U(-0.1, 0.1), obs noise 5e-4 / 2e-3, torque noise 0.05."""
    B = 4000
    inj = sc.injector_for_scenarios(["actuation_uncertainty"] * B, 0.005, seeds=range(B))
    assert 0.95 <= inj.a.min() and inj.a.max() <= 1.05 and abs(inj.a.mean() - 1.0) < 3e-3
    assert -0.1 <= inj.b.min() and inj.b.max() <= 0.1 and abs(inj.b.mean()) < 6e-3
    z = np.zeros((B, 7))
    d = inj.observation_for_controller(q=z, dq=z)
    assert abs(d["q"].std() / 5e-4 - 1) < 0.03 and abs(d["dq"].std() / 2e-3 - 1) < 0.03
    assert abs((d["tau_meas"] - inj.b[:, None]).std() / 0.05 - 1) < 0.03
    assert np.array_equal(d["tau_meas_filt"], 0.2 * d["tau_meas"])


def test_metrics_match_the_runner_formulas():
    rng = np.random.default_rng(0)
    B, T = 3, 400
    t = 0.005 * (1 + np.arange(T))
    et, e3, fn = np.abs(rng.normal(0, 0.01, (B, T))), np.abs(rng.normal(0, 0.02, (B, T))), rng.normal(20, 5, (B, T))
    contact = (fn > 0.5).astype(float)
    contact[1, 300:] = 0.0
    m = sc.closed_loop_metrics(t, et, e3, fn, contact, fn_des=22.0, t_contact_phase=0.8)
    for b in range(B):   # the runner's own expressions (run_classical.py:513-534) on one loop
        mask = t >= 0.8
        assert m["rms_tangential_error"][b] == float(np.sqrt(np.mean(et[b] ** 2)))
        assert m["rms_3d_error"][b] == float(np.sqrt(np.mean(e3[b] ** 2)))
        assert m["rms_tangential_error_contact_phase"][b] == float(np.sqrt(np.mean(et[b][mask] ** 2)))
        assert m["avg_abs_position_err"][b] == float(np.mean(np.abs(et[b])))
        assert m["avg_abs_force_err"][b] == float(np.mean(np.abs(fn[b] - 22.0)))
        assert m["max_fn"][b] == float(np.max(fn[b]))
        assert m["contact_loss_pct"][b] == float((1.0 - np.mean(contact[b])) * 100.0)
        assert m["contact_loss_contact_phase_pct"][b] == float((1.0 - np.mean(contact[b][mask])) * 100.0)
        assert m["fn_mean_contact_phase"][b] == float(np.mean(fn[b][mask]))
    late = sc.closed_loop_metrics(t, et, e3, fn, contact, 22.0, t_contact_phase=99.0)
    assert np.isnan(late["fn_mean_contact_phase"]).all() and np.isfinite(late["rms_3d_error"]).all()
    one = sc.closed_loop_metrics(t, et[0], e3[0], fn[0], contact[0], 22.0, 0.8)
    assert one["max_fn"].shape == (1,) and one["max_fn"][0] == m["max_fn"][0]


def test_run_log_files_equal_the_reference_logger(gold, tmp_path):
    """Loop 1 of a 3-loop log carries the rows the reference's RunLogger was given: its three files must match byte for
    byte (csv), value for value (npz: keys, dtypes, arrays) and key for key in order (meta)."""
    q, dq, bias, cmd = gold["in_q"], gold["in_dq"], gold["in_bias"], gold["in_cmd"]
    log = sc.BatchRunLog(["classical_tilted_5", "classical_flat", "classical_tilted_10"], tmp_path, stamp="20260101_000000",
                         notes=[{"scenario": "tilted_5"}, {"scenario": "flat", "scene": "assets/scenes/x.xml"}, None])
    off = np.array([1.0, 0.0, -2.0])
    for k in range(5):
        log.log(t=0.005 * (k + 1), ee_pos=q[k, :3][None] + off[:, None], tau_cmd=cmd[k][None] * (1 + off[:, None]),
                err_tan=abs(dq[k, 0]) + off, fn_meas=bias[k, 0] + off, fn_pred=(np.nan if k == 1 else bias[k, 1]) + off,
                contact=np.full(3, k % 2, dtype=np.int64), solver_iters=np.full(3, 3 + k, dtype=np.int32),
                big=(np.arange(12.0) + k)[None] + off[:, None])
    log.set_meta(total_time=0.025, dt=0.005)
    log.set_meta_rows(scenario_label=["Tilted table (5deg)", "Flat table", "Tilted table (10deg)"])
    log.set_meta(benchmark_mode=True, uncertainty_profile=None, torque_scale=np.ones(7))
    log.set_meta_rows(rms_tangential_error=np.array([1.0, np.sqrt(0.5), 2.0]))
    log.set_meta(cfg_summary={"horizon": 36, "w": [1.0, 2.0]})
    log.save()
    d = log.run_dirs[1]
    assert d.name == "20260101_000000_classical_flat" and d.parent.name == "logs"
    assert open(d / "data.csv", newline="").read() == str(gold["log_csv"])
    want_meta = json.loads(str(gold["log_meta"]))
    got_meta = json.loads(open(d / "meta.json").read())
    want_meta["timestamp"] = got_meta["timestamp"]
    assert got_meta == want_meta and list(got_meta) == list(want_meta)
    assert open(d / "meta.json").read() == json.dumps(want_meta, indent=2)
    z = np.load(d / "data.npz")
    assert sorted(z.files) == gold["log_npz_keys"].tolist()
    for k in z.files:
        assert z[k].dtype == gold[f"log_npz_{k}"].dtype and np.array_equal(z[k], gold[f"log_npz_{k}"], equal_nan=True), k
    assert np.array_equal(np.load(log.run_dirs[2] / "data.npz")["fn_meas"], gold["log_npz_fn_meas"] - 2.0)
    with pytest.raises(FileExistsError):
        sc.BatchRunLog(["classical_flat"], tmp_path, stamp="20260101_000000")
    with pytest.raises(ValueError):
        sc.BatchRunLog(["a", "a"], tmp_path)


def test_batched_latch_and_torque_state_equal_the_controller_mirror():
    """B loops through BatchedSurfaceLatch / torque_state vs B copies of the controller mirror's scalar methods (which
    tests/test_controllers.py pins to the reference source)."""
    import types

    from franka_force_feedback_mpc_b200 import config, controllers

    cfg = config.benchmark_ff_config(horizon=6, phase_source="force_latch", contact_release_steps=3)
    rng = np.random.default_rng(5)
    B, T = 16, 60
    ctls = []
    for _ in range(B):
        c = object.__new__(controllers.ForceFeedbackCrocoddylMPC)
        c.cfg, c._surface_latched, c._contact_loss_count = cfg, False, 0
        ctls.append(c)
    latch = sc.BatchedSurfaceLatch(B, cfg)
    for k in range(T):
        fn = np.where(rng.random(B) < 0.5, 0.0, rng.uniform(0, 3, B))
        ee_z = cfg.z_contact + rng.uniform(-0.01, 0.03, B)
        hint = rng.random(B) < 0.3
        want = [c._detect_surface(types.SimpleNamespace(f_contact_normal=fn[b], ee_pos=np.array([0, 0, ee_z[b]])), 0.0, bool(hint[b])) for b, c in enumerate(ctls)]
        assert latch.update(fn, ee_z, hint).tolist() == want, k
    obs = {"q": np.zeros((B, 7)), "tau_meas_act_filt": rng.normal(0, 1, (B, 7)), "tau_meas_act": rng.normal(0, 1, (B, 7)), "tau_cmd": rng.normal(0, 1, (B, 7))}
    obs["tau_meas_act_filt"][::3, 2] = np.nan
    obs["tau_meas_act"][::6, 0] = np.inf
    got = sc.torque_state(obs, cfg)
    for b, c in enumerate(ctls):
        o = types.SimpleNamespace(**{k: v[b] for k, v in obs.items()})
        assert np.array_equal(got[b], c._tau_state_from_obs(o))


class KinematicPlants:
    """B kinematic stand-ins for the MuJoCo plant (unavailable on this box): joints drift a little each step, the
    applied torque becomes the measured one, the 'contact force' is a spring on table penetration."""

    dt = 0.005

    def __init__(self, model, q0):
        from franka_force_feedback_mpc_b200 import panda, synthetic

        self.m, self.syn, self.R = model, synthetic, panda.R_MJ_FROM_PIN
        self.q, self.dq = np.array(q0, float), np.zeros_like(q0)
        self.tau = synthetic.batched_gravity_torque(model, self.q)
        self.k = 0

    def observe(self):
        p = np.stack([self.m.fk(q)[0] for q in self.q]) @ self.R
        return dict(q=self.q.copy(), dq=self.dq.copy(), tau_bias=self.syn.batched_gravity_torque(self.m, self.q), ee_pos=p,
                    ee_vel=np.zeros_like(p), fn_meas=np.maximum(0.0, 2.0e4 * (0.342 - p[:, 2])), tau_cmd=self.tau.copy(),
                    tau_meas_act_filt=self.tau.copy())

    def step(self, tau_applied):
        self.k += 1
        self.q = self.q + 1e-3 * np.sin(np.arange(7) + self.k)[None]
        self.dq = 0.02 * np.cos(np.arange(7) * 0.7 + self.k)[None] * np.ones_like(self.q)
        self.tau = np.array(tau_applied, float)
        return self.observe()


class TorqueDrivenPlants(KinematicPlants):
    """Stand-in plants that RESPOND to the applied torque, so that the loop is closed (the MuJoCo plant is unavailable on
    this box): decoupled joints with inertia I and viscous damping, qdd = (tau - g(q) - D dq) / I, semi-implicit Euler;
    the 'contact force' is the spring on table penetration of the parent class."""

    I = np.array([2.0, 2.0, 2.0, 2.0, 0.5, 0.5, 0.5])
    D = 8.0 * I

    def step(self, tau_applied):
        self.k += 1
        g = self.syn.batched_gravity_torque(self.m, self.q)
        self.dq = self.dq + self.dt * (np.asarray(tau_applied, float) - g - self.D * self.dq) / self.I
        self.q = self.q + self.dt * self.dq
        self.tau = np.array(tau_applied, float)
        return self.observe()


def _reference_loops(cfg, plant, traj, scen, steps, ff, benchmark_mode, q_nom, t_contact_phase=0.8, contact_threshold=0.5):
    """The loop of `scenarios.run_closed_loops` with the device tick replaced by B INDEPENDENT controller mirrors
    (controllers.py, pinned to the reference's own methods by tests/test_controllers.py), each solving its B = 1 problem
    with the CPU oracle: nothing of the engine on this side."""
    import types

    from franka_force_feedback_mpc_b200 import controllers
    from test_controllers import FakeSim
    from test_tick import OracleB1Solver

    B = len(scen)
    inj = sc.injector_for_scenarios(scen, plant.dt, None, 0.2) if benchmark_mode else None
    ctls = []
    for b in range(B):
        sim = FakeSim(plant.m, np.array(q_nom, float))   # nominal pose / frame offsets of the controller constructor
        sim.dt = plant.dt
        ctl = getattr(controllers, "ForceFeedbackCrocoddylMPC" if ff else "ClassicalCrocoddylMPC")(sim, lambda t: traj(t), config=cfg)
        ctl._solver = OracleB1Solver(*ctl._abi())
        ctls.append(ctl)
    obs = plant.observe()
    for b, ctl in enumerate(ctls):
        # the reference builds each controller on ITS plant: the last-command memory starts at that plant's gravity torque
        # (crocoddyl_classical.py:176-178); the mirrors above were built on the nominal pose
        ctl._tau_prev = np.array(obs["tau_bias"][b], float)
    cols = {k: [] for k in ("t", "err_tan", "err_3d", "fn_meas", "contact", "tau_cmd", "solver_cost", "solver_iters")}
    t = 0.0
    for k in range(steps):
        seen = inj.observation_for_controller(**{n: v for n, v in obs.items() if isinstance(v, np.ndarray) and v.ndim >= 1}) if inj else obs
        taus, costs, its = [], [], []
        for b, ctl in enumerate(ctls):
            o = types.SimpleNamespace(**{n: np.asarray(v)[b] for n, v in seen.items()})
            o.f_contact_normal = float(np.asarray(seen["fn_meas"])[b])
            o.ee_quat = np.array([0.0, 1.0, 0.0, 0.0])
            taus.append(ctl.compute_control(o, t))
            costs.append(ctl.last_info["cost"])
            its.append(ctl.last_info["iters"])
        tau_cmd = np.stack(taus)
        obs = plant.step(sc.plant_command(tau_cmd, scen, inj))
        t += plant.dt
        p_next, _, _ = traj(np.full(B, t))
        err = obs["ee_pos"] - p_next
        row = dict(t=np.full(B, t), err_tan=np.linalg.norm(err[:, :2], axis=1), err_3d=np.linalg.norm(err, axis=1), fn_meas=np.asarray(obs["fn_meas"], float),
                   contact=(np.asarray(obs["fn_meas"]) > contact_threshold).astype(float), tau_cmd=tau_cmd, solver_cost=np.array(costs), solver_iters=np.array(its))
        for n, v in row.items():
            cols[n].append(np.asarray(v, float))
    hist = {n: np.stack(v, axis=1) for n, v in cols.items()}
    return dict(hist, **sc.closed_loop_metrics(hist["t"], hist["err_tan"], hist["err_3d"], hist["fn_meas"], hist["contact"], float(cfg.fn_des), t_contact_phase))


METRIC_KEYS = ("rms_tangential_error", "rms_3d_error", "mean_abs_force_error", "max_normal_force", "contact_loss_pct")


def _closed_loop_parity(lib, ff, scen, benchmark_mode, steps, N, iters, t_contact_phase):
    """(engine sweep, oracle-mirror sweep) of the same closed loops on torque-driven stand-in plants."""
    from franka_force_feedback_mpc_b200 import config, engine, panda, synthetic

    m = panda.default_model()
    cfg = (config.benchmark_ff_config if ff else config.benchmark_classical_config)(horizon=N, max_iters=iters)
    traj = synthetic.benchmark_trajectory(m)
    B = len(scen)
    q0 = panda.Q_NEUTRAL[None] + np.random.default_rng(1).normal(0, 0.02, (B, 7))
    from franka_force_feedback_mpc_b200 import controllers
    from test_controllers import FakeSim

    tmpl = getattr(controllers, "ForceFeedbackCrocoddylMPC" if ff else "ClassicalCrocoddylMPC")(FakeSim(m, panda.Q_NEUTRAL), lambda t: traj(t), config=cfg,
                                                                                                 solver=object())
    eng = engine.FFMPCEngine(*tmpl._abi(), device=0, max_batch=B, _lib=lib)   # the very config the mirrors lower to the ABI
    got = sc.run_closed_loops(eng, cfg, TorqueDrivenPlants(m, q0), traj, scen, steps, ff=ff, q_nom=panda.Q_NEUTRAL, benchmark_mode=benchmark_mode,
                              t_contact_phase=t_contact_phase)
    eng.close()
    want = _reference_loops(cfg, TorqueDrivenPlants(m, q0), traj, scen, steps, ff, benchmark_mode, panda.Q_NEUTRAL, t_contact_phase=t_contact_phase)
    return got, want


def _assert_closed_loop_parity(got, want, label):
    assert np.array_equal(got["solver_iters"], want["solver_iters"]), label
    assert np.abs(got["tau_cmd"] - want["tau_cmd"]).max() < 1e-5, (label, "tau_cmd", float(np.abs(got["tau_cmd"] - want["tau_cmd"]).max()))
    for k in ("err_tan", "err_3d", "fn_meas"):
        assert np.abs(got[k] - want[k]).max() <= 1e-7 * max(1.0, np.abs(want[k]).max()), (label, k)
    assert np.array_equal(got["contact"], want["contact"]), label
    for k in sc_metric_keys(got):
        a, b = np.asarray(got[k], float), np.asarray(want[k], float)
        assert np.array_equal(np.isnan(a), np.isnan(b)), (label, k)
        ok = ~np.isnan(b)
        assert np.all(np.abs(a[ok] - b[ok]) <= 1e-6 * np.maximum(1.0, np.abs(b[ok]))), (label, k, a, b)


def sc_metric_keys(res):
    """the summary-metric entries of a sweep result (run_classical.py:513-534): everything that is one number per loop"""
    return [k for k, v in res.items() if np.asarray(v).ndim == 1 and k not in ("t",)]


def _sweep(lib, tmp_path, ff, scen, benchmark_mode, steps=4):
    from franka_force_feedback_mpc_b200 import config, engine, panda, synthetic

    m = panda.default_model()
    cfg = (config.benchmark_ff_config if ff else config.benchmark_classical_config)(horizon=6, max_iters=2)
    traj = synthetic.benchmark_trajectory(m)
    B = len(scen)
    plants = KinematicPlants(m, panda.Q_NEUTRAL[None] + np.random.default_rng(1).normal(0, 0.02, (B, 7)))
    eng = engine.FFMPCEngine(config.to_abi(cfg, m, panda.Q_NEUTRAL), m.to_abi(), device=0, max_batch=B, _lib=lib)
    log = sc.BatchRunLog([f"{'ff' if ff else 'classical'}_{s}_{i}" for i, s in enumerate(scen)], tmp_path, notes=[{"scenario": s} for s in scen])
    res = sc.run_closed_loops(eng, cfg, plants, traj, scen, steps, ff=ff, q_nom=panda.Q_NEUTRAL, benchmark_mode=benchmark_mode, log=log)
    eng.close()
    return res, log


def test_emulated_sweep_writes_reference_format_runs(tmp_path):
    import subprocess

    from franka_force_feedback_mpc_b200 import engine

    emu_dir = os.path.join(HERE, "host_emul")
    subprocess.check_call(["make", "-C", emu_dir, "all"], stdout=subprocess.DEVNULL)
    lib = engine.load_library(os.path.join(emu_dir, "libffmpc_emu.so"))
    res, log = _sweep(lib, tmp_path / "a", False, ["flat", "tilted_5", "actuation_uncertainty"], benchmark_mode=False)
    assert res["t"].shape == (3, 4) and np.isfinite(res["rms_3d_error"]).all()
    z = np.load(log.run_dirs[2] / "data.npz")
    scale = sc.scenario_settings("actuation_uncertainty")["torque_scale"]
    assert np.array_equal(z["tau_applied"], z["tau_cmd"] * scale) and z["tau_cmd"].shape == (4, 7)
    assert np.array_equal(np.load(log.run_dirs[0] / "data.npz")["tau_applied"], np.load(log.run_dirs[0] / "data.npz")["tau_cmd"])
    meta = json.load(open(log.run_dirs[1] / "meta.json"))
    assert meta["scenario_tilt_deg"] == 5.0 and meta["uncertainty_profile"] is None and meta["rms_3d_error"] == res["rms_3d_error"][1]
    head = open(log.run_dirs[0] / "data.csv").readline().strip().split(",")
    assert head[:4] == ["contact", "ee_pos[0]", "ee_pos[1]", "ee_pos[2]"] and "tau_cmd[6]" in head and head == sorted(head, key=lambda h: h.split("[")[0])
    # benchmark mode: injector streams with the scenario seed -> two loops with the same seed and plant are identical,
    # a third with another seed is not
    res2, log2 = _sweep(lib, tmp_path / "b", True, ["actuation_uncertainty"] * 2, benchmark_mode=True)
    m0, m1 = (json.load(open(d / "meta.json")) for d in log2.run_dirs)
    assert m0["uncertainty_profile"]["seed"] == 15 and m0["uncertainty_profile"]["a"] == m1["uncertainty_profile"]["a"]
    assert np.isfinite(res2["err_3d"]).all()


@pytest.mark.gpu
def test_gpu_sweep_of_all_scenarios(built, tmp_path):
    """The five scenarios x 4 seeds as one batch of loops on the real engine (classical and FF), plain mode: finite
    metrics, one run directory per loop, the gain-mismatch scenario scales the applied torque."""
    from franka_force_feedback_mpc_b200 import engine

    lib = engine.load_library()
    scen = [s for s in sc.SCENARIOS for _ in range(4)]
    for ff in (False, True):
        res, log = _sweep(lib, tmp_path / ("ff" if ff else "cl"), ff, scen, benchmark_mode=False, steps=6)
        assert res["t"].shape == (20, 6) and np.isfinite(res["err_3d"]).all() and np.isfinite(res["rms_tangential_error"]).all()
        assert all((d / "data.csv").exists() and (d / "meta.json").exists() for d in log.run_dirs)
        z = np.load(log.run_dirs[17] / "data.npz")
        assert np.array_equal(z["tau_applied"], z["tau_cmd"] * sc.scenario_settings("actuation_uncertainty")["torque_scale"])
        assert (z["solver_iters"] >= 1).all() and np.isfinite(z["solver_cost"]).all()
    res, log = _sweep(lib, tmp_path / "bm", True, ["actuation_uncertainty"] * 8, benchmark_mode=True, steps=6)
    assert np.isfinite(res["err_3d"]).all()


@pytest.mark.parametrize("ff,benchmark_mode", [(False, False), (True, True)])
def test_emulated_closed_loop_matches_oracle_backed_mirrors(ff, benchmark_mode):
    """north_star: closed-loop metrics (tangential / 3D tracking error, force error, max normal force, contact-loss %)
    must match the reference's.  The MuJoCo plant is unavailable on this box; what CAN be checked is that the batched
    closed loop around the device tick reproduces, tick by tick and in the end-of-run metrics, the same loops run by B
    independent controller mirrors whose solves are answered by the CPU oracle, on stand-in plants that respond to the
    commanded torque.  Emulator build of the kernels, small sizes; the GPU version runs all five scenarios."""
    import subprocess

    from franka_force_feedback_mpc_b200 import engine

    emu_dir = os.path.join(HERE, "host_emul")
    subprocess.check_call(["make", "-C", emu_dir, "all"], stdout=subprocess.DEVNULL)
    lib = engine.load_library(os.path.join(emu_dir, "libffmpc_emu.so"))
    scen = ["actuation_uncertainty"] * 2 if benchmark_mode else ["flat", "tilted_10", "actuation_uncertainty"]
    got, want = _closed_loop_parity(lib, ff, scen, benchmark_mode, steps=5, N=6, iters=2, t_contact_phase=0.0)
    _assert_closed_loop_parity(got, want, (ff, benchmark_mode))


@pytest.mark.gpu
@pytest.mark.parametrize("ff", [False, True])
def test_gpu_closed_loop_metrics_match_oracle_backed_mirrors(built, ff):
    """All five scenarios x 2 loops on the real engine against the oracle-backed mirrors (see the emulator test above):
    per-tick commands, error histories and the runner's end-of-run metrics."""
    from franka_force_feedback_mpc_b200 import engine

    lib = engine.load_library()
    scen = [s for s in sc.SCENARIOS for _ in range(2)]
    got, want = _closed_loop_parity(lib, ff, scen, False, steps=12, N=30, iters=5, t_contact_phase=0.0)
    _assert_closed_loop_parity(got, want, (ff, "plain"))
    got, want = _closed_loop_parity(lib, ff, ["actuation_uncertainty"] * 4, True, steps=12, N=30, iters=5, t_contact_phase=0.0)
    _assert_closed_loop_parity(got, want, (ff, "benchmark"))
