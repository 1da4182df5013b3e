"""Batched scenario / uncertainty harness around the tick driver.

The reference runs ONE closed loop per process: a scenario name picks a seed, a hidden table tilt and a torque scale
(`src/run/run_classical.py:30-91`), the `actuation_uncertainty` scenario wraps plant and controller in a
`ScenarioUncertaintyInjector` (`src/run/uncertainty_profiles.py:84-161`), every step is logged through `RunLogger`
(`src/utils/logging.py:33-151`) and the run ends with a block of summary metrics (`run_classical.py:513-534`).

Here the same things exist for B closed loops at once, laid out the way `FFMPCEngine.tick_host` wants its inputs
([B, ...] arrays, one row per loop):

  * `BatchedUncertaintyInjector`  -- B independent injector streams.  Row b reproduces, bit for bit, what the
    reference's injector built from `configs[b]` produces (same generator, same draw order, same delay queues), so a
    sweep over seeds is B reference runs, not a statistical look-alike.
  * `closed_loop_metrics`         -- the runner's end-of-run metrics, vectorised over loops.
  * `BatchRunLog`                 -- columnar logger; `save()` writes, per loop, the reference's
    `<stamp>_<run_name>/{data.npz,data.csv,meta.json}` with identical keys, dtypes, column flattening and meta order.

Pinned by tests/test_scenarios.py against fixtures produced by running the reference's own classes
(tools/make_golden.py::gen_scen).
"""
from __future__ import annotations

import csv
import json
import os
import time
from dataclasses import asdict, dataclass, is_dataclass
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence

import numpy as np

SCENARIOS = ("flat", "tilted_5", "tilted_10", "tilted_15", "actuation_uncertainty")   # run_classical.py:26

_SEEDS = {"flat": 11, "tilted_5": 12, "tilted_10": 13, "tilted_15": 14, "actuation_uncertainty": 15, "tilted": 16}
_TILT = {"flat": 0.0, "tilted_5": 5.0, "tilted_10": 10.0, "tilted_15": 15.0, "actuation_uncertainty": 0.0, "tilted": 8.0}
_LABEL = {"flat": "Flat table", "tilted_5": "Tilted table (5deg)", "tilted_10": "Tilted table (10deg)",
          "tilted_15": "Tilted table (15deg)", "actuation_uncertainty": "Actuation gain mismatch",
          "tilted": "Tilted table (8deg)"}
_GAIN_MISMATCH = (0.90, 1.08, 0.92, 1.05, 0.88, 1.10, 0.86)


def scenario_seed(name: str) -> int:
    """`_scenario_seed` (run_classical.py:29-38): unknown names fall back to 99."""
    return int(_SEEDS.get(name, 99))


def scenario_settings(name: str) -> Dict[str, Any]:
    """`_scenario_settings` (run_classical.py:53-91): hidden tilt, per-joint torque scale, label."""
    if name not in _TILT:
        raise ValueError(f"Unknown scenario '{name}'")
    scale = np.array(_GAIN_MISMATCH, dtype=float) if name == "actuation_uncertainty" else np.ones(7, dtype=float)
    return {"tilt_deg": _TILT[name], "torque_scale": scale, "label": _LABEL[name]}


@dataclass
class UncertaintyProfileConfig:
    """Same fields and defaults as the reference's dataclass (uncertainty_profiles.py:12-30)."""

    a_min: float = 0.95
    a_max: float = 1.05
    b_min: float = -0.1
    b_max: float = 0.1
    sigma_q: float = 5.0e-4
    sigma_dq: float = 2.0e-3
    sigma_tau: float = 5.0e-2
    delta_obs_cycles: int = 2
    delta_cmd_s: float = 1.0e-3
    seed: int = 0


def config_for_scenario(scenario: str, seed: int = 0) -> Optional[UncertaintyProfileConfig]:
    """Only `actuation_uncertainty` carries an injected profile (uncertainty_profiles.py:33-52)."""
    if str(scenario).strip().lower() == "actuation_uncertainty":
        return UncertaintyProfileConfig(seed=int(seed))
    return None


class BatchedUncertaintyInjector:
    """B injector streams side by side (uncertainty_profiles.py:84-161).

    Per loop b and tick k the reference does, in this order: delayed observation = the one from
    `max(k - obs_delay_steps, 0)`; `q += N(0, sigma_q)`, `dq += N(0, sigma_dq)`; `tau_hat = a*cmd_delayed + b +
    N(0, sigma_tau)`; EMA of tau_hat; then, after the controller ran, `command_for_plant` queues the nominal command
    and returns another `a*cmd_delayed + b + N(0, sigma_tau)`.  All draws of loop b come from
    `numpy.random.default_rng(configs[b].seed)` (a, b first): one generator per loop is the reference's seeding
    contract, so the draws are a host loop over generators while queues, delays and filters are [B, ...] arrays.
    """

    def __init__(self, dt: float, nu: int, configs: Sequence[UncertaintyProfileConfig], tau_lpf_alpha: float = 0.2):
        self.dt = float(max(dt, 1.0e-9))
        self.nu = int(nu)
        self.cfgs = list(configs)
        B = self.B = len(self.cfgs)
        if B == 0:
            raise ValueError("at least one loop is required")
        self.rngs = [np.random.default_rng(int(c.seed)) for c in self.cfgs]
        self.a = np.empty(B)
        self.b = np.empty(B)
        for i, (c, g) in enumerate(zip(self.cfgs, self.rngs)):
            self.a[i] = float(g.uniform(float(c.a_min), float(c.a_max)))
            self.b[i] = float(g.uniform(float(c.b_min), float(c.b_max)))
        self.sig = np.array([[float(c.sigma_q), float(c.sigma_dq), float(c.sigma_tau)] for c in self.cfgs])
        self.obs_delay_cycles_1khz = np.array([int(max(c.delta_obs_cycles, 0)) for c in self.cfgs])
        self.obs_delay_steps = np.maximum(np.round(self.obs_delay_cycles_1khz * 1.0e-3 / self.dt), 0).astype(int)
        self.cmd_delay_steps = np.maximum(np.round(np.array([float(c.delta_cmd_s) for c in self.cfgs]) / self.dt), 0).astype(int)
        self._k = 0                                             # observations seen
        self._n_cmd = 0                                         # commands queued
        self._obs_ring: Optional[Dict[str, np.ndarray]] = None  # name -> [Do+1, B, ...], slot = tick % (Do+1)
        self._cmd_ring = np.zeros((int(self.cmd_delay_steps.max()) + 1, B, self.nu))
        self._tau_hat_filt = np.zeros((B, self.nu))
        self._alpha = float(np.clip(tau_lpf_alpha, 0.0, 1.0))
        self._rows = np.arange(B)

    def meta(self, b: int = 0) -> dict:
        c = self.cfgs[b]
        return {
            "a": float(self.a[b]), "b": float(self.b[b]),
            "sigma_q": float(c.sigma_q), "sigma_dq": float(c.sigma_dq), "sigma_tau": float(c.sigma_tau),
            "delta_obs_cycles_1khz": int(self.obs_delay_cycles_1khz[b]), "delta_obs_steps": int(self.obs_delay_steps[b]),
            "delta_cmd_steps": int(self.cmd_delay_steps[b]), "delta_cmd_s": float(c.delta_cmd_s), "seed": int(c.seed),
        }

    # -- queues ----------------------------------------------------------------------------------------------------
    def _delayed_command(self) -> np.ndarray:
        # deque(maxlen=D+1) seeded with zeros: its head after n appends is command n-1-D (zero while negative)
        src = self._n_cmd - 1 - self.cmd_delay_steps
        L = self._cmd_ring.shape[0]
        out = self._cmd_ring[np.maximum(src, 0) % L, self._rows]
        return np.where((src >= 0)[:, None], out, 0.0)

    def _tau_hat(self, z_tau: np.ndarray) -> np.ndarray:
        return self.a[:, None] * self._delayed_command() + self.b[:, None] + z_tau

    def _draw(self, per_loop: int) -> np.ndarray:
        """[B, per_loop, nu] standard normals, loop b from its own generator, in the reference's draw order."""
        z = np.empty((self.B, per_loop, self.nu))
        for i, g in enumerate(self.rngs):
            z[i] = g.standard_normal((per_loop, self.nu))
        return z

    # -- the two calls of the runner's `_step_once` (run_classical.py:425-431) -------------------------------------
    def observation_for_controller(self, **obs: np.ndarray) -> Dict[str, np.ndarray]:
        """`obs`: [B, ...] arrays of the current plant observation; `q` and `dq` are required, anything else
        (tau_bias, ee_pos, ...) is delayed alongside.  Returns the delayed, noise-corrupted observation with the four
        torque-measurement channels replaced (uncertainty_profiles.py:137-156)."""
        if "q" not in obs or "dq" not in obs:
            raise ValueError("observation needs q and dq")
        obs = {k: np.array(v, dtype=float) for k, v in obs.items()}
        for k, v in obs.items():
            if v.shape[0] != self.B:
                raise ValueError(f"{k}: leading dimension {v.shape[0]} != {self.B} loops")
        L = int(self.obs_delay_steps.max()) + 1
        if self._obs_ring is None:
            self._obs_ring = {k: np.repeat(v[None], L, axis=0) for k, v in obs.items()}
        else:
            for k, v in obs.items():
                self._obs_ring[k][self._k % L] = v
        src = np.maximum(self._k - self.obs_delay_steps, 0) % L
        delayed = {k: ring[src, self._rows].copy() for k, ring in self._obs_ring.items()}
        self._k += 1

        z = self._draw(3)
        # numpy's normal(0, s) is 0.0 + s*z
        delayed["q"] = delayed["q"] + (0.0 + self.sig[:, 0, None] * z[:, 0])
        delayed["dq"] = delayed["dq"] + (0.0 + self.sig[:, 1, None] * z[:, 1])
        tau_hat = self._tau_hat(0.0 + self.sig[:, 2, None] * z[:, 2])
        self._tau_hat_filt = (1.0 - self._alpha) * self._tau_hat_filt + self._alpha * tau_hat
        for k in ("tau_meas", "tau_meas_act"):
            delayed[k] = tau_hat.copy()
        for k in ("tau_meas_filt", "tau_meas_act_filt"):
            delayed[k] = self._tau_hat_filt.copy()
        return delayed

    def command_for_plant(self, tau_cmd_nominal: np.ndarray) -> np.ndarray:
        tau = np.asarray(tau_cmd_nominal, dtype=float).reshape(self.B, self.nu)
        self._cmd_ring[self._n_cmd % self._cmd_ring.shape[0]] = tau
        self._n_cmd += 1
        return self._tau_hat(0.0 + self.sig[:, 2, None] * self._draw(1)[:, 0])


def injector_for_scenarios(scenarios: Sequence[str], dt: float, seeds: Optional[Sequence[int]] = None,
                           tau_lpf_alpha: float = 0.2) -> Optional[BatchedUncertaintyInjector]:
    """One stream per loop whose scenario carries a profile; `seeds[b]` defaults to the scenario's own seed, which is
    what the runner passes (run_classical.py:386).  Loops of the other scenarios must not be mixed in: the runner
    treats them differently (`tau_applied = tau_cmd * torque_scale`, :428-430) -- see `plant_command`."""
    cfgs = [config_for_scenario(s, seed=scenario_seed(s) if seeds is None else int(seeds[i])) for i, s in enumerate(scenarios)]
    if all(c is None for c in cfgs):
        return None
    if any(c is None for c in cfgs):
        raise ValueError("mixing loops with and without an uncertainty profile in one injector")
    return BatchedUncertaintyInjector(dt, 7, cfgs, tau_lpf_alpha)


def plant_command(tau_cmd: np.ndarray, scenarios: Sequence[str], injector: Optional[BatchedUncertaintyInjector]) -> np.ndarray:
    """What reaches the plant (run_classical.py:427-430): the injector's delayed/gained/biased command in benchmark
    mode, else the nominal command times the scenario's torque scale."""
    if injector is not None:
        return injector.command_for_plant(tau_cmd)
    scale = np.stack([scenario_settings(s)["torque_scale"] for s in scenarios])
    return np.asarray(tau_cmd, dtype=float) * scale


# ----------------------------------------------------------------------------------------------------------------
# end-of-run metrics (run_classical.py:513-534), vectorised over loops
# ----------------------------------------------------------------------------------------------------------------
def closed_loop_metrics(t: np.ndarray, err_tan: np.ndarray, err_3d: np.ndarray, fn_meas: np.ndarray, contact: np.ndarray,
                        fn_des: float, t_contact_phase: float) -> Dict[str, np.ndarray]:
    """Inputs [B, T] (or [T] -> one loop); `t` may be [T], shared by all loops.  Empty selections give NaN as in the
    reference.  Keys are the names `logger.set_meta` uses there."""
    err_tan, err_3d, fn_meas, contact = (np.atleast_2d(np.asarray(x, dtype=float)) for x in (err_tan, err_3d, fn_meas, contact))
    B, T = err_tan.shape
    t = np.broadcast_to(np.asarray(t, dtype=float), (B, T))
    phase = t >= float(t_contact_phase)
    nan = np.full(B, np.nan)

    def rows(fn, arr, mask=None):
        out = nan.copy()
        for b in range(B):
            sel = arr[b] if mask is None else arr[b][mask[b]]
            if sel.size > 0:
                out[b] = fn(sel)
        return out

    return {
        "rms_tangential_error": rows(lambda e: np.sqrt(np.mean(e ** 2)), err_tan),
        "rms_3d_error": rows(lambda e: np.sqrt(np.mean(e ** 2)), err_3d),
        "rms_tangential_error_contact_phase": rows(lambda e: np.sqrt(np.mean(e ** 2)), err_tan, phase),
        "avg_abs_position_err": rows(lambda e: np.mean(np.abs(e)), err_tan),
        "avg_abs_force_err": rows(lambda f: np.mean(np.abs(f - float(fn_des))), fn_meas),
        "max_fn": rows(np.max, fn_meas),
        "contact_loss_pct": rows(lambda c: (1.0 - np.mean(c)) * 100.0, contact),
        "contact_loss_contact_phase_pct": rows(lambda c: (1.0 - np.mean(c)) * 100.0, contact, phase),
        "fn_mean_contact_phase": rows(np.mean, fn_meas, phase),
    }


# ----------------------------------------------------------------------------------------------------------------
# on-disk run format (src/utils/logging.py:33-151), written from columnar [T, B, ...] storage
# ----------------------------------------------------------------------------------------------------------------
def _jsonable(x: Any) -> Any:
    """Conversion rules of the reference's meta writer (logging.py:13-30)."""
    if x is None or isinstance(x, (str, int, float, bool)):
        return x
    if is_dataclass(x):
        return asdict(x)
    if isinstance(x, Path):
        return str(x)
    if isinstance(x, dict):
        return {str(k): _jsonable(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [_jsonable(v) for v in x]
    if isinstance(x, np.ndarray):
        return x.tolist()
    return str(x)


class BatchRunLog:
    """Columnar log of B closed loops.  `log(**fields)` takes one tick: a field is a [B] array (a scalar per loop;
    integer dtypes stay integers in the CSV, as `int(...)` values do in the runner), a [B, n] array (a vector per
    loop) or a Python scalar (same for all loops).  `save()` writes one reference-format run directory per loop."""

    def __init__(self, run_names: Sequence[str], results_dir: Path | str = "results", notes: Optional[Sequence[Dict[str, Any]]] = None,
                 overwrite: bool = False, stamp: Optional[str] = None):
        self.B = len(run_names)
        self.logs_dir = Path(results_dir) / "logs"
        self.logs_dir.mkdir(parents=True, exist_ok=True)
        self.stamp = stamp or time.strftime("%Y%m%d_%H%M%S")
        self.run_dirs = [self.logs_dir / f"{self.stamp}_{n}" for n in run_names]
        if len(set(self.run_dirs)) != self.B:
            raise ValueError("run names must be distinct (one directory per loop)")
        for d in self.run_dirs:
            if d.exists() and not overwrite:
                raise FileExistsError(f"Run dir exists: {d}")
        for d in self.run_dirs:
            d.mkdir(parents=True, exist_ok=True)
        self._cols: Dict[str, List[np.ndarray]] = {}
        self._T = 0
        self.meta: List[Dict[str, Any]] = [
            {"run_name": n, "timestamp": self.stamp, "notes": _jsonable((notes[i] if notes else None) or {})}
            for i, n in enumerate(run_names)]

    def log(self, **fields: Any) -> None:
        if self._T and set(fields) != set(self._cols):
            raise ValueError("every tick must log the same fields")
        for k, v in fields.items():
            a = np.asarray(v)
            if a.ndim == 0:
                a = np.broadcast_to(a, (self.B,))
            if a.shape[0] != self.B or a.ndim > 2:
                raise ValueError(f"{k}: expected a scalar, [B] or [B, n] with B={self.B}, got {a.shape}")
            self._cols.setdefault(k, []).append(np.array(a))
        self._T += 1

    def set_meta(self, **kwargs: Any) -> None:
        """Meta entries shared by every loop (logging.py:84-85)."""
        for m in self.meta:
            m.update(_jsonable(kwargs))

    def set_meta_rows(self, **kwargs: Sequence[Any]) -> None:
        """Per-loop meta entries: each value has one row per loop (e.g. the arrays of `closed_loop_metrics`)."""
        for k, v in kwargs.items():
            if len(v) != self.B:
                raise ValueError(f"{k}: {len(v)} rows for {self.B} loops")
            for i, m in enumerate(self.meta):
                vi = v[i]
                m[k] = _jsonable(vi.item() if isinstance(vi, np.generic) else vi)

    def save(self) -> None:
        if self._T == 0:
            return
        keys = sorted(self._cols)
        data = {k: np.stack(self._cols[k], axis=0) for k in keys}          # [T, B] or [T, B, n]
        header: List[str] = []
        for k in keys:
            a = data[k]
            header.extend([k] if a.ndim == 2 or a.shape[2] > 10 else [f"{k}[{i}]" for i in range(a.shape[2])])
        for b, d in enumerate(self.run_dirs):
            np.savez_compressed(d / "data.npz", **{k: (a[:, b].astype(float) if a.ndim == 2 else a[:, b]) for k, a in data.items()})
            with open(d / "data.csv", "w", newline="") as f:
                w = csv.writer(f)
                w.writerow(header)
                cols = []
                for k in keys:
                    a = data[k][:, b]
                    if a.ndim == 1:
                        cols.append(a.tolist())
                    elif a.shape[1] <= 10:
                        cols.extend(a[:, i].tolist() for i in range(a.shape[1]))
                    else:
                        cols.append([r.tolist() for r in a])
                w.writerows(zip(*cols))
            with open(d / "meta.json", "w") as f:
                json.dump(self.meta[b], f, indent=2)


# ----------------------------------------------------------------------------------------------------------------
# B closed loops around the device-side tick (run_classical.py:425-510 once per loop in the reference)
# ----------------------------------------------------------------------------------------------------------------
class BatchedSurfaceLatch:
    """`_detect_surface` (crocoddyl_classical.py:286-303) for B loops: latch on measured normal force above
    `fn_contact_on` or on (trajectory says surface AND the tool is within `z_contact_band` of the table); release after
    `contact_release_steps` consecutive ticks below `fn_contact_off`."""

    def __init__(self, B: int, cfg):
        self.cfg = cfg
        self.latched = np.zeros(B, dtype=bool)
        self.loss = np.zeros(B, dtype=np.int64)

    def update(self, fn: np.ndarray, ee_z: Optional[np.ndarray], hint: np.ndarray) -> np.ndarray:
        c = self.cfg
        fn = np.asarray(fn, dtype=float)
        ee_z = np.full(fn.shape, np.inf) if ee_z is None else np.asarray(ee_z, dtype=float)
        near = np.isfinite(ee_z) & (ee_z <= float(c.z_contact) + float(c.z_contact_band))
        was = self.latched.copy()
        self.loss = np.where(was & (fn < c.fn_contact_off), self.loss + 1, 0)
        release = was & (self.loss >= int(c.contact_release_steps))
        grab = ~was & ((fn > c.fn_contact_on) | (np.asarray(hint, dtype=bool) & near))
        self.latched = (was & ~release) | grab
        self.loss = np.where(release | grab, 0, self.loss)
        return self.latched.copy()


def torque_state(obs: Dict[str, np.ndarray], cfg) -> np.ndarray:
    """[B, 7] torque part of the FF state: per loop the first finite source in the reference's order
    (`_tau_state_from_obs`, crocoddyl_force_feedback.py:512-536)."""
    src = str(cfg.ff_tau_state_source).strip().lower()
    if src == "auto":
        src = "tau_meas_filt" if bool(cfg.ff_use_tau_meas_filt) else "tau_meas"
    from .controllers import ForceFeedbackCrocoddylMPC as _FF   # same source table as the B = 1 controller mirror

    keys = _FF._TAU_SOURCES.get(src, ("tau_meas_act_filt", "tau_meas_act", "tau_cmd", "tau_meas")) + ("tau_cmd",)
    B = obs["q"].shape[0]
    out, done = np.zeros((B, 7)), np.zeros(B, dtype=bool)
    for k in keys:
        if k in obs and obs[k] is not None:
            v = np.asarray(obs[k], dtype=float).reshape(B, 7)
            take = ~done & np.isfinite(v).all(axis=1)
            out[take], done = v[take], done | take
    return out


def run_closed_loops(eng, cfg, plant, traj, scenarios: Sequence[str], steps: int, *, ff: bool, q_nom: np.ndarray,
                     benchmark_mode: bool = True, seeds: Optional[Sequence[int]] = None, log: Optional[BatchRunLog] = None,
                     t_contact_phase: float = 0.8, contact_threshold: float = 0.5, fixed_iters: bool = False) -> Dict[str, np.ndarray]:
    """The runner's loop for B loops at once: per tick one (optional) injector pass, ONE `tick_host` call (shift,
    solve, policy, guards on the device), one plant step; the runner's per-step log rows and its summary metrics.

    `plant`: B plants behind one object -- `.dt`, `.observe() -> dict` and `.step(tau_applied[B,7]) -> dict` with
    [B, ...] arrays `q, dq, tau_bias, ee_pos, ee_vel, fn_meas` plus whatever torque channels the FF state source
    reads (MuJoCo is unavailable on this box; tests use a kinematic stand-in).  `traj(t[B]) -> (p_mj, v_mj, surf)` is
    `synthetic.ApproachThenCircle`.  All loops of one call share one injector kind (see `injector_for_scenarios`).

    Same OCP as the B = 1 controller mirror: `posture_ref_mode` and `torque_ref_mode` are resolved per tick into the
    tick's x_reg_ref / tau_ref inputs; the surface latch sees what the controller sees (the injector-delayed
    observation).  `mpc_update_steps` must be 1 (the device tick solves on every tick; every preset of the reference
    uses 1).  The logged `fn_pred` is the RAW prediction: the reference's FF controller additionally passes it through
    its lag/affine alignment filter for logging only (`_align_logged_force_prediction`); that filter lives in the
    B = 1 mirror (controllers.py), not here."""
    from . import config as _config
    from .synthetic import batched_gravity_torque, refs_pin

    if int(cfg.mpc_update_steps) != 1:
        raise ValueError("run_closed_loops: the device tick solves on every tick; mpc_update_steps must be 1")
    B = len(scenarios)
    inj = injector_for_scenarios(scenarios, plant.dt, seeds, getattr(plant, "tau_meas_lpf_alpha", 0.2)) if benchmark_mode else None
    pol = _config.tick_policy(cfg, dt_mpc=plant.dt)
    latch = BatchedSurfaceLatch(B, cfg)
    force_latch = str(cfg.phase_source).strip().lower() == "force_latch"
    x_reg = np.concatenate([np.broadcast_to(np.asarray(q_nom, float), (B, 7)), np.zeros((B, 7))], axis=1)
    N, dt_ocp = int(cfg.horizon), _config.dt_ocp_of(cfg)
    tmode = str(cfg.torque_ref_mode).strip().lower()
    tau_ref = None                                       # "gravity_x0": g(q_now), evaluated on the device
    if tmode == "zero":
        tau_ref = np.zeros((B, 7))
    elif tmode == "gravity_qnom":
        tau_ref = np.repeat(batched_gravity_torque(getattr(plant, "m", None) or plant.model, np.asarray(q_nom, float)[None]), B, axis=0)
    obs = plant.observe()
    cols = {k: [] for k in ("t", "err_tan", "err_3d", "fn_meas", "fn_pred", "contact")}
    ctl_hist = {k: [] for k in ("tau_cmd", "solver_cost", "solver_iters")}   # what the controllers did, per tick
    t = 0.0
    for k in range(int(steps)):
        seen = inj.observation_for_controller(**{n: v for n, v in obs.items() if isinstance(v, np.ndarray) and v.ndim >= 1}) if inj else obs
        _, _, hint = traj(np.full(B, t))
        # the latch reads the controller's view of the force (obs.f_contact_normal after the injector's delay)
        mode = latch.update(seen.get("fn_meas", obs["fn_meas"]), np.asarray(seen["ee_pos"])[:, 2], hint) if force_latch else np.asarray(hint, dtype=bool)
        x_now = np.concatenate([seen["q"], seen["dq"]] + ([torque_state(seen, cfg)] if ff else []), axis=1)
        p_ref, v_ref = refs_pin(traj, np.full(B, t), N, dt_ocp)
        x_reg_k = x_reg if str(cfg.posture_ref_mode).strip().lower() == "q_nom" else x_now[:, :14]
        out = eng.tick_host(mode.astype(np.int32), x_now, p_ref, v_ref, x_reg_k, seen["tau_bias"], pol, int(cfg.max_iters),
                            fixed_iters=fixed_iters, reset=(k == 0), tau_ref=tau_ref)
        tau_applied = plant_command(out["tau_cmd"], scenarios, inj)
        obs = plant.step(tau_applied)
        t += plant.dt
        p_next, v_next, surf_next = traj(np.full(B, t))
        err = obs["ee_pos"] - p_next
        row = dict(t=t, err_tan=np.linalg.norm(err[:, :2], axis=1), err_3d=np.linalg.norm(err, axis=1), fn_meas=np.asarray(obs["fn_meas"], float),
                   fn_pred=out["fn_pred"], contact=(np.asarray(obs["fn_meas"]) > contact_threshold).astype(np.int64))
        for n, v in row.items():
            cols[n].append(np.broadcast_to(np.asarray(v, dtype=float), (B,)).copy())
        ctl_hist["tau_cmd"].append(out["tau_cmd"].copy())
        ctl_hist["solver_cost"].append(out["cost"].copy())
        ctl_hist["solver_iters"].append(out["iters"].copy())
        if log is not None:
            log.log(**row, ee_pos=obs["ee_pos"], ee_ref=p_next, ee_vel=obs["ee_vel"], ee_vel_ref=v_next, fn_des=float(cfg.fn_des),
                    tau_cmd=out["tau_cmd"], tau_applied=tau_applied, surface_ref=np.asarray(surf_next).astype(np.int64),
                    solver_iters=out["iters"].astype(np.int64), solver_cost=out["cost"], solver_success=out["ok"].astype(np.int64),
                    solver_unstable=out["unstable"].astype(np.int64), tau_raw_inf=np.abs(out["tau_raw"]).max(axis=1),
                    tau_cmd_inf=np.abs(out["tau_cmd"]).max(axis=1))
    hist = {n: np.stack(v, axis=1) for n, v in cols.items()}        # [B, T]
    metrics = closed_loop_metrics(hist["t"], hist["err_tan"], hist["err_3d"], hist["fn_meas"], hist["contact"], float(cfg.fn_des), t_contact_phase)
    if log is not None:
        log.set_meta(total_time=float(steps * plant.dt), dt=float(plant.dt))
        log.set_meta_rows(scenario_label=[scenario_settings(s)["label"] for s in scenarios],
                          scenario_tilt_deg=[float(scenario_settings(s)["tilt_deg"]) for s in scenarios])
        log.set_meta(contact_definition="in_contact = (fn_meas > 0.5 N)", benchmark_mode=bool(benchmark_mode))
        log.set_meta_rows(uncertainty_profile=[inj.meta(b) if inj else None for b in range(B)],
                          torque_scale=[scenario_settings(s)["torque_scale"] for s in scenarios])
        log.set_meta(fn_des=float(cfg.fn_des))
        log.set_meta_rows(**metrics)
        log.set_meta(contact_phase_start_s=float(t_contact_phase))
        log.save()
    return dict(hist, **metrics, **{n: np.stack(v, axis=1) for n, v in ctl_hist.items()})
