// ffmpc_tick.cuh -- device side of the receding-horizon tick (include/ffmpc.h, ffmpc_tick_batch): the O(7) per-tick logic
// the reference runs in Python around every solve, batched:
//   k_tick_prepare : `_shift_guess` (crocoddyl_classical.py:733-757, FF :1014-1039) + plan invalidation on a mode switch (:321-329)
//   k_tick_policy  : `_policy_control` (:759-779; FF Eq. 14-18, :1041-1093), unstable guard (:393-403), `_safe_tau` (:260-284),
//                    predicted normal force (:905-942; FF :1219-1243)
// HBM-trivial kernels (a few KB per instance); the solve between them is the hot path.
#pragma once

#include "ffmpc_kernels.cuh"

namespace ffmpc {

struct TickState {
  int *plan_valid;    // [B] a plan of the previous tick exists (xs/us in workspace slot cur[b], K in ws.K)
  int *prev_mode;     // [B]
  double *tau_prev;   // [B,7] last command (`_tau_prev`)
  double *xs_init;    // [B,N+1,nx] warm start handed to the solve
  double *us_init;    // [B,N,7]
};

struct TickArgs {
  TickState st;
  int reset;
  const int *mode;
  const double *x_now, *tau_bias;
  ffmpc_tick_policy pol;
  // outputs (nullable)
  double *tau_cmd, *tau_raw, *cost, *fn_pred;
  int *iters, *ok, *unstable;
};

FFMPC_DEV bool tick_plan_usable(const TickArgs &A, int b) {
  return !A.reset && A.st.plan_valid[b] && A.st.prev_mode[b] == A.mode[b];
}

static __global__ void k_tick_prepare(Workspace ws, TickArgs A) {
  const int N = ws.N, nx = ws.nx;
  const size_t nxs = (size_t)ws.B * (N + 1) * nx, nus = (size_t)ws.B * N * NJ;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)global_tid(); i < nxs; i += stride) {
    const int b = (int)(i / ((size_t)(N + 1) * nx));
    const int r = (int)(i - (size_t)b * (N + 1) * nx);
    const int t = r / nx, c = r - t * nx;
    double v = A.x_now[(size_t)b * nx + c];
    if (t > 0 && tick_plan_usable(A, b)) v = ws.xs[(size_t)ws.cur[b] * nxs + i];  // [x0] + xs_prev[1:]
    A.st.xs_init[i] = v;
  }
  for (size_t i = (size_t)global_tid(); i < nus; i += stride) {
    const int b = (int)(i / ((size_t)N * NJ));
    const int r = (int)(i - (size_t)b * N * NJ);
    const int t = r / NJ, a = r - t * NJ;
    double v;
    if (tick_plan_usable(A, b)) {
      const int tp = (t + 1 < N) ? t + 1 : N - 1;  // us_prev[1:] + [us_prev[-1]]
      v = ws.us[(size_t)ws.cur[b] * nus + ((size_t)b * N + tp) * NJ + a];
    } else if (nx == 21) {
      v = A.x_now[(size_t)b * 21 + 14 + a];        // FF cold start holds tau_hat
    } else {
      v = (A.reset ? A.tau_bias : A.st.tau_prev)[(size_t)b * NJ + a];  // classical cold start holds the last command
    }
    A.st.us_init[i] = v;
  }
}

static __global__ void k_tick_policy(const __grid_constant__ Params P, TickArgs A) {
  const Workspace &ws = P.ws;
  const ffmpc_config &cfg = P.cfg;
  const ffmpc_tick_policy &pol = A.pol;
  const int b = global_tid();
  if (b >= ws.B) return;
  const int N = ws.N, nx = ws.nx;
  const bool ff = nx == 21;
  const size_t slot = (size_t)ws.cur[b] * ws.B + b;
  const double *xs0 = ws.xs + slot * (N + 1) * nx, *us0 = ws.us + slot * N * NJ;
  const double *K0 = ws.K + (size_t)b * N * NJ * nx;
  const double *xn = A.x_now + (size_t)b * nx;
  const double *bias = A.tau_bias + (size_t)b * NJ;
  double *tprev = A.st.tau_prev + (size_t)b * NJ;
  bool planok = true;
  for (int a = 0; a < NJ; ++a) planok = planok && isfinite(us0[a]);
  const double s = pol.feedback_gain_scale;
  const double eps_full = fmin(fmax(pol.dt_mpc / cfg.dt_ocp, 0.0), 1.0);
  double tau[NJ];
  if (!planok) {
    for (int a = 0; a < NJ; ++a) tau[a] = ff ? xn[14 + a] : (A.reset ? bias[a] : tprev[a]);
  } else if (!ff) {
    for (int a = 0; a < NJ; ++a) {
      double u = us0[a];
      if (pol.use_feedback_policy) {
        double kd = 0.0;
        for (int c = 0; c < 14; ++c) kd += K0[a * 14 + c] * (xn[c] - xs0[c]);
        u += s * kd;  // PLUS: the reference's sign (crocoddyl_classical.py:775-777)
      }
      tau[a] = u;
    }
  } else {
    const double alpha = fmin(fmax(cfg.ff_alpha, 0.0), 0.999999);
    const double eps = pol.ff_use_tau_interpolation ? eps_full : 0.0;
    const double gain = eps * (1.0 - alpha);
    for (int a = 0; a < NJ; ++a) {
      const double tau0 = xs0[14 + a], tau1 = xs0[21 + 14 + a];
      double u = tau0 + eps * (tau1 - tau0);
      if (pol.use_feedback_policy) {
        double fb = tau0 - xn[14 + a];  // identity part of K_tilde_tau
        for (int c = 0; c < 14; ++c) fb += gain * K0[a * 21 + c] * (xs0[c] - xn[c]);
        for (int c = 0; c < NJ; ++c) fb += gain * (K0[a * 21 + 14 + c] - (c == a ? 1.0 : 0.0)) * (xs0[14 + c] - xn[14 + c]);
        u += s * fb;
      }
      tau[a] = u;
    }
  }
  double raw[NJ], raw_inf = 0.0;
  for (int a = 0; a < NJ; ++a) {
    raw[a] = tau[a];
    if (ff && pol.ff_inverse_actuation_model) raw[a] = (tau[a] - pol.ff_alpha_ctrl * xn[14 + a]) / fmax(1.0e-6, 1.0 - pol.ff_alpha_ctrl);
    raw_inf = fmax(raw_inf, fabs(raw[a]));
    if (isnan(raw[a])) raw_inf = NAN;  // np.max(np.abs(.)) propagates NaN
  }
  const double cost = ws.cost[b];
  // the reference's guard (crocoddyl_classical.py:393-403): `tau_raw_inf > max` is False for NaN -- a non-finite command
  // does not trip the guard, `_safe_tau` below then holds the last command
  const bool unstable = !isfinite(cost) || cost > pol.max_solver_cost || raw_inf > pol.max_tau_raw_inf;
  if (A.tau_raw) for (int a = 0; a < NJ; ++a) A.tau_raw[(size_t)b * NJ + a] = tau[a];
  if (unstable)
    for (int a = 0; a < NJ; ++a) raw[a] = bias[a] - pol.fallback_dq_damping * xn[NJ + a];
  // _safe_tau (the reference replaces the WHOLE target by the last command when any entry is non-finite)
  bool allfin = true;
  for (int c = 0; c < NJ; ++c) allfin = allfin && isfinite(raw[c]);
  for (int a = 0; a < NJ; ++a) {
    const double lim = cfg.tau_limits[a];
    const double prev = A.reset ? bias[a] : tprev[a];
    double t = allfin ? raw[a] : prev;
    t = fmin(fmax(t, -lim), lim);
    if (pol.apply_command_filter) {
      double d = fmin(fmax(t - prev, -pol.tau_trust_inf), pol.tau_trust_inf);
      const double slew = pol.tau_rate_limit[a] * pol.dt_mpc;
      d = fmin(fmax(d, -slew), slew);
      const double sm = fmin(fmax(pol.tau_smoothing_alpha, 0.0), 1.0);
      t = (1.0 - sm) * prev + sm * (prev + d);
      t = fmin(fmax(t, -lim), lim);
    }
    raw[a] = t;
  }
  for (int a = 0; a < NJ; ++a) {
    tprev[a] = raw[a];
    if (A.tau_cmd) A.tau_cmd[(size_t)b * NJ + a] = raw[a];
  }
  const bool contact = A.mode[b] == FFMPC_MODE_CONTACT;
  if (A.fn_pred) {
    double f = NAN;
    if (contact) {
      const double f0 = ws.f_pred[b * 6 + 2], f1 = ws.f_pred[b * 6 + 5];
      if (!ff) f = f0;
      else if (N == 1 || !isfinite(f1)) f = fabs(f0);
      else if (!isfinite(f0)) f = fabs(f1);
      else if (pol.dt_mpc >= cfg.dt_ocp - 1.0e-9) f = fabs(f0);
      else f = (1.0 - eps_full) * fabs(f0) + eps_full * fabs(f1);
    }
    A.fn_pred[b] = f;
  }
  if (A.cost) A.cost[b] = cost;
  if (A.iters) A.iters[b] = ws.iter[b];
  if (A.ok) A.ok[b] = ws.ok[b];
  if (A.unstable) A.unstable[b] = unstable ? 1 : 0;
  A.st.plan_valid[b] = (planok && !unstable) ? 1 : 0;
  A.st.prev_mode[b] = A.mode[b];
}

}  // namespace ffmpc
