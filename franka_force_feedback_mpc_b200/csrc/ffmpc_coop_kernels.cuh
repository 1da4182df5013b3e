// ffmpc_coop_kernels.cuh -- kernels built on the lane-cooperative knot evaluation of ffmpc_coop.cuh:
//   k_node_diff_coop : octet (8 lanes) per (instance, knot)   ShootingProblem.calc + calcDiff -> records, gaps
//   k_forward_coop   : octet per instance                      forwardPass + sequential line search + acceptance
//   k_node_eval_coop : octet per evaluation point              parity-test hook (records -> k_node_expand -> dense)
// Block = 128 threads = 16 octets; dynamic shared memory = constant table + 16 octet scratch areas.
#pragma once

#include "ffmpc_coop.cuh"
#include "ffmpc_kernels.cuh"

namespace ffmpc {

#ifndef FFMPC_COOP_BLOCK
#define FFMPC_COOP_BLOCK 128   // threads per block of the knot-evaluation kernels (16 knots)
#endif
constexpr int kCoopBlock = FFMPC_COOP_BLOCK;
#ifndef FFMPC_COOP_MINBLOCKS
#define FFMPC_COOP_MINBLOCKS 3  // resident blocks per SM the register allocation is sized for
#endif
#ifndef FFMPC_FWD_MINBLOCKS
#define FFMPC_FWD_MINBLOCKS 3   // the rollout kernel's register allocation: 128-thread equivalents per SM
#endif
constexpr size_t kCoopSmemBytes = (size_t)(CT_SIZE + (kCoopBlock / 8) * CS_SIZE) * sizeof(double);
// the rollout is latency-bound (51 dependent knots per instance): one warp (4 instances) per block spreads a small
// line-search queue over all SMs instead of packing it into a few 4-warp blocks
constexpr int kFwdCoopBlock = 32;
constexpr size_t kFwdCoopSmemBytes = (size_t)(CT_SIZE + (kFwdCoopBlock / 8) * CS_SIZE) * sizeof(double);

// bring one 128-byte line into L1 ahead of use (no register cost); the rollout issues these for knot t+1 while knot t
// is evaluated: its global operands (K_t, xs_t, fs_t, Vxx fs_t, us_t, k_t, references) do not depend on the rollout state
FFMPC_DEV void prefetch_l1(const void *p) {
#if !defined(FFMPC_EMU) && !defined(FFMPC_NO_PREFETCH)
#ifdef FFMPC_PREFETCH_L2
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#else
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#endif
#else
  (void)p;
#endif
}

FFMPC_DEV double *coop_smem_base() {
#ifdef FFMPC_EMU
  return (double *)emu::S().smem;
#else
  extern __shared__ __align__(16) double coop_dyn_smem[];
  return coop_dyn_smem;
#endif
}

// ------------------------------------------------------------------------------------------------
template <int NC, bool FF>
__global__ void __launch_bounds__(kCoopBlock, FFMPC_COOP_MINBLOCKS) k_node_diff_coop(const __grid_constant__ Params P) {
  const Workspace &ws = P.ws;
  constexpr int nx = FF ? 21 : 14;
  const int N = ws.N;
  const int nfree = ws.counts[0];
  const int base = (NC == 0) ? 0 : nfree;
  const int cnt = (NC == 0) ? nfree : ws.counts[1];
  if ((long long)blockIdx.x * (kCoopBlock / 8) >= (long long)cnt * (N + 1)) return;  // block without work (block-uniform)
  double *ct = coop_smem_base();
  double *sm = ct + CT_SIZE + (size_t)(threadIdx.x >> 3) * CS_SIZE;
  const int g = global_tid();
  const int o = g >> 3, j = g & 7;
  const int jj = j < NJ ? j : NJ - 1;
  const bool act = j < NJ;
  int slot = o / (N + 1), t = o - slot * (N + 1);
  bool valid = slot < cnt;
  if (!valid) { slot = 0; t = 0; }
  // every global operand of the knot (the dependent index chain perm -> done / recalc / cur -> xs, us, references) is
  // requested before the constant table is built, so that the memory latency hides behind the table fill
  const int b = ws.perm[base + slot];
  const int b_done = ws.done[b], b_recalc = ws.recalc[b], b_cur = ws.cur[b];
  prefetch_l1(ws.p_ref + ((size_t)b * (N + 1) + t) * 3);
  prefetch_l1(ws.v_ref + ((size_t)b * (N + 1) + t) * 3);
  const size_t bufoff = (size_t)b_cur * ws.B;
  const double *xg = ws.xs + ((bufoff + b) * (N + 1) + t) * nx;
  CoopX x, xn;
  x.q = xg[jj];
  x.v = xg[NJ + jj];
  x.t = FF ? xg[2 * NJ + jj] : 0.0;
  const double u = (t < N) ? ws.us[((bufoff + b) * N + t) * NJ + jj] : 0.0;
  CoopCtx ctx;
  ctx.p_ref = ws.p_ref + ((size_t)b * (N + 1) + t) * 3;
  ctx.v_ref = ws.v_ref + ((size_t)b * (N + 1) + t) * 3;
  ctx.xr_q = ws.x_reg_ref[(size_t)b * 14 + jj];
  ctx.xr_v = ws.x_reg_ref[(size_t)b * 14 + NJ + jj];
  ctx.tau_ref = ws.tau_ref[(size_t)b * NJ + jj];
  const double *x0 = ws.x0 + (size_t)b * nx;
  ctx.yr_q = x0[jj];
  ctx.yr_v = x0[NJ + jj];
  ctx.yr_t = FF ? x0[2 * NJ + jj] : 0.0;
  ctx.terminal = (t == N);
  coop_fill_table(P.mdl, P.cfg, ct);
  __syncthreads();
  if (valid && (b_done || !b_recalc)) valid = false;
  if (!__any_sync(kFull, valid)) return;
  double cost, lam[3];
  double *rec = ws.rec + ((size_t)b * (N + 1) + t) * REC_SIZE;
  coop_action_node<NC, FF, true>(P.mdl, P.cfg, ct, sm, j, ctx, x, u, xn, cost, lam, rec, valid);
  if (!valid) return;
  if (j == 0) {
    ws.ncost[(size_t)b * (N + 1) + t] = cost;
    if (NC != 0 && t < 2) {
      ws.f_pred[b * 6 + t * 3 + 0] = lam[0];
      ws.f_pred[b * 6 + t * 3 + 1] = lam[1];
      ws.f_pred[b * 6 + t * 3 + 2] = lam[2];
    }
  }
  if (!act) return;
  // gaps: fs[0] = x0 - xs[0]; fs[t+1] = xnext_t - xs[t+1]
  double *fsb = ws.fs + (size_t)b * (N + 1) * nx;
  if (t == 0) {
    fsb[j] = x0[j] - x.q;
    fsb[NJ + j] = x0[NJ + j] - x.v;
    if (FF) fsb[2 * NJ + j] = x0[2 * NJ + j] - x.t;
  }
  if (t < N) {
    const double *xnx = xg + nx;
    double *f = fsb + (size_t)(t + 1) * nx;
    f[j] = xn.q - xnx[j];
    f[NJ + j] = xn.v - xnx[NJ + j];
    if (FF) f[2 * NJ + j] = xn.t - xnx[2 * NJ + j];
  }
}

// End of one line search for instance b (SolverFDDP::solve after the step-length loop): accept the trial or keep the
// candidate, regularisation update, stopping test.  `steplength` is the accepted one, or the last one tried.
FFMPC_DEV void finish_line_search(const Workspace &ws, const ffmpc_config &cfg, int b, int it, bool accepted, bool feasible,
                                   double steplength, double cost_try, int trial_slot, double d1, double d2) {
  double preg = ws.preg[b];
  int wasfeas = ws.wasfeas[b];
  if (accepted) {
    wasfeas = feasible ? 1 : 0;
    ws.wasfeas[b] = wasfeas;
    ws.feas[b] = (wasfeas || steplength == 1.0) ? 1 : 0;
    ws.cur[b] = trial_slot;
    ws.cost[b] = cost_try;
    ws.recalc[b] = 1;
  } else {
    ws.recalc[b] = 0;
  }
  bool finished = false;
  int okflag = 0;
  if (steplength > cfg.th_stepdec) {
    preg /= cfg.reg_decfactor;
    if (preg < cfg.reg_min) preg = cfg.reg_min;
  }
  if (steplength <= cfg.th_stepinc) {
    preg *= cfg.reg_incfactor;
    if (preg > cfg.reg_max) preg = cfg.reg_max;
    if (preg == cfg.reg_max) finished = true;
  }
  ws.preg[b] = preg;
  if (!finished) {
    const double stop = fabs(d1 + 0.5 * d2);
    ws.stop[b] = stop;
    if (!ws.fixed_iters && wasfeas && stop < cfg.th_stop) {
      finished = true;
      okflag = 1;
    }
  }
  if (finished) {
    ws.done[b] = 1;
    ws.ok[b] = okflag;
    ws.iter[b] = it;
  } else {
    ws.iter[b] = it + 1;
  }
}

FFMPC_DEV bool accept_step(const ffmpc_config &cfg, bool feasible, double steplength, double dV, double d1, double d2) {
  const double dVexp = steplength * (d1 + 0.5 * steplength * d2);
  if (dVexp >= 0.0) return d1 < cfg.th_grad || dV > cfg.th_acceptstep * dVexp;
  return !feasible && dV > cfg.th_acceptnegstep * dVexp;
}

constexpr int TRIAL_REC = 10;  // per (instance, step length) result of a speculative rollout: ok, cost, dv, lam01[6], pad

// ------------------------------------------------------------------------------------------------
// k_forward_coop: SolverFDDP::forwardPass / tryStep / expectedImprovement + acceptance + reg update.
// The step lengths 2^-ia are tried in PASSES; pass p rolls out the group of `size` consecutive step lengths that starts
// at `first` (the schedule: Workspace::gfirst / gsize, chosen on the host).  Pass 0 takes every live instance of the mode
// bucket (ws.perm), pass p > 0 the compacted queue written by pass p - 1 (instances that rejected every step length so
// far are appended with an atomic counter), so warps stay full of useful octets instead of waiting for the slowest of
// their four instances (measured: 2.4 rollouts per instance and iteration on average, 48 % need more than one).
//   size == 1: the pass decides itself (accept / queue for the next pass / give up), as the loop of SolverFDDP::solve.
//   size  > 1: SPECULATIVE group: gridDim.y = size, every step length of the group is rolled out at once into its own
//              trial slot and only recorded (ws.trial); k_forward_finalize then takes, per instance, exactly the decisions
//              of the sequential loop over the group: trials after the first accepted one "never ran".
// Same decisions and the same arithmetic for every schedule: a trial depends only on its own step length.
// One octet per (instance, step length); trial slot e of a group = the e-th trajectory slot that is not `cur`.
// ------------------------------------------------------------------------------------------------
FFMPC_DEV int trial_slot(int e, int cur) { return e + (e >= cur ? 1 : 0); }

template <int NC, bool FF>
__global__ void __launch_bounds__(128, FFMPC_FWD_MINBLOCKS) k_forward_coop(const __grid_constant__ Params P, int it, int pass, int first, int size) {
  const Workspace &ws = P.ws;
  const bool spec = size > 1;
  const int ia = first + (int)blockIdx.y;
  const ffmpc_config &cfg = P.cfg;
  constexpr int nx = FF ? 21 : 14;
  constexpr int cls = (NC == 0) ? 0 : 1;
  const int nfree = ws.counts[0];
  const int base = (NC == 0) ? 0 : nfree;
  const int cnt = (pass == 0) ? ((NC == 0) ? nfree : ws.counts[1]) : ws.qcount[pass * 2 + cls];
  if ((int)(blockIdx.x * (blockDim.x >> 3)) >= cnt) return;  // whole block without work (block-uniform)
  double *ct = coop_smem_base();
  double *sm = ct + CT_SIZE + (size_t)(threadIdx.x >> 3) * CS_SIZE;
  coop_fill_table(P.mdl, P.cfg, ct);
  __syncthreads();
  const int N = ws.N;
  const int g = global_tid();  // (blockIdx.x, threadIdx.x) only: blockIdx.y is the step length within the group
  const int o = g >> 3, j = g & 7;
  const int jj = j < NJ ? j : NJ - 1;
  const bool act = j < NJ;
  const int lane = (int)(threadIdx.x & 31);
  const unsigned omask = 0xffu << (lane & ~7);
  const int *items = (pass == 0) ? (ws.perm + base) : (ws.queue + ((size_t)((pass & 1) * 2 + cls)) * ws.B);
  bool valid = o < cnt;
  const int b = items[valid ? o : 0];
  if (valid && pass == 0 && ws.done[b]) valid = false;
  if (!__any_sync(kFull, valid)) return;
  const int cur = ws.cur[b];
  const int trial = trial_slot(ia - first, cur);
  const bool feasible = ws.feas[b] != 0;
  double cost = ws.cost[b];
  if (pass == 0 && !spec && !ws.cost_valid[b]) {
    cost = 0.0;  // cost_ = problem.calcDiff(xs, us): sum of the knot costs in order
    for (int t = 0; t <= N; ++t) cost += ws.ncost[(size_t)b * (N + 1) + t];
  }
  const double dg = ws.dg[b], dq = ws.dq[b];
  const double *xs = ws.xs + ((size_t)cur * ws.B + b) * (N + 1) * nx;
  const double *us = ws.us + ((size_t)cur * ws.B + b) * N * NJ;
  double *xt_all = ws.xs + ((size_t)trial * ws.B + b) * (N + 1) * nx;
  double *ut_all = ws.us + ((size_t)trial * ws.B + b) * N * NJ;
  const double *fs = ws.fs + (size_t)b * (N + 1) * nx;
  const double *Vfs = ws.Vfs + (size_t)b * (N + 1) * nx;
  const double *Kb = ws.K + (size_t)b * N * NJ * nx;
  const double *kb = ws.k + (size_t)b * N * NJ;
  const double *x0 = ws.x0 + (size_t)b * nx;
  CoopCtx ctx;
  ctx.xr_q = ws.x_reg_ref[(size_t)b * 14 + jj];
  ctx.xr_v = ws.x_reg_ref[(size_t)b * 14 + NJ + jj];
  ctx.tau_ref = ws.tau_ref[(size_t)b * NJ + jj];
  ctx.yr_q = x0[jj];
  ctx.yr_v = x0[NJ + jj];
  ctx.yr_t = FF ? x0[2 * NJ + jj] : 0.0;
  const double ulim = ct[jj * CT_STRIDE + CT_TAULIM];
  const double steplength = ldexp(1.0, -ia);
  const bool close = feasible || (steplength == 1.0);
  const bool wlane = valid && act;

  // ---- rollout ----
  CoopX xh;
  xh.q = ctx.yr_q; xh.v = ctx.yr_v; xh.t = ctx.yr_t;  // x0
  double ct_sum = 0.0, dv_part = 0.0;
  double lam01[6] = {NAN, NAN, NAN, NAN, NAN, NAN};
  bool failed = false;
  for (int t = 0; t <= N; ++t) {
    CoopX xt;
    const double *fst = fs + (size_t)t * nx, *xst = xs + (size_t)t * nx, *vft = Vfs + (size_t)t * nx;
    // gaps enter only while the candidate is infeasible and the step is partial (`close` is octet-uniform): no load otherwise
    xt = xh;
    if (!close) {
      xt.q = xh.q + fst[jj] * (steplength - 1.0);
      xt.v = xh.v + fst[NJ + jj] * (steplength - 1.0);
      if (FF) xt.t = xh.t + fst[2 * NJ + jj] * (steplength - 1.0);
    }
    if (!FF) xt.t = 0.0;
    const double dxq = xt.q - xst[jj], dxv = xt.v - xst[NJ + jj], dxt = FF ? xt.t - xst[2 * NJ + jj] : 0.0;
    if (wlane) {
      double *xw = xt_all + (size_t)t * nx;
      xw[j] = xt.q;
      xw[NJ + j] = xt.v;
      if (FF) xw[2 * NJ + j] = xt.t;
      // -fs_t^T Vxx_t (xs_t - xs_try_t): enters the expected improvement only while gaps are open (the sweep stores
      // Vxx fs only then)
      if (!feasible) dv_part += vft[j] * dxq + vft[NJ + j] * dxv + (FF ? vft[2 * NJ + j] * dxt : 0.0);
    }
    double u = 0.0;
    if (t < N) {
      // K dx: lane j multiplies its own three columns of every row, the reduce-scatter hands row a to lane a
      double pr[8];
      const double *Kt = Kb + (size_t)t * NJ * nx;
      // the padding lane contributes nothing: zero its dx once instead of selecting every partial product
      const double mq = act ? dxq : 0.0, mv = act ? dxv : 0.0, mt = act ? dxt : 0.0;
#pragma unroll
      for (int a = 0; a < NJ; ++a) {
        double s = Kt[a * nx + jj] * mq + Kt[a * nx + NJ + jj] * mv;
        if (FF) s += Kt[a * nx + 2 * NJ + jj] * mt;
        pr[a] = s;
      }
      pr[7] = 0.0;
      const double kd = oct_reduce_scatter8(pr, j);
      double s = us[t * NJ + jj] - kb[t * NJ + jj] * steplength;
      s -= kd;
      if (cfg.use_box_fddp) s = fmin(fmax(s, -ulim), ulim);
      u = s;
      if (wlane) ut_all[t * NJ + j] = s;
    }
    ctx.p_ref = ws.p_ref + ((size_t)b * (N + 1) + t) * 3;
    ctx.v_ref = ws.v_ref + ((size_t)b * (N + 1) + t) * 3;
    ctx.terminal = (t == N);
    if (t < N) {
      // next knot's operands -> L1 (lines spread over the 8 lanes of the octet)
      const char *Kn = (const char *)(Kb + (size_t)(t + 1) * NJ * nx);
      if (t + 1 < N) {
        prefetch_l1(Kn + 128 * j);
        if (128 * (8 + j) < NJ * nx * 8 + 128) prefetch_l1(Kn + 128 * (8 + j));
      }
      const size_t ro = (size_t)(t + 1) * nx * 8 + 128 * (j & 1);
      if (j < 2) prefetch_l1((const char *)xs + ro);
      else if (j < 4) { if (!close) prefetch_l1((const char *)fs + ro); }
      else if (j < 6) { if (!feasible) prefetch_l1((const char *)Vfs + ro); }
      else if (j == 6) prefetch_l1(us + (t + 1) * NJ);
      else prefetch_l1(kb + (t + 1) * NJ);
      if (j == 0) prefetch_l1(ctx.p_ref + 3);
      if (j == 1) prefetch_l1(ctx.v_ref + 3);
    }
    double cst, lam[3];
    coop_action_node<NC, FF, false>(P.mdl, cfg, ct, sm, j, ctx, xt, u, xh, cst, lam, nullptr, false);
    // contact forces of knots 0 / 1 as left by the last rollout that reached them (crocoddyl keeps them in the data)
    if (NC != 0 && t < 2 && valid && j == 0 && !failed) {
      if (spec) {
        lam01[t * 3 + 0] = lam[0]; lam01[t * 3 + 1] = lam[1]; lam01[t * 3 + 2] = lam[2];
      } else {
        ws.f_pred[b * 6 + t * 3 + 0] = lam[0];
        ws.f_pred[b * 6 + t * 3 + 1] = lam[1];
        ws.f_pred[b * 6 + t * 3 + 2] = lam[2];
      }
    }
    ct_sum += cst;
    bool bad = false;
    if (t < N) bad = act && !(isfinite(xh.q) && isfinite(xh.v) && isfinite(xh.t));
    const bool anybad = (__ballot_sync(kFull, bad) & omask) != 0u;
    if (!isfinite(ct_sum) || anybad) failed = true;  // "forward_error": this step length is skipped
  }
  const double dv_all = oct_sum(dv_part);  // unconditional: every lane of the warp takes part in the shuffles
  if (!valid || j != 0) return;
  const double dv = feasible ? 0.0 : dv_all;
  if (spec) {
    // speculative group: only record the trial; k_forward_finalize replays the sequential decisions
    double *tr = ws.trial + ((size_t)b * cfg.n_alphas + ia) * TRIAL_REC;
    tr[0] = failed ? 0.0 : 1.0;
    tr[1] = ct_sum;
    tr[2] = dv;
#pragma unroll
    for (int i = 0; i < 6; ++i) tr[3 + i] = lam01[i];
    return;
  }
  bool accepted = false;
  double d1 = ws.d1[b], d2 = ws.d2[b];  // keep the values of the last forwardPass that did not throw
  if (!failed) {
    d1 = dg + dv;
    d2 = dq - 2.0 * dv;
    accepted = accept_step(cfg, feasible, steplength, cost - ct_sum, d1, d2);
  }
  if (pass == 0) {
    ws.cost_valid[b] = 1;
    ws.cost[b] = cost;
  }
  ws.d1[b] = d1;
  ws.d2[b] = d2;
  if (!accepted && ia + 1 < cfg.n_alphas) {
    // rejected: queue the instance for the next pass
    const int pos = atomicAdd(&ws.qcount[(pass + 1) * 2 + cls], 1);
    ws.queue[((size_t)(((pass + 1) & 1) * 2 + cls)) * ws.B + pos] = b;
    return;
  }
  finish_line_search(ws, cfg, b, it, accepted, feasible, steplength, ct_sum, trial, d1, d2);
}

// Decisions of a speculative group (k_forward_coop with size > 1): per instance of pass `pass`, replay the sequential loop
// over the recorded trials first .. first + size - 1.  Accepted -> finish; every one rejected -> queue for the next pass,
// or give up after the last step length.  blockIdx.y = mode bucket.
static __global__ void k_forward_finalize(const __grid_constant__ Params P, int it, int pass, int first, int size) {
  const Workspace &ws = P.ws;
  const ffmpc_config &cfg = P.cfg;
  const int cls = (int)blockIdx.y;
  const int cnt = (pass == 0) ? ws.counts[cls] : ws.qcount[pass * 2 + cls];
  const int i = global_tid();
  if (i >= cnt) return;
  const int *items = (pass == 0) ? (ws.perm + (cls ? ws.counts[0] : 0)) : (ws.queue + ((size_t)((pass & 1) * 2 + cls)) * ws.B);
  const int b = items[i];
  if (pass == 0 && ws.done[b]) return;
  const int N = ws.N;
  const int cur = ws.cur[b];
  const bool feasible = ws.feas[b] != 0;
  double cost = ws.cost[b];
  if (pass == 0 && !ws.cost_valid[b]) {
    cost = 0.0;
    for (int t = 0; t <= N; ++t) cost += ws.ncost[(size_t)b * (N + 1) + t];
  }
  const double dg = ws.dg[b], dq = ws.dq[b];
  double d1 = ws.d1[b], d2 = ws.d2[b], cost_try = 0.0;
  bool accepted = false;
  int ia_acc = first + size - 1;
  for (int ia = first; ia < first + size && !accepted; ++ia) {
    const double *tr = ws.trial + ((size_t)b * cfg.n_alphas + ia) * TRIAL_REC;
    if (cls)
      for (int c = 0; c < 6; ++c)
        if (!isnan(tr[3 + c])) ws.f_pred[b * 6 + c] = tr[3 + c];
    if (tr[0] != 0.0) {
      const double steplength = ldexp(1.0, -ia);
      const double dv = feasible ? 0.0 : tr[2];
      d1 = dg + dv;
      d2 = dq - 2.0 * dv;
      if (accept_step(cfg, feasible, steplength, cost - tr[1], d1, d2)) {
        accepted = true;
        ia_acc = ia;
        cost_try = tr[1];
      }
    }
  }
  if (pass == 0) {
    ws.cost_valid[b] = 1;
    ws.cost[b] = cost;
  }
  ws.d1[b] = d1;
  ws.d2[b] = d2;
  if (!accepted && first + size < cfg.n_alphas) {
    const int pos = atomicAdd(&ws.qcount[(pass + 1) * 2 + cls], 1);
    ws.queue[((size_t)(((pass + 1) & 1) * 2 + cls)) * ws.B + pos] = b;
    return;
  }
  finish_line_search(ws, cfg, b, it, accepted, feasible, ldexp(1.0, -ia_acc), cost_try, trial_slot(ia_acc - first, cur), d1, d2);
}

// ------------------------------------------------------------------------------------------------
// parity-test hook: ActionModel.calc + calcDiff at arbitrary points.  The cooperative kernel writes the knot record
// (A.rec, zero-initialised by the host), k_node_expand turns it into the dense Crocoddyl-shaped matrices.
// ------------------------------------------------------------------------------------------------
struct NodeEvalCoopArgs {
  NodeEvalArgs a;
  double *rec;  // [M, REC_SIZE]
};
template <int NC, bool FF>
__global__ void __launch_bounds__(kCoopBlock) k_node_eval_coop(const __grid_constant__ Params P, NodeEvalCoopArgs AA) {
  const Workspace &ws = P.ws;
  const NodeEvalArgs &A = AA.a;
  constexpr int nx = FF ? 21 : 14;
  double *ct = coop_smem_base();
  double *sm = ct + CT_SIZE + (size_t)(threadIdx.x >> 3) * CS_SIZE;
  coop_fill_table(P.mdl, P.cfg, ct);
  __syncthreads();
  const int N = ws.N;
  const int g = global_tid();
  int m = g >> 3;
  const int j = g & 7;
  const int jj = j < NJ ? j : NJ - 1;
  bool valid = m < A.M;
  if (!valid) m = 0;
  const int b = A.inst[m], t = A.knot[m];
  const int nc_b = (ws.mode[b] == FFMPC_MODE_CONTACT) ? P.cfg.contact_nc : 0;
  if (nc_b != NC) valid = false;
  if (!__any_sync(kFull, valid)) return;
  CoopX x, xn;
  x.q = A.x[(size_t)m * nx + jj];
  x.v = A.x[(size_t)m * nx + NJ + jj];
  x.t = FF ? A.x[(size_t)m * nx + 2 * NJ + jj] : 0.0;
  const double u = A.u[(size_t)m * NJ + jj];
  CoopCtx ctx;
  ctx.p_ref = ws.p_ref + ((size_t)b * (N + 1) + t) * 3;
  ctx.v_ref = ws.v_ref + ((size_t)b * (N + 1) + t) * 3;
  ctx.xr_q = ws.x_reg_ref[(size_t)b * 14 + jj];
  ctx.xr_v = ws.x_reg_ref[(size_t)b * 14 + NJ + jj];
  ctx.tau_ref = ws.tau_ref[(size_t)b * NJ + jj];
  const double *x0 = ws.x0 + (size_t)b * nx;
  ctx.yr_q = x0[jj];
  ctx.yr_v = x0[NJ + jj];
  ctx.yr_t = FF ? x0[2 * NJ + jj] : 0.0;
  ctx.terminal = (t == N);
  double cost, lam[3];
  coop_action_node<NC, FF, true>(P.mdl, P.cfg, ct, sm, j, ctx, x, u, xn, cost, lam, AA.rec + (size_t)m * REC_SIZE, valid);
  if (!valid) return;
  if (j == 0) {
    if (A.cost) A.cost[m] = cost;
    if (A.force) for (int i = 0; i < 3; ++i) A.force[(size_t)m * 3 + i] = lam[i];
  }
  if (j < NJ && A.xnext) {
    A.xnext[(size_t)m * nx + j] = xn.q;
    A.xnext[(size_t)m * nx + NJ + j] = xn.v;
    if (FF) A.xnext[(size_t)m * nx + 2 * NJ + j] = xn.t;
  }
}

template <bool FF>
__global__ void k_node_expand(const __grid_constant__ Params P, NodeEvalCoopArgs AA) {
  const Workspace &ws = P.ws;
  const NodeEvalArgs &A = AA.a;
  constexpr int nx = FF ? 21 : 14;
  const int m = global_tid();
  if (m >= A.M) return;
  const int t = A.knot[m];
  const double *rec = AA.rec + (size_t)m * REC_SIZE;
  const double dt = P.cfg.dt_ocp;
  const double alpha = fmin(fmax(P.cfg.ff_alpha, 0.0), 0.999999), beta = 1.0 - alpha;
  const bool running = FF || (t != ws.N);
  // the record holds Gh = dt * d a/d z: row r of the position block of Fz is dt * Gh[r], of the velocity block Gh[r]
  auto gh = [&](int r, int j) { return rec[rec_gh(r, j / NJ, j % NJ)]; };
  for (int i = 0; i < nx; ++i) {
    for (int j = 0; j < nx; ++j) {
      double fx = (i == j) ? 1.0 : 0.0;
      if (i < 14) {
        const int r = i % NJ;
        const double sc = (i < NJ) ? dt : 1.0;
        if (running) {
          fx += sc * gh(r, j);
          if (i < NJ && j == NJ + i) fx += dt;
        }
        if (j >= 14) fx = running ? sc * gh(r, j) : 0.0;
      } else {
        fx = (i == j) ? alpha : 0.0;
      }
      if (A.Fx) A.Fx[((size_t)m * nx + i) * nx + j] = fx;
      if (A.Lxx) A.Lxx[((size_t)m * nx + i) * nx + j] = rec[rec_lzz_any(i, j)];
    }
    for (int a = 0; a < NJ; ++a) {
      double fu, lxu;
      if (FF) {
        fu = (i == 14 + a) ? beta : 0.0;
        lxu = 0.0;
      } else {
        const int r = i % NJ;
        const double sc = (i < NJ) ? dt : 1.0;
        fu = running ? sc * gh(r, 14 + a) : 0.0;
        lxu = rec[rec_lzz_any(i, 14 + a)];
      }
      if (A.Fu) A.Fu[((size_t)m * nx + i) * NJ + a] = fu;
      if (A.Lxu) A.Lxu[((size_t)m * nx + i) * NJ + a] = lxu;
    }
    if (A.Lx) A.Lx[(size_t)m * nx + i] = rec[rec_lz_any(i)];
  }
  for (int a = 0; a < NJ; ++a) {
    if (A.Lu) A.Lu[(size_t)m * NJ + a] = FF ? rec[REC_LW + a] : rec[rec_lz_any(14 + a)];
    for (int c = 0; c < NJ; ++c)
      if (A.Luu) A.Luu[((size_t)m * NJ + a) * NJ + c] = FF ? ((a == c) ? rec[REC_LWW + a] : 0.0) : rec[rec_lzz_any(14 + a, 14 + c)];
  }
}

}  // namespace ffmpc
