// ffmpc_kernels.cuh -- CUDA kernels of the batched FDDP / Box-FDDP solver (sm_100a, fp64).
//
// One solver iteration = three launches over the whole batch (no host synchronisation; the
// per-instance FDDP state machine -- regularisation, feasibility flags, acceptance, termination --
// lives in HBM and every kernel checks it):
//   k_node_diff : thread per (instance, knot)  ShootingProblem.calc/calcDiff -> node records, gaps
//   k_backward  : warp per instance            SolverFDDP::backwardPass + computeGains (LLT | BoxQP)
//   k_forward   : thread per instance          forwardPass + line search + acceptance + reg update
// Algorithm spec: SURVEY.md App. A.7 / A.8 (crocoddyl SolverFDDP / SolverBoxFDDP / BoxQP); the call
// being replaced is `solver.solve(xs_init, us_init, max_iters, False)`, crocoddyl_classical.py:367.
#pragma once

#include "ffmpc_node.cuh"

namespace ffmpc {

struct Workspace {
  int B, N, nx;
  int max_iters, fixed_iters;
  // inputs (device)
  const int *mode;
  const double *x0, *p_ref, *v_ref, *x_reg_ref, *tau_ref_in, *xs_init, *us_init;
  // resolved / internal
  double *tau_ref;  // [B,7]
  int *perm;        // [B] instance ids, free-mode first then contact-mode
  int *counts;      // [2] = {#free, #contact}
  int nslots;       // trajectory slots per instance = 1 (accepted) + lanes of the parallel line search
  double *xs;       // [nslots,B,N+1,nx]  slot cur[b] = accepted candidate, the others = trial buffers
  double *us;       // [nslots,B,N,7]
  double *fs;       // [B,N+1,nx]
  double *rec;      // [B,N+1,REC_SIZE]
  double *ncost;    // [B,N+1]
  double *K;        // [B,N,7,nx]
  double *k;        // [B,N,7]
  double *Vfs;      // [B,N+1,nx]   Vxx_t fs_t
  double *preg, *cost, *dg, *dq, *stop, *d1, *d2;
  int *feas, *wasfeas, *recalc, *done, *ok, *iter, *cur, *cost_valid;
  double *f_pred;   // [B,2,3]
  int *queue;       // [2 parity][2 mode buckets][B] instances still in the line search (k_forward_coop)
  int *qcount;      // [32][2] number of queued instances per step-length pass and mode bucket
  int spec;         // 1: speculative line search (all step lengths at once, small batches), trial slots 1..n_alphas
  double *trial;    // [B, n_alphas, 10] results of the speculative rollouts
};

struct Params {
  ffmpc_model mdl;
  ffmpc_config cfg;
  Workspace ws;
};

FFMPC_DEV int global_tid() { return (int)(blockIdx.x * blockDim.x + threadIdx.x); }

// ------------------------------------------------------------------------------------------------
// gravity torque g(q) = RNEA(q,0,0)  (`_gravity_torque`, crocoddyl_classical.py:447-451)
// ------------------------------------------------------------------------------------------------
FFMPC_DEV void gravity_torque(const ffmpc_model &mdl, const double *q, double *tau) {
  v3 p[NJ], z[NJ], cw[NJ];
  m3 R;
#pragma unroll
  for (int i = 0; i < NJ; ++i) {
    double s, c;
    sincos(q[i], &s, &c);
    m3 Rl;
    const double *jr = mdl.joint_rot[i];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      Rl.m[3 * r + 0] = jr[3 * r] * c + jr[3 * r + 1] * s;
      Rl.m[3 * r + 1] = -jr[3 * r] * s + jr[3 * r + 1] * c;
      Rl.m[3 * r + 2] = jr[3 * r + 2];
    }
    const v3 pl = mk3(mdl.joint_pos[i][0], mdl.joint_pos[i][1], mdl.joint_pos[i][2]);
    if (i == 0) { p[0] = pl; R = Rl; }
    else { p[i] = p[i - 1] + mulv(R, pl); R = mulm(R, Rl); }
    z[i] = mk3(R.m[2], R.m[5], R.m[8]);
    cw[i] = p[i] + mulv(R, mk3(mdl.com[i][0], mdl.com[i][1], mdl.com[i][2]));
  }
  // f_k = m_k (-g) at c_k ; tau_j = sum_{k>=j} (z_j x (c_k - p_j)) . f_k = z_j . ((c_k - p_j) x f_k)
  const v3 ng = mk3(-mdl.gravity[0], -mdl.gravity[1], -mdl.gravity[2]);
  v3 fsum = mk3(0, 0, 0), nsum = mk3(0, 0, 0);  // force and moment (about the origin) of the subtree
#pragma unroll
  for (int j = NJ - 1; j >= 0; --j) {
    const v3 f = ng * mdl.mass[j];
    fsum = fsum + f;
    nsum = nsum + cross(cw[j], f);
    tau[j] = dot(z[j], nsum - cross(p[j], fsum));
  }
}

// ------------------------------------------------------------------------------------------------
// set-up kernels
// ------------------------------------------------------------------------------------------------
// stable partition of the instances by mode; single block
static __global__ void k_build_perm(Workspace ws) {
  extern __shared__ int sh_counts[];  // [blockDim.x + 1]
#ifdef FFMPC_EMU
  int *shc = (int *)emu::S().smem;
#else
  int *shc = sh_counts;
#endif
  const int T = (int)blockDim.x, tid = (int)threadIdx.x;
  const int chunk = (ws.B + T - 1) / T;
  const int lo = tid * chunk, hi = (lo + chunk < ws.B) ? lo + chunk : ws.B;
  int nfree = 0;
  for (int b = lo; b < hi; ++b) nfree += (ws.mode[b] == FFMPC_MODE_FREE);
  shc[tid + 1] = nfree;
  __syncthreads();
  if (tid == 0) {
    shc[0] = 0;
    for (int i = 1; i <= T; ++i) shc[i] += shc[i - 1];
    ws.counts[0] = shc[T];
    ws.counts[1] = ws.B - shc[T];
  }
  __syncthreads();
  const int total_free = shc[T];
  int fpos = shc[tid];
  int cpos = total_free + (lo < ws.B ? lo - shc[tid] : 0);
  for (int b = lo; b < hi; ++b) {
    if (ws.mode[b] == FFMPC_MODE_FREE) ws.perm[fpos++] = b;
    else ws.perm[cpos++] = b;
  }
}

// per-instance solver state + tau_ref; thread per instance
static __global__ void k_init_state(const __grid_constant__ Params P) {
  const Workspace &ws = P.ws;
  const int b = global_tid();
  if (b >= ws.B) return;
  ws.preg[b] = P.cfg.reg_min;
  ws.cost[b] = 0.0;
  ws.dg[b] = ws.dq[b] = ws.stop[b] = ws.d1[b] = ws.d2[b] = 0.0;
  ws.feas[b] = 0;
  ws.wasfeas[b] = 0;
  ws.recalc[b] = 1;
  ws.done[b] = 0;
  ws.ok[b] = 0;
  ws.iter[b] = 0;
  ws.cur[b] = 0;
  ws.cost_valid[b] = 0;
  for (int i = 0; i < 6; ++i) ws.f_pred[b * 6 + i] = NAN;
  double tr[NJ];
  if (ws.tau_ref_in) {
    for (int i = 0; i < NJ; ++i) tr[i] = ws.tau_ref_in[b * NJ + i];
  } else {
    double q[NJ];
    for (int i = 0; i < NJ; ++i) q[i] = ws.x0[(size_t)b * ws.nx + i];
    gravity_torque(P.mdl, q, tr);
  }
  for (int i = 0; i < NJ; ++i) ws.tau_ref[b * NJ + i] = tr[i];
}

// warm / cold start of xs, us (`_shift_guess` output is the caller's; NULL -> cold start) and k = 0
static __global__ void k_init_traj(Workspace ws) {
  const size_t nxs = (size_t)ws.B * (ws.N + 1) * ws.nx, nus = (size_t)ws.B * ws.N * NJ;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)global_tid(); i < nxs; i += stride) {
    const size_t b = i / ((size_t)(ws.N + 1) * ws.nx);
    const int c = (int)(i % ws.nx);
    ws.xs[i] = ws.xs_init ? ws.xs_init[i] : ws.x0[b * ws.nx + c];
  }
  for (size_t i = (size_t)global_tid(); i < nus; i += stride) {
    const size_t b = i / ((size_t)ws.N * NJ);
    const int c = (int)(i % NJ);
    double v;
    if (ws.us_init) v = ws.us_init[i];
    else v = (ws.nx == 21) ? ws.x0[b * 21 + 14 + c] : ws.tau_ref[b * NJ + c];
    ws.us[i] = v;
    ws.k[i] = 0.0;
  }
}

// ------------------------------------------------------------------------------------------------
// k_node_diff: ShootingProblem::calc + calcDiff at (xs, us); gaps fs; knot costs
// ------------------------------------------------------------------------------------------------
template <int NC, bool FF>
__global__ void __launch_bounds__(128) k_node_diff(const __grid_constant__ Params P) {
  const Workspace &ws = P.ws;
  constexpr int nx = FF ? 21 : 14;
  const int N = ws.N;
  const int g = global_tid();
  const int slot = g / (N + 1), t = g - slot * (N + 1);
  const int nfree = ws.counts[0];
  const int cnt = (NC == 0) ? nfree : ws.counts[1];
  if (slot >= cnt) return;
  const int b = ws.perm[(NC == 0 ? 0 : nfree) + slot];
  if (ws.done[b] || !ws.recalc[b]) return;
  const size_t bufoff = (size_t)ws.cur[b] * ws.B;
  const double *xg = ws.xs + ((bufoff + b) * (N + 1) + t) * nx;
  double x[nx], u[NJ];
#pragma unroll
  for (int i = 0; i < nx; ++i) x[i] = xg[i];
  if (t < N) {
    const double *ug = ws.us + ((bufoff + b) * N + t) * NJ;
#pragma unroll
    for (int i = 0; i < NJ; ++i) u[i] = ug[i];
  } else {
#pragma unroll
    for (int i = 0; i < NJ; ++i) u[i] = 0.0;
  }
  KnotCtx ctx;
  ctx.p_ref = ws.p_ref + ((size_t)b * (N + 1) + t) * 3;
  ctx.v_ref = ws.v_ref + ((size_t)b * (N + 1) + t) * 3;
  ctx.x_reg_ref = ws.x_reg_ref + (size_t)b * 14;
  ctx.tau_ref = ws.tau_ref + (size_t)b * NJ;
  ctx.y_ref = ws.x0 + (size_t)b * nx;  // y_ref = y0 (crocoddyl_force_feedback.py:782)
  ctx.terminal = (t == N);
  double xnext[nx], cost, lam[3];
  double *rec = ws.rec + ((size_t)b * (N + 1) + t) * REC_SIZE;
  action_node<NC, FF, true>(P.mdl, P.cfg, ctx, x, u, xnext, &cost, lam, rec);
  ws.ncost[(size_t)b * (N + 1) + t] = cost;
  if (NC != 0 && t < 2) {
    ws.f_pred[b * 6 + t * 3 + 0] = lam[0];
    ws.f_pred[b * 6 + t * 3 + 1] = lam[1];
    ws.f_pred[b * 6 + t * 3 + 2] = lam[2];
  }
  // gaps: fs[0] = x0 - xs[0]; fs[t+1] = xnext_t - xs[t+1]
  double *fsb = ws.fs + (size_t)b * (N + 1) * nx;
  if (t == 0) {
    const double *x0 = ws.x0 + (size_t)b * nx;
#pragma unroll
    for (int i = 0; i < nx; ++i) fsb[i] = x0[i] - x[i];
  }
  if (t < N) {
    const double *xn = xg + nx;
#pragma unroll
    for (int i = 0; i < nx; ++i) fsb[(t + 1) * nx + i] = xnext[i] - xn[i];
  }
}

// ------------------------------------------------------------------------------------------------
// Warp-cooperative 7x7 algebra of the backward pass.  Lane i < 7 owns ROW i of the symmetric 7x7
// matrix (7 registers); the other lanes execute the same instructions on dummy rows (SIMT: a
// warp-wide instruction costs the same with 7 or 32 useful lanes, but 7 registers instead of 49
// per lane is what lets 16 warps live on an SM).  Every result that steers control flow is made
// warp-uniform with a broadcast from lane 0.
// ------------------------------------------------------------------------------------------------
FFMPC_DEV double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }
FFMPC_DEV double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// sum / max over lanes 0..7 (callers make lane 7 contribute the neutral element), broadcast to the warp
FFMPC_DEV double sum8(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  return shfl_d(v, 0);
}
FFMPC_DEV double max8(double v) {
  v = fmax(v, __shfl_xor_sync(0xffffffffu, v, 1));
  v = fmax(v, __shfl_xor_sync(0xffffffffu, v, 2));
  v = fmax(v, __shfl_xor_sync(0xffffffffu, v, 4));
  return shfl_d(v, 0);
}
// every lane receives the 7-vector whose component j lives on lane j
FFMPC_DEV void gather7(double mine, double *out) {
#pragma unroll
  for (int j = 0; j < NJ; ++j) out[j] = shfl_d(mine, j);
}

// Cholesky of the 7x7 SPD matrix whose row `lane` is a[0..6]; rows / columns outside `mask` are replaced
// by identity (BoxQP free set).  On exit a[j], j <= lane, is L[lane][j] and dinv = 1 / L[lane][lane].
// Returns false (uniformly) on a non-positive pivot (Eigen::LLT NumericalIssue).
FFMPC_DEV bool dist_chol7(double *a, unsigned mask, int lane, double &dinv) {
  bool ok = true;
  const bool fi = (mask >> lane) & 1u;
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const bool fj = (mask >> j) & 1u;
    if (!(fi && fj)) a[j] = (lane == j) ? 1.0 : 0.0;
  }
  dinv = 1.0;
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const double d = shfl_d(a[j], j);
    if (!(d > 0.0)) ok = false;
    const double r = rsqrt(d);
    const double lij = a[j] * r;
    if (lane == j) dinv = r;
    a[j] = lij;
#pragma unroll
    for (int k = j + 1; k < NJ; ++k) a[k] -= lij * shfl_d(lij, k);
  }
  return ok;
}

// Row `lane` (= column, symmetric) of (L L^T)^-1 restricted to the free set, from the factor stored in shared
// memory: LL[i*8+j] = L[i][j] (j <= i), DI[i] = 1 / L[i][i].  Clamped rows / columns come out as zero.
FFMPC_DEV void dist_inverse_row(const double *LL, const double *DI, unsigned mask, int lane, double *x) {
  const int c = lane < NJ ? lane : NJ - 1;
#pragma unroll
  for (int i = 0; i < NJ; ++i) {
    double s = (i == c) ? 1.0 : 0.0;
#pragma unroll
    for (int k = 0; k < i; ++k) s -= LL[i * 8 + k] * x[k];
    x[i] = s * DI[i];
  }
#pragma unroll
  for (int i = NJ - 1; i >= 0; --i) {
    double s = x[i];
#pragma unroll
    for (int k = i + 1; k < NJ; ++k) s -= LL[k * 8 + i] * x[k];
    x[i] = s * DI[i];
  }
  const bool fc = (mask >> c) & 1u;
#pragma unroll
  for (int i = 0; i < NJ; ++i)
    if (!(fc && ((mask >> i) & 1u))) x[i] = 0.0;
}

// factor + inverse: hrow = row `lane` of H (+ reg on the diagonal), hi = row `lane` of the masked inverse.
// LL / DI are 64 + 8 doubles of warp-private shared memory.
FFMPC_DEV bool dist_masked_inverse(const double *hrow, unsigned mask, double reg, int lane, double *LL, double *DI, double *hi) {
  double a[NJ], dinv;
#pragma unroll
  for (int j = 0; j < NJ; ++j) a[j] = hrow[j] + ((lane == j) ? reg : 0.0);
  const bool ok = dist_chol7(a, mask, lane, dinv);
  __syncwarp();
  if (lane < NJ) {
#pragma unroll
    for (int j = 0; j < NJ; ++j) LL[lane * 8 + j] = a[j];
    DI[lane] = dinv;
  }
  __syncwarp();
  if (!ok) return false;
  dist_inverse_row(LL, DI, mask, lane, hi);
  return true;
}

// crocoddyl::BoxQP::solve (projected Newton, Tassa 2014), SURVEY App. A.8, rows distributed over lanes 0..6.
//   hrow: row of H; q, lb, ub, x: component `lane`;  x: in = warm start, out = solution;
//   mask_out = free set;  hi = row `lane` of the free-block inverse (zero on clamped rows / columns)
FFMPC_DEV bool dist_boxqp7(const ffmpc_config &cfg, const double *hrow, double q, double lb, double ub, double &x, int lane,
                           double *LL, double *DI, unsigned &mask_out, double *hi) {
  const bool act = lane < NJ;
  x = fmax(fmin(x, ub), lb);
  unsigned inv_mask = 0xffffffffu, mask = 0;
  bool ok = true;
  for (int it = 0; it < cfg.boxqp_maxiter; ++it) {
    double xv[NJ];
    gather7(x, xv);
    double hx = 0.0;
#pragma unroll
    for (int j = 0; j < NJ; ++j) hx += hrow[j] * xv[j];
    const double g = q + hx;
    const bool clamped = (x == lb && g > 0.0) || (x == ub && g < 0.0);
    mask = __ballot_sync(0xffffffffu, act && !clamped) & 0x7fu;
    const double gnorm = max8(act ? fabs(g) : 0.0);
    if (gnorm <= cfg.boxqp_th_grad || mask == 0) {
      if (it == 0 || mask != inv_mask) {
        if (!dist_masked_inverse(hrow, mask, cfg.boxqp_reg, lane, LL, DI, hi)) ok = false;
        inv_mask = mask;
      }
      break;
    }
    if (!dist_masked_inverse(hrow, mask, cfg.boxqp_reg, lane, LL, DI, hi)) { ok = false; break; }
    inv_mask = mask;
    const bool fr = (mask >> (act ? lane : 0)) & 1u;
    // r = -q - H[:, clamped] x_clamped on the free rows
    double s = -q;
#pragma unroll
    for (int j = 0; j < NJ; ++j)
      if (!((mask >> j) & 1u)) s -= hrow[j] * xv[j];
    double rv[NJ];
    gather7((act && fr) ? s : 0.0, rv);
    double hr = 0.0;
#pragma unroll
    for (int j = 0; j < NJ; ++j) hr += hi[j] * rv[j];
    const double dx = (act && fr) ? hr - x : 0.0;
    const double fold = sum8(act ? (0.5 * x * hx + q * x) : 0.0);
    bool changed = false;
    double steplength = 1.0;
    for (int ia = 0; ia < cfg.n_alphas; ++ia, steplength *= 0.5) {
      const double xn = fmax(fmin(x + steplength * dx, ub), lb);
      double xnv[NJ];
      gather7(xn, xnv);
      double hy = 0.0;
#pragma unroll
      for (int j = 0; j < NJ; ++j) hy += hrow[j] * xnv[j];
      const double fnew = sum8(act ? (0.5 * xn * hy + q * xn) : 0.0);
      const double gd = sum8(act ? g * (x - xn) : 0.0);
      if (fold - fnew > cfg.boxqp_th_acceptstep * gd) {
        changed = __any_sync(0xffffffffu, act && xn != x) != 0;
        x = xn;
        break;
      }
    }
    // the iteration map is a deterministic function of x: nothing changes any more (see oracle)
    if (!changed) break;
  }
  mask_out = mask;
  return ok;
}

// ------------------------------------------------------------------------------------------------
// k_backward: Riccati sweep, one warp per instance, lane c owns column c of the nx-wide matrices.
// With Fx = Pt + Et Gh (Pt: Euler/LPF structure, Et = [dt I; I; 0], Gh = dt * d a/dz) the products
// Fx^T V Fx etc. are evaluated in the factored form of DESIGN.md section "Backward kernel".
// The 7x7 gain algebra (LLT / BoxQP) is row-distributed over lanes 0..6 (dist_* above).
// ------------------------------------------------------------------------------------------------
constexpr int RS = 10;  // row stride (doubles) of the 21x7 operand arrays: 16-byte aligned rows for 128-bit broadcast loads,
                        // and 80-byte stride so that a quarter-warp's 128-bit row stores hit 8 distinct bank groups
template <bool FF> struct BwdLayout {
  static constexpr int n = FF ? 21 : 14;
  static constexpr int LD = FF ? 21 : 15;  // odd leading dimension: conflict-free column AND row reads
  static constexpr int oREC = 0;
  static constexpr int oV = oREC + REC_SIZE;
  static constexpr int oVE = oV + n * LD + (n * LD) % 2;
  static constexpr int oYS = oVE + n * RS;
  static constexpr int oGT = oYS + n * RS;
  static constexpr int oQXU = oGT + n * RS;
  static constexpr int oSS = oQXU + n * RS;
  static constexpr int oLL = oVE;        // Cholesky factor rows: aliases VE (dead once Qxu is formed)
  static constexpr int oHI = oSS;        // masked inverse of Quu: aliases S (dead once Y and Quu are formed)
  static constexpr int oVX = oSS + 56;
  static constexpr int oFS = oVX + 24;
  static constexpr int oDI = oFS + 24;   // reciprocal pivots
  static constexpr int oTT = oDI + 8;    // Tt rows (Pt^T V Et), operand of the blocked Qxx assembly
  static constexpr int oKT = oTT;        // K^T rows (K[:, c]), operand of the blocked Vxx update: aliases TT (dead once Qxx is assembled)
  static constexpr int TOTAL = oKT + n * RS;  // doubles per warp (even: every warp region stays 16-byte aligned)
  static_assert(n * RS >= 56 && TOTAL % 2 == 0 && oVE % 2 == 0, "layout");
};

struct alignas(16) dbl2 { double x, y; };
// 7-vector rows in shared memory, moved as four 128-bit accesses (element 7 is padding)
FFMPC_DEV void st7(double *row, const double *v) {
  dbl2 *r = reinterpret_cast<dbl2 *>(row);
  dbl2 a; a.x = v[0]; a.y = v[1]; r[0] = a;
  a.x = v[2]; a.y = v[3]; r[1] = a;
  a.x = v[4]; a.y = v[5]; r[2] = a;
  a.x = v[6]; a.y = 0.0; r[3] = a;
}
FFMPC_DEV void ld7(const double *row, double *v) {
  const dbl2 *r = reinterpret_cast<const dbl2 *>(row);
  const dbl2 a = r[0], b2 = r[1], c2 = r[2], d = r[3];
  v[0] = a.x; v[1] = a.y; v[2] = b2.x; v[3] = b2.y; v[4] = c2.x; v[5] = c2.y; v[6] = d.x;
}
// node record of knot t -> shared memory, asynchronously (LDGSTS, 16 bytes per copy): issued right after the last
// read of the current record so that the load overlaps the gain computation and the value-function update
FFMPC_DEV void rec_prefetch(double *dst, const double *src, int lane) {
#if defined(FFMPC_EMU)
  for (int i = lane; i < REC_SIZE; i += 32) dst[i] = src[i];
#else
  const unsigned d0 = (unsigned)__cvta_generic_to_shared(dst);
  for (int i = lane; i < REC_SIZE / 2; i += 32)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + 16u * (unsigned)i), "l"(src + 2 * i) : "memory");
  asm volatile("cp.async.commit_group;" ::: "memory");
#endif
}
FFMPC_DEV void rec_wait() {
#if !defined(FFMPC_EMU)
  asm volatile("cp.async.wait_group 0;" ::: "memory");
#endif
  __syncwarp();
}

template <bool FF>
FFMPC_DEV bool backward_sweep(const Params &P, int b, double *sm, double preg, bool feasible, double &dg_out, double &dq_out) {
  using LYT = BwdLayout<FF>;
  constexpr int n = LYT::n, LD = LYT::LD;
  const Workspace &ws = P.ws;
  const ffmpc_config &cfg = P.cfg;
  const int N = ws.N;
  const int lane = (int)(threadIdx.x & 31);
  const int c = lane < n ? lane : n - 1;  // inactive lanes shadow the last column (no writes)
  const int r7 = lane < NJ ? lane : NJ - 1;
  const bool act = lane < n, act7 = lane < NJ;
  const double dt = cfg.dt_ocp;
  const double alpha = fmin(fmax(cfg.ff_alpha, 0.0), 0.999999), beta = 1.0 - alpha;
  const bool box = cfg.use_box_fddp && feasible;
  double *REC = sm + LYT::oREC, *V = sm + LYT::oV, *VEs = sm + LYT::oVE, *YS = sm + LYT::oYS, *GT = sm + LYT::oGT,
         *QXU = sm + LYT::oQXU, *SS = sm + LYT::oSS, *LL = sm + LYT::oLL, *HI = sm + LYT::oHI, *VX = sm + LYT::oVX,
         *FS = sm + LYT::oFS, *DI = sm + LYT::oDI, *TT = sm + LYT::oTT, *KT = sm + LYT::oKT;
  // 3x3 register blocks of the symmetric nx x nx matrices (Qxx, Vxx): lane l < NBLK owns block (bI, bJ), bI <= bJ, of the
  // upper triangle.  Every operand row loaded from shared memory then feeds three FMAs (the column-per-lane form fed one and
  // was bound by the shared-memory return path), and only the upper triangle is computed.
  constexpr int NB = n / 3;                 // 7 (FF) -- classical n = 14 is padded: see NB3 below
  constexpr int NB3 = (n + 2) / 3;          // block rows: 7 (n = 21), 5 (n = 14, last block is 2 wide)
  constexpr int NBLK = NB3 * (NB3 + 1) / 2; // 28 / 15
  (void)NB;
  int bI = 0, bJ = 0;
  {
    int idx = lane < NBLK ? lane : NBLK - 1, len = NB3;
    while (idx >= len) { idx -= len; ++bI; --len; }
    bJ = bI + idx;
  }
  const bool blk = lane < NBLK;
  const int i0 = 3 * bI, c0 = 3 * bJ;
  const double *recb = ws.rec + (size_t)b * (N + 1) * REC_SIZE;
  const double *fsb = ws.fs + (size_t)b * (N + 1) * n;
  const double *usb = ws.us + ((size_t)ws.cur[b] * ws.B + b) * N * NJ;
  double *Kb = ws.K + (size_t)b * N * NJ * n;
  double *kb = ws.k + (size_t)b * N * NJ;
  double *Vfsb = ws.Vfs + (size_t)b * (N + 1) * n;
  double dg = 0.0, dq = 0.0;
  // packed-upper index of Lzz(i, c): i <= c -> base_i + c (base_i compile-time), i > c -> colbase + i
  const int colbase = c * NZ - (c * (c - 1)) / 2 - c;
#define LZZ_AT(i) REC[REC_LZZ + (((i) <= c) ? ((i) * NZ - ((i) * ((i) - 1)) / 2 - (i) + c) : (colbase + (i)))]

  // ---- terminal knot: Vxx = Lxx + preg, Vx = Lx (+ Vxx fs) ----
  for (int i = lane; i < REC_SIZE; i += 32) REC[i] = recb[(size_t)N * REC_SIZE + i];
  if (act) FS[lane] = fsb[N * n + lane];
  __syncwarp();
  {
    double vf = 0.0;
#pragma unroll
    for (int i = 0; i < n; ++i) {
      const double vi = LZZ_AT(i) + ((i == c) ? preg : 0.0);
      if (act) V[i * LD + c] = vi;
      vf += vi * FS[i];
    }
    double vx = REC[REC_LZ + c];
    if (!feasible) vx += vf;
    if (act) {
      VX[lane] = vx;
      Vfsb[N * n + lane] = vf;
    }
    if (!feasible) {
      dg -= warp_sum(act ? vx * FS[c] : 0.0);
      dq += warp_sum(act ? vf * FS[c] : 0.0);
    }
  }
  __syncwarp();
  rec_prefetch(REC, recb + (size_t)(N - 1) * REC_SIZE, lane);
  if (act) FS[lane] = fsb[(N - 1) * n + lane];

  for (int t = N - 1; t >= 0; --t) {
    const double fs_next = (act && t > 0) ? fsb[(t - 1) * n + lane] : 0.0;
    // BoxQP operands of this knot (nominal control, warm start = stored k_t): issued here so that their global-memory
    // latency hides behind the Qxx assembly instead of stalling the gain computation
    const double u0_pre = box ? usb[t * NJ + r7] : 0.0;
    const double kq_pre = box ? kb[t * NJ + r7] : 0.0;
    rec_wait();
    const double *G = REC + REC_G;
    // (a) Gh[:,c] = dt * G[:,c] ; (b) VE = V Et : row c
    double g[NJ], ve[NJ];
#pragma unroll
    for (int k = 0; k < NJ; ++k) {
      g[k] = dt * G[k * NZ + c];
      ve[k] = dt * V[k * LD + c] + V[(NJ + k) * LD + c];
    }
    if (act) {
      st7(VEs + c * RS, ve);
      st7(GT + c * RS, g);
    }
    __syncwarp();
    // (c) S = Et^T V Et
    for (int e = lane; e < 49; e += 32) {
      const int k = e / NJ, l = e - k * NJ;
      SS[k * 8 + l] = dt * VEs[k * RS + l] + VEs[(NJ + k) * RS + l];
    }
    __syncwarp();
    // (d) Tt = Pt^T V Et (row c), Y = Tt + Gh^T S (row c)
    double tt[NJ];
#pragma unroll
    for (int k = 0; k < NJ; ++k) {
      if (c < NJ) tt[k] = ve[k];
      else if (c < 2 * NJ) tt[k] = SS[(c - NJ) * 8 + k];
      else tt[k] = alpha * ve[k];
    }
    {
      double y[NJ];
#pragma unroll
      for (int l = 0; l < NJ; ++l) y[l] = tt[l];
#pragma unroll
      for (int k = 0; k < NJ; ++k)
#pragma unroll
        for (int l = 0; l < NJ; ++l) y[l] += g[k] * SS[k * 8 + l];
      if (act) {
        st7(YS + c * RS, y);
        st7(TT + c * RS, tt);
      }
      // (g) Qxu row c (classical: Lxu + Y Gu)
      if (!FF) {
#pragma unroll
        for (int a = 0; a < NJ; ++a) {
          double s = REC[REC_LZZ + sym_idx(c, 2 * NJ + a)];
#pragma unroll
          for (int k = 0; k < NJ; ++k) s += y[k] * (dt * G[k * NZ + 2 * NJ + a]);
          if (act) QXU[c * RS + a] = s;
        }
      }
    }
    // (f) Qx
    double qx = REC[REC_LZ + c];
    {
      if (c < NJ) qx += VX[c];
      else if (c < 2 * NJ) qx += VX[c] + dt * VX[c - NJ];
      else qx += alpha * VX[c];
#pragma unroll
      for (int k = 0; k < NJ; ++k) qx += g[k] * (dt * VX[k] + VX[NJ + k]);
    }
    // (g) Qxu row c (FF: beta * (Fx^T V)[:, tau])
    if (FF) {
#pragma unroll
      for (int a = 0; a < NJ; ++a) {
        double pc = V[(2 * NJ + a) * LD + c];
        if (c >= NJ && c < 2 * NJ) pc += dt * V[(2 * NJ + a) * LD + c - NJ];
        if (c >= 2 * NJ) pc *= alpha;
#pragma unroll
        for (int k = 0; k < NJ; ++k) pc += g[k] * VEs[(2 * NJ + a) * RS + k];
        if (act) QXU[c * RS + a] = beta * pc;
      }
    }
    // (h) Quu row r7 (+ preg on the diagonal), Qu[r7]  -- lanes 0..6, the others shadow row 6
    double qrow[NJ], qu;
    if (FF) {
#pragma unroll
      for (int bb = 0; bb < NJ; ++bb)
        qrow[bb] = ((r7 == bb) ? REC[REC_LWW + r7] + preg : 0.0) + beta * beta * V[(2 * NJ + r7) * LD + 2 * NJ + bb];
      qu = REC[REC_LW + r7] + beta * VX[2 * NJ + r7];
    } else {
      double sb[NJ];
#pragma unroll
      for (int k = 0; k < NJ; ++k) {
        double s = 0.0;
#pragma unroll
        for (int m = 0; m < NJ; ++m) s += SS[k * 8 + m] * (dt * G[m * NZ + 2 * NJ + r7]);
        sb[k] = s;
      }
      qu = REC[REC_LZ + 2 * NJ + r7];
#pragma unroll
      for (int bb = 0; bb < NJ; ++bb) {
        double s = REC[REC_LZZ + sym_idx_any(2 * NJ + bb, 2 * NJ + r7)] + ((r7 == bb) ? preg : 0.0);
#pragma unroll
        for (int k = 0; k < NJ; ++k) s += (dt * G[k * NZ + 2 * NJ + bb]) * sb[k];
        qrow[bb] = s;
      }
#pragma unroll
      for (int k = 0; k < NJ; ++k) qu += (dt * G[k * NZ + 2 * NJ + r7]) * (dt * VX[k] + VX[NJ + k]);
    }
    __syncwarp();  // YS / GT / QXU visible to every lane
    // (e) Qxx = Pt^T V Pt + Lxx + Y Gh + Gh^T Tt^T.
    // (e1) W = Pt^T V Pt + Lxx, column c per lane (the Euler / LPF structure of Pt makes this O(nx) per column with
    //      compile-time row predicates), written IN PLACE over V: every reader of the old V is past the barrier above.
    {
      double wcol[n];
      // per-lane factors instead of per-entry selects (x 1.0 and + 0.0 * finite are exact: same bits as the branchy form)
      const bool vcol = c >= NJ && c < 2 * NJ;
      const double m_dt = vcol ? dt : 0.0, m_al = (FF && c >= 2 * NJ) ? alpha : 1.0;
      const int c2 = vcol ? c - NJ : c;
#pragma unroll
      for (int i = 0; i < n; ++i) wcol[i] = fma(m_dt, V[i * LD + c2], V[i * LD + c]) * m_al;
#pragma unroll
      for (int i = n - 1; i >= 0; --i) {
        double acc = wcol[i];
        if (i >= NJ && i < 2 * NJ) acc += dt * wcol[i - NJ];
        if (FF && i >= 2 * NJ) acc *= alpha;
        wcol[i] = acc + LZZ_AT(i);
      }
      __syncwarp();
      if (act) {
#pragma unroll
        for (int i = 0; i < n; ++i) V[i * LD + c] = wcol[i];
      }
      __syncwarp();
    }
    // (e2) upper-triangular 3x3 block (bI, bJ) per lane: W block + rank-14 update
    double qb[3][3];
    {
      const int ir[3] = {i0, (i0 + 1 < n) ? i0 + 1 : n - 1, (i0 + 2 < n) ? i0 + 2 : n - 1};
      const int cr[3] = {c0, (c0 + 1 < n) ? c0 + 1 : n - 1, (c0 + 2 < n) ? c0 + 2 : n - 1};
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int s2 = 0; s2 < 3; ++s2) qb[r][s2] = V[ir[r] * LD + cr[s2]];
#pragma unroll
      for (int k = 0; k < 8; k += 2) {
        double yI[3][2], gI[3][2], gJ[3][2], tJ[3][2];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          const dbl2 a1 = *reinterpret_cast<const dbl2 *>(YS + ir[r] * RS + k);
          const dbl2 a2 = *reinterpret_cast<const dbl2 *>(GT + ir[r] * RS + k);
          const dbl2 a3 = *reinterpret_cast<const dbl2 *>(GT + cr[r] * RS + k);
          const dbl2 a4 = *reinterpret_cast<const dbl2 *>(TT + cr[r] * RS + k);
          yI[r][0] = a1.x; yI[r][1] = a1.y;
          gI[r][0] = a2.x; gI[r][1] = a2.y;
          gJ[r][0] = a3.x; gJ[r][1] = a3.y;
          tJ[r][0] = a4.x; tJ[r][1] = a4.y;
        }
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int s2 = 0; s2 < 3; ++s2) {
            qb[r][s2] += yI[r][0] * gJ[s2][0] + gI[r][0] * tJ[s2][0];
            if (k + 1 < NJ) qb[r][s2] += yI[r][1] * gJ[s2][1] + gI[r][1] * tJ[s2][1];
          }
      }
    }
    __syncwarp();  // every lane is done with this knot's record
    if (t > 0) rec_prefetch(REC, recb + (size_t)(t - 1) * REC_SIZE, lane);
    // (i) gains.  LLT path (crocoddyl computeGains: Quu_llt.solve): factor row-distributed, then lane c solves
    //     Quu K[:,c] = Qxu[c,:]^T by substitution with the factor broadcast from shared memory; lane n does the same
    //     for Qu.  BoxQP path: K = (Quu_ff)^-1 Qxu^T with the masked inverse, k = -x*.
    double kc[NJ], kvv[NJ], kv;
    if (!box) {
      double a[NJ], dinv;
#pragma unroll
      for (int j = 0; j < NJ; ++j) a[j] = qrow[j];
      const bool okc = dist_chol7(a, 0x7fu, lane, dinv);
      if (act7) {
#pragma unroll
        for (int j = 0; j < NJ; ++j) LL[lane * 8 + j] = a[j];
        DI[lane] = dinv;
      }
      __syncwarp();
      if (!okc) { rec_wait(); return false; }
      double quv[NJ];
      gather7(qu, quv);
#pragma unroll
      for (int i = 0; i < NJ; ++i) {
        double s2 = (lane == n) ? quv[i] : QXU[c * RS + i];
#pragma unroll
        for (int k = 0; k < i; ++k) s2 -= LL[i * 8 + k] * kc[k];
        kc[i] = s2 * DI[i];
      }
#pragma unroll
      for (int i = NJ - 1; i >= 0; --i) {
        double s2 = kc[i];
#pragma unroll
        for (int k = i + 1; k < NJ; ++k) s2 -= LL[k * 8 + i] * kc[k];
        kc[i] = s2 * DI[i];
      }
#pragma unroll
      for (int i = 0; i < NJ; ++i) kvv[i] = shfl_d(kc[i], n);
      kv = kvv[0];
#pragma unroll
      for (int i = 1; i < NJ; ++i) kv = (r7 == i) ? kvv[i] : kv;
    } else {
      double hi[NJ];
      const double u0 = u0_pre;
      const double lb = -cfg.tau_limits[r7] - u0, ub = cfg.tau_limits[r7] - u0;
      double xq = kq_pre;  // warm start = stored k_t (crocoddyl passes k_[t])
      unsigned mask;
      if (!dist_boxqp7(cfg, qrow, qu, lb, ub, xq, lane, LL, DI, mask, hi)) { rec_wait(); return false; }
      kv = -xq;
      if (!((mask >> r7) & 1u)) qu = 0.0;  // clamped directions do not count for the stopping measure
      __syncwarp();
      if (act7) {
#pragma unroll
        for (int bb = 0; bb < NJ; ++bb) HI[lane * 8 + bb] = hi[bb];
      }
      __syncwarp();
#pragma unroll
      for (int a = 0; a < NJ; ++a) kc[a] = 0.0;
#pragma unroll
      for (int bb = 0; bb < NJ; ++bb) {
        const double qxb = QXU[c * RS + bb];
#pragma unroll
        for (int a = 0; a < NJ; ++a) kc[a] += HI[a * 8 + bb] * qxb;
      }
      gather7(kv, kvv);
    }
    if (act7) kb[t * NJ + lane] = kv;
    if (act) {
#pragma unroll
      for (int a = 0; a < NJ; ++a) Kb[((size_t)t * NJ + a) * n + c] = kc[a];
      st7(KT + c * RS, kc);
    }
    // value function: wv = Quu k - 2 Qu (row r7), dg += Qu.k, dq -= k.Quu k
    double wv[NJ];
    double qk = 0.0;
#pragma unroll
    for (int bb = 0; bb < NJ; ++bb) qk += qrow[bb] * kvv[bb];
    gather7(qk - 2.0 * qu, wv);
    dg += sum8(act7 ? qu * kv : 0.0);
    dq -= sum8(act7 ? kv * qk : 0.0);
    double vx = qx;
#pragma unroll
    for (int a = 0; a < NJ; ++a) vx += kc[a] * wv[a];
    // Vxx = Qxx - Qxu K (+ preg), symmetric: upper 3x3 block per lane, mirrored into shared memory
    __syncwarp();  // KT visible; every reader of the old V is done
    bool bad = false;
    {
      const int ir[3] = {i0, (i0 + 1 < n) ? i0 + 1 : n - 1, (i0 + 2 < n) ? i0 + 2 : n - 1};
      const int cr[3] = {c0, (c0 + 1 < n) ? c0 + 1 : n - 1, (c0 + 2 < n) ? c0 + 2 : n - 1};
#pragma unroll
      for (int k = 0; k < 8; k += 2) {
        double qI[3][2], kJ[3][2];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          const dbl2 a1 = *reinterpret_cast<const dbl2 *>(QXU + ir[r] * RS + k);
          const dbl2 a2 = *reinterpret_cast<const dbl2 *>(KT + cr[r] * RS + k);
          qI[r][0] = a1.x; qI[r][1] = a1.y;
          kJ[r][0] = a2.x; kJ[r][1] = a2.y;
        }
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int s2 = 0; s2 < 3; ++s2) {
            qb[r][s2] -= qI[r][0] * kJ[s2][0];
            if (k + 1 < NJ) qb[r][s2] -= qI[r][1] * kJ[s2][1];
          }
      }
      if (blk) {
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int s2 = 0; s2 < 3; ++s2) {
            const int i = i0 + r, cc = c0 + s2;
            if (i < n && cc < n && i <= cc) {
              const double v = qb[r][s2] + ((i == cc) ? preg : 0.0);
              bad = bad || !isfinite(v);
              V[i * LD + cc] = v;
              V[cc * LD + i] = v;
            }
          }
      }
    }
    __syncwarp();
    double vf = 0.0;
#pragma unroll
    for (int i = 0; i < n; ++i) vf += V[i * LD + c] * FS[i];
    if (act) Vfsb[t * n + lane] = vf;
    if (!feasible) vx += vf;
    bad = bad || (act && !isfinite(vx));
    if (act) VX[lane] = vx;
    if (!feasible) {
      dg -= warp_sum(act ? vx * FS[c] : 0.0);
      dq += warp_sum(act ? vf * FS[c] : 0.0);
    }
    if (__any_sync(0xffffffffu, bad)) {
      rec_wait();
      return false;
    }
    __syncwarp();
    if (act) FS[lane] = fs_next;
  }
#undef LZZ_AT
  dg_out = dg;
  dq_out = dq;
  return true;
}

// Residency of the Riccati kernel: FFMPC_BWD_WARPS warps (= instances) per block, FFMPC_BWD_MINBLOCKS blocks per SM.
// Shared memory (16.1 KB per FF warp) allows 14 warps per SM, the register file then allows 144 registers per thread.
#ifndef FFMPC_BWD_WARPS
#define FFMPC_BWD_WARPS 4
#endif
#ifndef FFMPC_BWD_MINBLOCKS
#define FFMPC_BWD_MINBLOCKS 3
#endif
template <bool FF>
__global__ void __launch_bounds__(FFMPC_BWD_WARPS * 32, FFMPC_BWD_MINBLOCKS) k_backward(const __grid_constant__ Params P) {
#ifdef FFMPC_EMU
  double *smem = (double *)emu::S().smem;
#else
  extern __shared__ __align__(16) double smem[];
#endif
  const Workspace &ws = P.ws;
  const int warp = (int)(threadIdx.x >> 5), lane = (int)(threadIdx.x & 31);
  const int b = (int)(blockIdx.x * (blockDim.x >> 5)) + warp;
  if (b >= ws.B) return;
  if (ws.done[b]) return;
  double *sm = smem + (size_t)warp * BwdLayout<FF>::TOTAL;
  double preg = ws.preg[b];
  const bool feasible = ws.feas[b] != 0;
  const int it = ws.iter[b];
  // computeDirection with the regularisation-escalation loop of SolverFDDP::solve
  while (true) {
    double dg, dq;
    const bool okb = backward_sweep<FF>(P, b, sm, preg, feasible, dg, dq);
    __syncwarp();
    if (okb) {
      if (lane == 0) {
        ws.dg[b] = dg;
        ws.dq[b] = dq;
        ws.preg[b] = preg;
      }
      break;
    }
    preg *= P.cfg.reg_incfactor;
    if (preg > P.cfg.reg_max) preg = P.cfg.reg_max;
    if (preg == P.cfg.reg_max) {
      if (lane == 0) {
        ws.preg[b] = preg;
        ws.done[b] = 1;
        ws.ok[b] = 0;
        ws.iter[b] = it;
      }
      break;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// k_forward: SolverFDDP::forwardPass / tryStep / expectedImprovement + acceptance + reg update
//
// Parallel line search: instance b is served by A = 2^a adjacent lanes of one warp; lane j rolls out step
// length alpha = 2^-(r A + j) in round r into its OWN trial buffer, then the group elects the FIRST (largest)
// accepted alpha -- the same decision the sequential loop of SolverFDDP::solve takes, because a trial never
// depends on another trial.  xs / us live in (A_max + 1) trajectory slots per instance; `cur[b]` names the
// accepted one and lane j writes slot j + (j >= cur).
// ------------------------------------------------------------------------------------------------
template <int NC, bool FF>
FFMPC_DEV bool rollout(const Params &P, int b, int cur, int slot, double steplength, bool feasible, double &cost_try, double &dv_out,
                       double *lam01) {
  const Workspace &ws = P.ws;
  const ffmpc_config &cfg = P.cfg;
  constexpr int nx = FF ? 21 : 14;
  const int N = ws.N;
  const double *xs = ws.xs + ((size_t)cur * ws.B + b) * (N + 1) * nx;
  const double *us = ws.us + ((size_t)cur * ws.B + b) * N * NJ;
  double *xt_all = ws.xs + ((size_t)slot * ws.B + b) * (N + 1) * nx;
  double *ut_all = ws.us + ((size_t)slot * ws.B + b) * N * NJ;
  const double *fs = ws.fs + (size_t)b * (N + 1) * nx;
  const double *Vfs = ws.Vfs + (size_t)b * (N + 1) * nx;
  const double *Kb = ws.K + (size_t)b * N * NJ * nx;
  const double *kb = ws.k + (size_t)b * N * NJ;
  const double *x0 = ws.x0 + (size_t)b * nx;
  KnotCtx ctx;
  ctx.x_reg_ref = ws.x_reg_ref + (size_t)b * 14;
  ctx.tau_ref = ws.tau_ref + (size_t)b * NJ;
  ctx.y_ref = x0;
  const bool close = feasible || (steplength == 1.0);
  double xh[nx];
#pragma unroll
  for (int i = 0; i < nx; ++i) xh[i] = x0[i];
  cost_try = 0.0;
  double dv = 0.0;
  for (int t = 0; t <= N; ++t) {
    double xt[nx], dx[nx];
    double dvt = 0.0;
#pragma unroll
    for (int i = 0; i < nx; ++i) {
      xt[i] = close ? xh[i] : xh[i] + fs[t * nx + i] * (steplength - 1.0);
      xt_all[t * nx + i] = xt[i];
      dx[i] = xt[i] - xs[t * nx + i];
      dvt += Vfs[t * nx + i] * dx[i];  // = -fs_t^T Vxx_t (xs_t - xs_try_t)
    }
    if (!feasible) dv += dvt;
    double u[NJ];
    if (t < N) {
#pragma unroll
      for (int a = 0; a < NJ; ++a) {
        double s = us[t * NJ + a] - kb[t * NJ + a] * steplength;
        double kd = 0.0;
        const double *Krow = Kb + ((size_t)t * NJ + a) * nx;
#pragma unroll
        for (int j = 0; j < nx; ++j) kd += Krow[j] * dx[j];
        s -= kd;
        if (cfg.use_box_fddp) s = fmin(fmax(s, -cfg.tau_limits[a]), cfg.tau_limits[a]);
        u[a] = s;
        ut_all[t * NJ + a] = s;
      }
    } else {
#pragma unroll
      for (int a = 0; a < NJ; ++a) u[a] = 0.0;
    }
    ctx.p_ref = ws.p_ref + ((size_t)b * (N + 1) + t) * 3;
    ctx.v_ref = ws.v_ref + ((size_t)b * (N + 1) + t) * 3;
    ctx.terminal = (t == N);
    double cst, lam[3];
    action_node<NC, FF, false>(P.mdl, cfg, ctx, xt, u, xh, &cst, lam, nullptr);
    if (NC != 0 && t < 2) {
      lam01[t * 3 + 0] = lam[0];
      lam01[t * 3 + 1] = lam[1];
      lam01[t * 3 + 2] = lam[2];
    }
    cost_try += cst;
    bool fin = isfinite(cost_try);
    if (t < N) {
#pragma unroll
      for (int i = 0; i < nx; ++i) fin = fin && isfinite(xh[i]);
    }
    if (!fin) return false;  // "forward_error": this step length is skipped
  }
  dv_out = dv;
  return true;
}

template <int NC, bool FF>
__global__ void __launch_bounds__(128) k_forward(const __grid_constant__ Params P, int it, int lgA) {
  const Workspace &ws = P.ws;
  const ffmpc_config &cfg = P.cfg;
  const int N = ws.N;
  const int A = 1 << lgA;
  const int g = global_tid();
  const int lane = (int)(threadIdx.x & 31);
  const int slot_id = g >> lgA, j = g & (A - 1);
  const int gbase = lane & ~(A - 1);
  const unsigned gmask = (A == 32) ? 0xffffffffu : (((1u << A) - 1u) << gbase);
  const int nfree = ws.counts[0];
  const int cnt = (NC == 0) ? nfree : ws.counts[1];
  bool valid = slot_id < cnt;
  const int b = valid ? ws.perm[(NC == 0 ? 0 : nfree) + slot_id] : 0;
  if (valid && ws.done[b]) valid = false;
  if (!__any_sync(0xffffffffu, valid)) return;
  const int cur = valid ? ws.cur[b] : 0;
  const bool feasible = valid ? (ws.feas[b] != 0) : false;
  double cost = 0.0, dg = 0.0, dq = 0.0;
  if (valid) {
    cost = ws.cost[b];
    if (!ws.cost_valid[b]) {
      // cost_ = problem.calcDiff(xs, us): sum of the knot costs in order
      cost = 0.0;
      for (int t = 0; t <= N; ++t) cost += ws.ncost[(size_t)b * (N + 1) + t];
    }
    dg = ws.dg[b];
    dq = ws.dq[b];
  }
  const int myslot = j + (j >= cur ? 1 : 0);
  double d1 = 0.0, d2 = 0.0, cost_try = 0.0;   // of this lane's latest successful rollout
  double lam01[6] = {NAN, NAN, NAN, NAN, NAN, NAN};
  int ia_ok = -1;                               // step index of this lane's latest successful rollout
  int ia_acc = -1;                              // accepted step index of the group
  int win = -1;                                 // lane (within the warp) that holds the accepted trial
  const int rounds = (cfg.n_alphas + A - 1) >> lgA;
  for (int r = 0; r < rounds; ++r) {
    const int ia = (r << lgA) + j;
    const bool has = valid && ia_acc < 0 && ia < cfg.n_alphas;
    bool acc = false;
    if (has) {
      const double steplength = ldexp(1.0, -ia);
      double ct, dv;
      if (rollout<NC, FF>(P, b, cur, myslot, steplength, feasible, ct, dv, lam01)) {
        cost_try = ct;
        ia_ok = ia;
        const double dV = cost - ct;
        d1 = dg + dv;
        d2 = dq - 2.0 * dv;
        const double dVexp = steplength * (d1 + 0.5 * steplength * d2);
        if (dVexp >= 0.0) acc = (d1 < cfg.th_grad || dV > cfg.th_acceptstep * dVexp);
        else acc = (!feasible && dV > cfg.th_acceptnegstep * dVexp);
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, acc) & gmask;
    if (bal != 0u && ia_acc < 0) {
      win = __ffs((int)bal) - 1;
      ia_acc = (r << lgA) + (win - gbase);
    }
    if (!__any_sync(0xffffffffu, valid && ia_acc < 0)) break;
  }
  // the lane that speaks for the instance: the accepted trial, else the LAST successful trial (d1, d2 keep the
  // values of the last forwardPass that did not throw), else lane 0 of the group with the stored d1, d2
  int best = ia_ok;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int other = __shfl_xor_sync(0xffffffffu, best, o);
    if (o < A && other > best) best = other;
  }
  if (!valid) return;
  const bool accepted = ia_acc >= 0;
  bool writer;
  if (accepted) writer = (lane == win);
  else if (best >= 0) writer = (ia_ok == best);
  else writer = (j == 0);
  // contact forces left by the last rollout that ran: the accepted one, else step n_alphas - 1
  if (NC != 0) {
    const int ia_last = accepted ? ia_acc : cfg.n_alphas - 1;
    const bool mine = accepted ? (lane == win) : (ia_last - ((ia_last >> lgA) << lgA) == j);
    if (mine) {
#pragma unroll
      for (int i = 0; i < 6; ++i)
        if (!isnan(lam01[i]) || accepted) ws.f_pred[b * 6 + i] = lam01[i];
    }
  }
  if (!writer) return;
  if (!accepted && best < 0) {
    d1 = ws.d1[b];
    d2 = ws.d2[b];
  }
  const double steplength = accepted ? ldexp(1.0, -ia_acc) : ldexp(1.0, -(cfg.n_alphas - 1));
  double preg = ws.preg[b];
  int wasfeas = ws.wasfeas[b];
  if (accepted) {
    wasfeas = feasible ? 1 : 0;
    ws.wasfeas[b] = wasfeas;
    ws.feas[b] = (wasfeas || steplength == 1.0) ? 1 : 0;
    ws.cur[b] = myslot;
    cost = cost_try;
    ws.recalc[b] = 1;
  } else {
    ws.recalc[b] = 0;
  }
  ws.cost_valid[b] = 1;
  ws.cost[b] = cost;
  ws.d1[b] = d1;
  ws.d2[b] = d2;
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

// ------------------------------------------------------------------------------------------------
// export: gather results into the caller's row-major buffers
// ------------------------------------------------------------------------------------------------
struct ExportPtrs {
  double *xs, *us, *K, *k, *cost, *stop, *gaps, *f_pred;
  int *iters, *ok;
};
static __global__ void k_export(Workspace ws, ExportPtrs o) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t g0 = (size_t)global_tid();
  const int N = ws.N, nx = ws.nx;
  const size_t nxs = (size_t)ws.B * (N + 1) * nx, nus = (size_t)ws.B * N * NJ, nK = nus * nx;
  if (o.xs)
    for (size_t i = g0; i < nxs; i += stride) {
      const size_t b = i / ((size_t)(N + 1) * nx);
      o.xs[i] = ws.xs[(size_t)ws.cur[b] * nxs + i];
    }
  if (o.us)
    for (size_t i = g0; i < nus; i += stride) {
      const size_t b = i / ((size_t)N * NJ);
      o.us[i] = ws.us[(size_t)ws.cur[b] * nus + i];
    }
  if (o.K)
    for (size_t i = g0; i < nK; i += stride) o.K[i] = ws.K[i];
  if (o.k)
    for (size_t i = g0; i < nus; i += stride) o.k[i] = ws.k[i];
  if (o.gaps)
    for (size_t i = g0; i < nxs; i += stride) o.gaps[i] = ws.fs[i];
  for (size_t b = g0; b < (size_t)ws.B; b += stride) {
    if (o.cost) o.cost[b] = ws.cost[b];
    if (o.stop) o.stop[b] = ws.stop[b];
    if (o.iters) o.iters[b] = ws.iter[b];
    if (o.ok) o.ok[b] = ws.ok[b];
    if (o.f_pred)
      for (int i = 0; i < 6; ++i) o.f_pred[b * 6 + i] = ws.f_pred[b * 6 + i];
  }
}

// ------------------------------------------------------------------------------------------------
// k_node_eval: ActionModel.calc + calcDiff at arbitrary points, dense outputs (parity-test hook)
// ------------------------------------------------------------------------------------------------
struct NodeEvalArgs {
  int M;
  const int *inst, *knot;
  const double *x, *u;
  double *xnext, *cost, *force, *Fx, *Fu, *Lx, *Lu, *Lxx, *Lxu, *Luu;
};
template <int NC, bool FF>
__global__ void __launch_bounds__(128) k_node_eval(const __grid_constant__ Params P, NodeEvalArgs A) {
  const Workspace &ws = P.ws;
  constexpr int nx = FF ? 21 : 14;
  const int N = ws.N;
  const int m = global_tid();
  if (m >= A.M) return;
  const int b = A.inst[m], t = A.knot[m];
  const int nc_b = (ws.mode[b] == FFMPC_MODE_CONTACT) ? P.cfg.contact_nc : 0;
  if (nc_b != NC) return;
  double x[nx], u[NJ];
  for (int i = 0; i < nx; ++i) x[i] = A.x[(size_t)m * nx + i];
  for (int i = 0; i < NJ; ++i) u[i] = A.u[(size_t)m * NJ + i];
  KnotCtx ctx;
  ctx.p_ref = ws.p_ref + ((size_t)b * (N + 1) + t) * 3;
  ctx.v_ref = ws.v_ref + ((size_t)b * (N + 1) + t) * 3;
  ctx.x_reg_ref = ws.x_reg_ref + (size_t)b * 14;
  ctx.tau_ref = ws.tau_ref + (size_t)b * NJ;
  ctx.y_ref = ws.x0 + (size_t)b * nx;
  ctx.terminal = (t == N);
  double xnext[nx], cost, lam[3];
  double rec[REC_SIZE];
  for (int i = 0; i < REC_SIZE; ++i) rec[i] = 0.0;
  action_node<NC, FF, true>(P.mdl, P.cfg, ctx, x, u, xnext, &cost, lam, rec);
  const double dt = P.cfg.dt_ocp;
  const double alpha = fmin(fmax(P.cfg.ff_alpha, 0.0), 0.999999), beta = 1.0 - alpha;
  const bool running = FF || !ctx.terminal;
  if (A.xnext) for (int i = 0; i < nx; ++i) A.xnext[(size_t)m * nx + i] = xnext[i];
  if (A.cost) A.cost[m] = cost;
  if (A.force) for (int i = 0; i < 3; ++i) A.force[(size_t)m * 3 + i] = lam[i];
  // dense Fz (14 x 21): Euler structure around G
  for (int i = 0; i < nx; ++i) {
    for (int j = 0; j < nx; ++j) {
      double fx = (i == j) ? 1.0 : 0.0;
      if (i < 14) {
        const int r = i % NJ;
        const double sc = (i < NJ) ? dt * dt : dt;
        if (running) {
          fx += sc * rec[REC_G + r * NZ + j];
          if (i < NJ && j == NJ + i) fx += dt;
        }
        if (j >= 14) fx = running ? sc * rec[REC_G + r * NZ + j] : 0.0;
      } else {
        fx = (i == j) ? alpha : 0.0;
      }
      if (A.Fx) A.Fx[((size_t)m * nx + i) * nx + j] = fx;
      if (A.Lxx) A.Lxx[((size_t)m * nx + i) * nx + j] = rec[REC_LZZ + sym_idx_any(i, j)];
    }
    for (int a = 0; a < NJ; ++a) {
      double fu, lxu;
      if (FF) {
        fu = (i == 14 + a) ? beta : 0.0;
        lxu = 0.0;
      } else {
        const int r = i % NJ;
        const double sc = (i < NJ) ? dt * dt : dt;
        fu = running ? sc * rec[REC_G + r * NZ + 14 + a] : 0.0;
        lxu = rec[REC_LZZ + sym_idx_any(i, 14 + a)];
      }
      if (A.Fu) A.Fu[((size_t)m * nx + i) * NJ + a] = fu;
      if (A.Lxu) A.Lxu[((size_t)m * nx + i) * NJ + a] = lxu;
    }
    if (A.Lx) A.Lx[(size_t)m * nx + i] = rec[REC_LZ + i];
  }
  for (int a = 0; a < NJ; ++a) {
    if (A.Lu) A.Lu[(size_t)m * NJ + a] = FF ? rec[REC_LW + a] : rec[REC_LZ + 14 + a];
    for (int c = 0; c < NJ; ++c)
      if (A.Luu) A.Luu[((size_t)m * NJ + a) * NJ + c] = FF ? ((a == c) ? rec[REC_LWW + a] : 0.0) : rec[REC_LZZ + sym_idx_any(14 + a, 14 + c)];
  }
}

}  // namespace ffmpc
