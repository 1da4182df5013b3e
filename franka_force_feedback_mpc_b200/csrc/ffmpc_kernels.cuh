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
  double *xs;       // [2,B,N+1,nx]  parity double buffer (accepted / trial)
  double *us;       // [2,B,N,7]
  double *fs;       // [B,N+1,nx]
  double *rec;      // [B,N+1,REC_SIZE]
  double *ncost;    // [B,N+1]
  double *K;        // [B,N,7,nx]
  double *k;        // [B,N,7]
  double *Vfs;      // [B,N+1,nx]   Vxx_t fs_t
  double *preg, *cost, *dg, *dq, *stop, *d1, *d2;
  int *feas, *wasfeas, *recalc, *done, *ok, *iter, *cur, *cost_valid;
  double *f_pred;   // [B,2,3]
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
// 7x7 helpers used redundantly by all lanes of the backward warp (uniform control flow)
// ------------------------------------------------------------------------------------------------
// Cholesky of the lower triangle of A; `mask` bit i clear -> row/col i replaced by identity (BoxQP free set).
// Returns false on a non-positive pivot (Eigen::LLT NumericalIssue).  Li = L^-1 (lower).
FFMPC_DEV bool chol7(const double A[NJ][NJ], unsigned mask, double L[NJ][NJ], double Li[NJ][NJ]) {
  bool ok = true;
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const bool fj = (mask >> j) & 1u;
    double s = fj ? A[j][j] : 1.0;
#pragma unroll
    for (int k = 0; k < j; ++k) s -= L[j][k] * L[j][k];
    if (!(s > 0.0)) ok = false;
    const double l = sqrt(s);
    const double il = 1.0 / l;
    L[j][j] = l;
    Li[j][j] = il;
#pragma unroll
    for (int i = j + 1; i < NJ; ++i) {
      const bool fi = (mask >> i) & 1u;
      double t = (fi && fj) ? A[i][j] : 0.0;
#pragma unroll
      for (int k = 0; k < j; ++k) t -= L[i][k] * L[j][k];
      L[i][j] = t * il;
    }
  }
  return ok;
}
// x = (L L^T)^-1 b with divisions replaced by the stored reciprocal diagonal
FFMPC_DEV void chol7_solve(const double L[NJ][NJ], const double Li[NJ][NJ], double *b) {
#pragma unroll
  for (int i = 0; i < NJ; ++i) {
    double s = b[i];
#pragma unroll
    for (int k = 0; k < i; ++k) s -= L[i][k] * b[k];
    b[i] = s * Li[i][i];
  }
#pragma unroll
  for (int i = NJ - 1; i >= 0; --i) {
    double s = b[i];
#pragma unroll
    for (int k = i + 1; k < NJ; ++k) s -= L[k][i] * b[k];
    b[i] = s * Li[i][i];
  }
}
// masked inverse: Hinv = (H_ff)^-1 embedded in 7x7, zero on clamped rows/cols
FFMPC_DEV bool masked_inverse7(const double H[NJ][NJ], unsigned mask, double reg, double Hinv[NJ][NJ]) {
  double A[NJ][NJ], L[NJ][NJ], Li[NJ][NJ];
#pragma unroll
  for (int i = 0; i < NJ; ++i)
#pragma unroll
    for (int j = 0; j < NJ; ++j) A[i][j] = H[i][j] + ((i == j) ? reg : 0.0);
  if (!chol7(A, mask, L, Li)) return false;
#pragma unroll
  for (int j = 0; j < NJ; ++j)
#pragma unroll
    for (int i = j + 1; i < NJ; ++i) {
      double s = 0.0;
#pragma unroll
      for (int k = j; k < i; ++k) s -= L[i][k] * Li[k][j];
      Li[i][j] = s * Li[i][i];
    }
#pragma unroll
  for (int i = 0; i < NJ; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      double s = 0.0;
#pragma unroll
      for (int k = i; k < NJ; ++k) s += Li[k][i] * Li[k][j];
      const bool f = ((mask >> i) & 1u) && ((mask >> j) & 1u);
      Hinv[i][j] = f ? s : 0.0;
      Hinv[j][i] = Hinv[i][j];
    }
  return true;
}

// crocoddyl::BoxQP::solve (projected Newton, Tassa 2014), SURVEY App. A.8.
//   x: in = warm start, out = solution;  mask out = free set;  Hinv = free-block inverse (masked)
FFMPC_DEV bool boxqp7(const ffmpc_config &cfg, const double H[NJ][NJ], const double *q, const double *lb, const double *ub,
                      double *x, unsigned &mask_out, double Hinv[NJ][NJ]) {
#pragma unroll
  for (int i = 0; i < NJ; ++i) x[i] = fmax(fmin(x[i], ub[i]), lb[i]);
  unsigned inv_mask = 0xffffffffu, mask = 0;
  bool ok = true;
  for (int it = 0; it < cfg.boxqp_maxiter; ++it) {
    double g[NJ], gnorm = 0.0;
    mask = 0;
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      double gi = q[i];
#pragma unroll
      for (int j = 0; j < NJ; ++j) gi += H[i][j] * x[j];
      g[i] = gi;
      gnorm = fmax(gnorm, fabs(gi));
      const bool clamped = (x[i] == lb[i] && gi > 0.0) || (x[i] == ub[i] && gi < 0.0);
      if (!clamped) mask |= 1u << i;
    }
    if (gnorm <= cfg.boxqp_th_grad || mask == 0) {
      if (it == 0 || mask != inv_mask) {
        if (!masked_inverse7(H, mask, cfg.boxqp_reg, Hinv)) ok = false;
        inv_mask = mask;
      }
      break;
    }
    if (!masked_inverse7(H, mask, cfg.boxqp_reg, Hinv)) { ok = false; break; }
    inv_mask = mask;
    double r[NJ], dx[NJ];
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      double s = -q[i];
#pragma unroll
      for (int j = 0; j < NJ; ++j)
        if (!((mask >> j) & 1u)) s -= H[i][j] * x[j];
      r[i] = ((mask >> i) & 1u) ? s : 0.0;
    }
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      double s = 0.0;
#pragma unroll
      for (int j = 0; j < NJ; ++j) s += Hinv[i][j] * r[j];
      dx[i] = ((mask >> i) & 1u) ? s - x[i] : 0.0;
    }
    double fold = 0.0;
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      double hy = 0.0;
#pragma unroll
      for (int j = 0; j < NJ; ++j) hy += H[i][j] * x[j];
      fold += 0.5 * x[i] * hy + q[i] * x[i];
    }
    bool changed = false;
    double steplength = 1.0;
    for (int ia = 0; ia < cfg.n_alphas; ++ia, steplength *= 0.5) {
      double xn[NJ], fnew = 0.0, gd = 0.0;
#pragma unroll
      for (int i = 0; i < NJ; ++i) xn[i] = fmax(fmin(x[i] + steplength * dx[i], ub[i]), lb[i]);
#pragma unroll
      for (int i = 0; i < NJ; ++i) {
        double hy = 0.0;
#pragma unroll
        for (int j = 0; j < NJ; ++j) hy += H[i][j] * xn[j];
        fnew += 0.5 * xn[i] * hy + q[i] * xn[i];
        gd += g[i] * (x[i] - xn[i]);
      }
      if (fold - fnew > cfg.boxqp_th_acceptstep * gd) {
#pragma unroll
        for (int i = 0; i < NJ; ++i) {
          if (xn[i] != x[i]) changed = true;
          x[i] = xn[i];
        }
        break;
      }
    }
    // iteration map is a deterministic function of x: nothing changes any more (see oracle)
    if (!changed) break;
  }
  mask_out = mask;
  return ok;
}

FFMPC_DEV double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------------------------------------
// k_backward: Riccati sweep, one warp per instance, lane c owns column c of the nx-wide matrices.
// With Fx = Pt + Et Gh (Pt: Euler/LPF structure, Et = [dt I; I; 0], Gh = dt * d a/dz) the products
// Fx^T V Fx etc. are evaluated in the factored form of DESIGN.md section "Backward kernel".
// ------------------------------------------------------------------------------------------------
template <bool FF> struct BwdLayout {
  static constexpr int n = FF ? 21 : 14;
  static constexpr int LD = FF ? 21 : 15;  // odd leading dimension: conflict-free column reads
  static constexpr int oREC = 0;
  static constexpr int oV = oREC + REC_SIZE;
  static constexpr int oVE = oV + n * LD + (n * LD) % 2;
  static constexpr int oYS = oVE + n * 8;
  static constexpr int oGT = oYS + n * 8;
  static constexpr int oQXU = oGT + n * 8;
  static constexpr int oSS = oQXU + n * 8;
  static constexpr int oQUU = oSS + 56;
  static constexpr int oVX = oQUU + 56;
  static constexpr int oFS = oVX + 24;
  static constexpr int oQU = oFS + 24;
  static constexpr int TOTAL = oQU + 8;  // doubles per warp
};

template <bool FF>
FFMPC_DEV bool backward_sweep(const Params &P, int b, double *sm, double preg, bool feasible, double &dg_out, double &dq_out) {
  using LYT = BwdLayout<FF>;
  constexpr int n = LYT::n, LD = LYT::LD;
  const Workspace &ws = P.ws;
  const ffmpc_config &cfg = P.cfg;
  const int N = ws.N;
  const int lane = (int)(threadIdx.x & 31);
  const int c = lane < n ? lane : n - 1;  // inactive lanes shadow the last column (no writes)
  const bool act = lane < n;
  const double dt = cfg.dt_ocp;
  const double alpha = fmin(fmax(cfg.ff_alpha, 0.0), 0.999999), beta = 1.0 - alpha;
  const bool box = cfg.use_box_fddp && feasible;
  double *REC = sm + LYT::oREC, *V = sm + LYT::oV, *VEs = sm + LYT::oVE, *YS = sm + LYT::oYS, *GT = sm + LYT::oGT,
         *QXU = sm + LYT::oQXU, *SS = sm + LYT::oSS, *QUU = sm + LYT::oQUU, *VX = sm + LYT::oVX, *FS = sm + LYT::oFS,
         *QU = sm + LYT::oQU;
  const double *recb = ws.rec + (size_t)b * (N + 1) * REC_SIZE;
  const double *fsb = ws.fs + (size_t)b * (N + 1) * n;
  const double *usb = ws.us + ((size_t)ws.cur[b] * ws.B + b) * N * NJ;
  double *Kb = ws.K + (size_t)b * N * NJ * n;
  double *kb = ws.k + (size_t)b * N * NJ;
  double *Vfsb = ws.Vfs + (size_t)b * (N + 1) * n;
  double dg = 0.0, dq = 0.0;

  // ---- terminal knot: Vxx = Lxx + preg, Vx = Lx (+ Vxx fs) ----
  for (int i = lane; i < REC_SIZE; i += 32) REC[i] = recb[(size_t)N * REC_SIZE + i];
  if (act) FS[lane] = fsb[N * n + lane];
  __syncwarp();
  {
    double vcol[n];
#pragma unroll
    for (int i = 0; i < n; ++i) {
      vcol[i] = REC[REC_LZZ + sym_idx_any(i, c)] + ((i == c) ? preg : 0.0);
      if (act) V[i * LD + c] = vcol[i];
    }
    double vx = REC[REC_LZ + c];
    double vf = 0.0;
#pragma unroll
    for (int i = 0; i < n; ++i) vf += vcol[i] * FS[i];
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

  for (int t = N - 1; t >= 0; --t) {
    for (int i = lane; i < REC_SIZE; i += 32) REC[i] = recb[(size_t)t * REC_SIZE + i];
    if (act) FS[lane] = fsb[t * n + lane];
    __syncwarp();
    const double *G = REC + REC_G;
    // Gh[:,c] = dt * G[:,c]
    double g[NJ];
#pragma unroll
    for (int k = 0; k < NJ; ++k) g[k] = dt * G[k * NZ + c];
    // (b) VE = V Et : row c
    double ve[NJ];
#pragma unroll
    for (int k = 0; k < NJ; ++k) {
      ve[k] = dt * V[k * LD + c] + V[(NJ + k) * LD + c];
      if (act) VEs[c * 8 + k] = ve[k];
    }
    __syncwarp();
    // (c) S = Et^T V Et
    for (int e = lane; e < 49; e += 32) {
      const int k = e / NJ, l = e - k * NJ;
      SS[k * 8 + l] = dt * VEs[k * 8 + l] + VEs[(NJ + k) * 8 + l];
    }
    __syncwarp();
    // (d) Tt = Pt^T V Et (row c), Y = Tt + Gh^T S (row c)
    double tt[NJ], y[NJ];
#pragma unroll
    for (int k = 0; k < NJ; ++k) {
      if (c < NJ) tt[k] = ve[k];
      else if (c < 2 * NJ) tt[k] = SS[(c - NJ) * 8 + k];
      else tt[k] = alpha * ve[k];
    }
#pragma unroll
    for (int l = 0; l < NJ; ++l) {
      double s = tt[l];
#pragma unroll
      for (int k = 0; k < NJ; ++k) s += g[k] * SS[k * 8 + l];
      y[l] = s;
    }
    if (act) {
#pragma unroll
      for (int k = 0; k < NJ; ++k) {
        YS[c * 8 + k] = y[k];
        GT[c * 8 + k] = g[k];
      }
    }
    // (f) Qx
    double e7[NJ];
#pragma unroll
    for (int k = 0; k < NJ; ++k) e7[k] = dt * VX[k] + VX[NJ + k];
    double qx = REC[REC_LZ + c];
    if (c < NJ) qx += VX[c];
    else if (c < 2 * NJ) qx += VX[c] + dt * VX[c - NJ];
    else qx += alpha * VX[c];
#pragma unroll
    for (int k = 0; k < NJ; ++k) qx += g[k] * e7[k];
    // (g) Qxu row c
    double qxu[NJ];
    if (FF) {
#pragma unroll
      for (int a = 0; a < NJ; ++a) {
        double pc = V[(2 * NJ + a) * LD + c];
        if (c >= NJ && c < 2 * NJ) pc += dt * V[(2 * NJ + a) * LD + c - NJ];
        if (c >= 2 * NJ) pc *= alpha;
#pragma unroll
        for (int k = 0; k < NJ; ++k) pc += g[k] * VEs[(2 * NJ + a) * 8 + k];
        qxu[a] = beta * pc;
      }
    } else {
#pragma unroll
      for (int a = 0; a < NJ; ++a) {
        double s = REC[REC_LZZ + sym_idx(c, 2 * NJ + a)];
#pragma unroll
        for (int k = 0; k < NJ; ++k) s += y[k] * (dt * G[k * NZ + 2 * NJ + a]);
        qxu[a] = s;
      }
    }
    if (act) {
#pragma unroll
      for (int a = 0; a < NJ; ++a) QXU[c * 8 + a] = qxu[a];
    }
    // (h) Quu column `lane`, Qu[lane] for lane < 7
    if (lane < NJ) {
      const int a = lane;
      if (FF) {
#pragma unroll
        for (int bb = 0; bb < NJ; ++bb)
          QUU[bb * 8 + a] = ((a == bb) ? REC[REC_LWW + a] : 0.0) + beta * beta * V[(2 * NJ + bb) * LD + 2 * NJ + a];
        QU[a] = REC[REC_LW + a] + beta * VX[2 * NJ + a];
      } else {
        double sb[NJ];
#pragma unroll
        for (int k = 0; k < NJ; ++k) {
          double s = 0.0;
#pragma unroll
          for (int m = 0; m < NJ; ++m) s += SS[k * 8 + m] * (dt * G[m * NZ + 2 * NJ + a]);
          sb[k] = s;
        }
        double qu = REC[REC_LZ + 2 * NJ + a];
#pragma unroll
        for (int bb = 0; bb < NJ; ++bb) {
          double s = REC[REC_LZZ + sym_idx_any(2 * NJ + bb, 2 * NJ + a)];
#pragma unroll
          for (int k = 0; k < NJ; ++k) s += (dt * G[k * NZ + 2 * NJ + bb]) * sb[k];
          QUU[bb * 8 + a] = s;
        }
#pragma unroll
        for (int k = 0; k < NJ; ++k) qu += (dt * G[k * NZ + 2 * NJ + a]) * e7[k];
        QU[a] = qu;
      }
    }
    __syncwarp();
    // (e) Qxx column c
    double qxx[n];
    {
      double w[n];
#pragma unroll
      for (int i = 0; i < n; ++i) {
        double wi = V[i * LD + c];
        if (c >= NJ && c < 2 * NJ) wi += dt * V[i * LD + c - NJ];
        if (FF && c >= 2 * NJ) wi *= alpha;
        w[i] = wi;
      }
#pragma unroll
      for (int i = 0; i < n; ++i) {
        double acc = w[i];
        if (i >= NJ && i < 2 * NJ) acc += dt * w[i - NJ];
        if (FF && i >= 2 * NJ) acc *= alpha;
        acc += REC[REC_LZZ + sym_idx_any(i, c)];
#pragma unroll
        for (int k = 0; k < NJ; ++k) acc += YS[i * 8 + k] * g[k] + GT[i * 8 + k] * tt[k];
        qxx[i] = acc;
      }
    }
    // (i) gains -- every lane runs the 7x7 algebra redundantly (uniform data)
    double Q[NJ][NJ], qu[NJ], kv[NJ], kc[NJ];
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
#pragma unroll
      for (int j = 0; j < NJ; ++j) Q[i][j] = QUU[i * 8 + j] + ((i == j) ? preg : 0.0);
      qu[i] = QU[i];
    }
    if (!box) {
      double L[NJ][NJ], Li[NJ][NJ];
      if (!chol7(Q, 0x7fu, L, Li)) return false;
#pragma unroll
      for (int a = 0; a < NJ; ++a) { kc[a] = qxu[a]; kv[a] = qu[a]; }
      chol7_solve(L, Li, kc);
      chol7_solve(L, Li, kv);
    } else {
      double lb[NJ], ub[NJ], xq[NJ], Hinv[NJ][NJ];
#pragma unroll
      for (int a = 0; a < NJ; ++a) {
        const double u0 = usb[t * NJ + a];
        lb[a] = -cfg.tau_limits[a] - u0;
        ub[a] = cfg.tau_limits[a] - u0;
        xq[a] = kb[t * NJ + a];  // warm start = stored k_t (crocoddyl passes k_[t])
      }
      unsigned mask;
      if (!boxqp7(cfg, Q, qu, lb, ub, xq, mask, Hinv)) return false;
#pragma unroll
      for (int a = 0; a < NJ; ++a) {
        double s = 0.0;
#pragma unroll
        for (int bb = 0; bb < NJ; ++bb) s += Hinv[a][bb] * qxu[bb];
        kc[a] = s;
        kv[a] = -xq[a];
        if (!((mask >> a) & 1u)) qu[a] = 0.0;  // clamped directions do not count for the stopping measure
      }
    }
    __syncwarp();  // all lanes have read k_t / QUU / QU before they are overwritten
    if (act) {
#pragma unroll
      for (int a = 0; a < NJ; ++a) Kb[((size_t)t * NJ + a) * n + c] = kc[a];
    }
    if (lane < NJ) kb[t * NJ + lane] = kv[lane];
    // value function
    double wv[NJ];
    double dgt = 0.0, dqt = 0.0;
#pragma unroll
    for (int a = 0; a < NJ; ++a) {
      double s = 0.0;
#pragma unroll
      for (int bb = 0; bb < NJ; ++bb) s += Q[a][bb] * kv[bb];
      wv[a] = s - 2.0 * qu[a];
      dgt += qu[a] * kv[a];
      dqt += kv[a] * s;
    }
    dg += dgt;
    dq -= dqt;
    double vx = qx;
#pragma unroll
    for (int a = 0; a < NJ; ++a) vx += kc[a] * wv[a];
    // Vxx = Qxx - Qxu K, column c
#pragma unroll
    for (int i = 0; i < n; ++i) {
      double s = qxx[i];
#pragma unroll
      for (int a = 0; a < NJ; ++a) s -= QXU[i * 8 + a] * kc[a];
      qxx[i] = s;
      if (act) V[i * LD + c] = s;
    }
    __syncwarp();
    bool bad = false;
    double vf = 0.0;
#pragma unroll
    for (int i = 0; i < n; ++i) {
      const double s = 0.5 * (qxx[i] + V[c * LD + i]) + ((i == c) ? preg : 0.0);
      qxx[i] = s;
      vf += s * FS[i];
      bad = bad || !isfinite(s);
    }
    __syncwarp();
    if (act) {
#pragma unroll
      for (int i = 0; i < n; ++i) V[i * LD + c] = qxx[i];
      Vfsb[t * n + lane] = vf;
    }
    if (!feasible) vx += vf;
    bad = bad || !isfinite(vx);
    if (act) VX[lane] = vx;
    if (!feasible) {
      dg -= warp_sum(act ? vx * FS[c] : 0.0);
      dq += warp_sum(act ? vf * FS[c] : 0.0);
    }
    if (__any_sync(0xffffffffu, bad && act)) return false;
    __syncwarp();
  }
  dg_out = dg;
  dq_out = dq;
  return true;
}

template <bool FF>
__global__ void __launch_bounds__(128) k_backward(const __grid_constant__ Params P) {
#ifdef FFMPC_EMU
  double *smem = (double *)emu::S().smem;
#else
  extern __shared__ double smem[];
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
// ------------------------------------------------------------------------------------------------
template <int NC, bool FF>
__global__ void __launch_bounds__(128) k_forward(const __grid_constant__ Params P, int it) {
  const Workspace &ws = P.ws;
  const ffmpc_config &cfg = P.cfg;
  constexpr int nx = FF ? 21 : 14;
  const int N = ws.N;
  const int slot = global_tid();
  const int nfree = ws.counts[0];
  const int cnt = (NC == 0) ? nfree : ws.counts[1];
  if (slot >= cnt) return;
  const int b = ws.perm[(NC == 0 ? 0 : nfree) + slot];
  if (ws.done[b]) return;
  const int cur = ws.cur[b];
  const double *xs = ws.xs + ((size_t)cur * ws.B + b) * (N + 1) * nx;
  const double *us = ws.us + ((size_t)cur * ws.B + b) * N * NJ;
  double *xt_all = ws.xs + ((size_t)(1 - cur) * ws.B + b) * (N + 1) * nx;
  double *ut_all = ws.us + ((size_t)(1 - cur) * ws.B + b) * N * NJ;
  const double *fs = ws.fs + (size_t)b * (N + 1) * nx;
  const double *Vfs = ws.Vfs + (size_t)b * (N + 1) * nx;
  const double *Kb = ws.K + (size_t)b * N * NJ * nx;
  const double *kb = ws.k + (size_t)b * N * NJ;
  const double *x0 = ws.x0 + (size_t)b * nx;
  const bool feasible = ws.feas[b] != 0;
  double cost = ws.cost[b];
  if (!ws.cost_valid[b]) {
    // cost_ = problem.calcDiff(xs, us): sum of the knot costs in order
    cost = 0.0;
    for (int t = 0; t <= N; ++t) cost += ws.ncost[(size_t)b * (N + 1) + t];
    ws.cost_valid[b] = 1;
  }
  const double dg = ws.dg[b], dq = ws.dq[b];
  KnotCtx ctx;
  ctx.x_reg_ref = ws.x_reg_ref + (size_t)b * 14;
  ctx.tau_ref = ws.tau_ref + (size_t)b * NJ;
  ctx.y_ref = x0;

  double steplength = 1.0, d1 = ws.d1[b], d2 = ws.d2[b], cost_try = 0.0;
  bool accepted = false;
  for (int ia = 0; ia < cfg.n_alphas; ++ia) {
    steplength = ldexp(1.0, -ia);
    const bool close = feasible || (steplength == 1.0);
    double xh[nx];
#pragma unroll
    for (int i = 0; i < nx; ++i) xh[i] = x0[i];
    cost_try = 0.0;
    double dv = 0.0;
    bool okr = true;
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
        ws.f_pred[b * 6 + t * 3 + 0] = lam[0];
        ws.f_pred[b * 6 + t * 3 + 1] = lam[1];
        ws.f_pred[b * 6 + t * 3 + 2] = lam[2];
      }
      cost_try += cst;
      bool fin = isfinite(cost_try);
      if (t < N) {
#pragma unroll
        for (int i = 0; i < nx; ++i) fin = fin && isfinite(xh[i]);
      }
      if (!fin) { okr = false; break; }
    }
    if (!okr) continue;  // "forward_error": try the next step length
    const double dV = cost - cost_try;
    d1 = dg + dv;
    d2 = dq - 2.0 * dv;
    const double dVexp = steplength * (d1 + 0.5 * steplength * d2);
    if (dVexp >= 0.0) {
      if (d1 < cfg.th_grad || dV > cfg.th_acceptstep * dVexp) accepted = true;
    } else {
      if (!feasible && dV > cfg.th_acceptnegstep * dVexp) accepted = true;
    }
    if (accepted) break;
  }
  double preg = ws.preg[b];
  int wasfeas = ws.wasfeas[b];
  if (accepted) {
    wasfeas = feasible ? 1 : 0;
    ws.wasfeas[b] = wasfeas;
    ws.feas[b] = (wasfeas || steplength == 1.0) ? 1 : 0;
    ws.cur[b] = 1 - cur;
    cost = cost_try;
    ws.recalc[b] = 1;
  } else {
    ws.recalc[b] = 0;
  }
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
