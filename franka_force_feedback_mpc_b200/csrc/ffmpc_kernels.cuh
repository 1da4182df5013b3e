// ffmpc_kernels.cuh -- workspace of the batched FDDP / Box-FDDP solver and its HBM-trivial kernels (sm_100a, fp64).
//
// One solver iteration = a fixed launch sequence over the whole batch (no host synchronisation; the
// per-instance FDDP state machine -- regularisation, feasibility flags, acceptance, termination --
// lives in HBM and every kernel checks it):
//   k_node_diff_coop : 8 lanes per (instance, knot)  ShootingProblem.calc/calcDiff -> knot records, gaps   (ffmpc_coop_kernels.cuh)
//   k_backward       : warp per instance             SolverFDDP::backwardPass + computeGains (LLT | BoxQP) (ffmpc_riccati.cuh)
//   k_forward_coop   : 8 lanes per instance          forwardPass + line search + acceptance + reg update    (ffmpc_coop_kernels.cuh)
// Algorithm spec: SURVEY.md App. A.7 / A.8 (crocoddyl SolverFDDP / SolverBoxFDDP / BoxQP); the call
// being replaced is `solver.solve(xs_init, us_init, max_iters, False)`, crocoddyl_classical.py:367.
#pragma once

#include "ffmpc_math.cuh"

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
  int *qcount;      // [32][2] number of queued instances per line-search pass and mode bucket
  double *trial;    // [B, n_alphas, 10] results of the speculative rollouts
  // line-search schedule (ffmpc_engine.cu: set_schedule): pass p tries the step lengths 2^-ia, ia = gfirst[p] .. gfirst[p] + gsize[p] - 1;
  // groups of more than one are rolled out speculatively (needs gsize <= nslots - 1 trial slots)
  int ls_mode, n_alphas, ngroups;
  signed char gfirst[32], gsize[32];
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

// tau_ref of instance b: the caller's, or g(q0) = RNEA(q0, 0, 0) (`_compute_tau_reference` "gravity_x0", crocoddyl_classical.py:447-460)
FFMPC_DEV void resolve_tau_ref(const ffmpc_model &mdl, const Workspace &ws, int b) {
  double tr[NJ];
  if (ws.tau_ref_in) {
    for (int i = 0; i < NJ; ++i) tr[i] = ws.tau_ref_in[b * NJ + i];
  } else {
    double q[NJ];
    for (int i = 0; i < NJ; ++i) q[i] = ws.x0[(size_t)b * ws.nx + i];
    gravity_torque(mdl, q, tr);
  }
  for (int i = 0; i < NJ; ++i) ws.tau_ref[b * NJ + i] = tr[i];
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
  resolve_tau_ref(P.mdl, ws, b);
}

// tau_ref alone (the knot-evaluation test hook needs nothing else of the per-instance state)
static __global__ void k_resolve_tau_ref(const __grid_constant__ Params P) {
  const int b = global_tid();
  if (b < P.ws.B) resolve_tau_ref(P.mdl, P.ws, b);
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

// parity-test hook (k_node_eval_coop / k_node_expand): evaluation points and dense outputs
struct NodeEvalArgs {
  int M;
  const int *inst, *knot;
  const double *x, *u;
  double *xnext, *cost, *force, *Fx, *Fu, *Lx, *Lu, *Lxx, *Lxu, *Luu;
};

}  // namespace ffmpc
