// ffmpc_riccati.cuh -- k_backward: SolverFDDP::backwardPass + computeGains (LLT | BoxQP) + the regularisation-
// escalation loop of SolverFDDP::solve, one WARP per instance, on the fp64 tensor-core MMA (mma.sync m8n8k4, "DMMA").
// Replaces what `solver.solve(...)` (crocoddyl_classical.py:367, crocoddyl_force_feedback.py:605) spends in
// backwardPass; algorithm: SURVEY.md App. A.7 / A.8, restated by oracle/ffmpc_oracle.cpp Solver::backward_pass.
//
// Why DMMA.  Measured on B200 (tools/dmma_probe.cu, profiles/r2a_dmma_probe.txt): one m8n8k4 DMMA (256 FMA) issues
// every 16 cycles per SM sub-partition = 37.1 TFLOP/s, the same pipe and the same peak as DFMA (2.2 cycles per
// 32-FMA instruction), with 27 cycles of latency.  The round-1 sweep (3x3 register blocks, operands exchanged through
// shared memory) was bound by shared-memory wavefronts (70 % of peak) and issue slots, not by that pipe; the MMA
// moves 8x more FMAs per instruction and does the operand broadcast inside the tensor core.
//
// Formulation.  z = (q, v, u_inner) in three blocks of 7 padded to 8; every matrix of the sweep is a grid of 8x8
// tiles held in the MMA accumulator layout ("C layout": lane 4g+t holds M[g][2t], M[g][2t+1]); the operand layout
// ("fragment": lane holds M[g][t], M[g][4+t]) is the A operand of M and the B operand of M^T, so every product is
// written D += X Y^T with two fragments.  With Fx = Pt + Et Gh (Pt: Euler / LPF structure, Et = [dt I; I; 0],
// Gh = dt da/dz from the knot record):
//     VE = V Et (tile-wise axpy), S = Et^T V Et, Tt = Pt^T V Et, W = Pt^T V Pt     -- elementwise on tiles
//     Z  = Tt + 1/2 Gh^T S                                                         -- 6 DMMA
//     Q  = Lzz + W + Z Gh + Gh^T Z^T  (six upper tiles)                            -- 24 DMMA
//     FF: Qxu = beta (Pt^T V)[:, tau] + Gh^T (Et^T V Fu), Quu = beta^2 V_tt + Lww  -- 6 DMMA   (classical: tiles of Q)
//     K^T = Qxu Quu^-1  (masked inverse of the BoxQP free set, or the full inverse) -- 6 DMMA
//     Vxx = Qxx - Qxu K (upper tiles), lower tiles by a shuffle transpose           -- 12 DMMA
// The gradient rides along in homogeneous coordinates: row 7 of the tile row 0 (the padding row of the q block) holds
// Vx^T / Lx^T / Qx^T, so Qx, Qu and Vx = Qx + K^T (Quu k - 2 Qu) come out of the same MMAs.  The 7x7 gain algebra
// (Cholesky, BoxQP) stays row-distributed over lanes 0..6 (dist_* below).  Knot records are streamed by one
// cp.async.bulk per knot into a two-stage shared-memory ring, completion on an mbarrier.
#pragma once

#include "ffmpc_kernels.cuh"

namespace ffmpc {

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
// 8x8 tiles on the fp64 MMA
// ------------------------------------------------------------------------------------------------
struct T2 { double c0, c1; };  // C layout: lane 4g+t holds M[g][2t], M[g][2t+1]
struct F2 { double k0, k1; };  // fragment: lane 4g+t holds M[g][t], M[g][4+t]
FFMPC_DEV T2 mkt(double a, double b) { T2 r; r.c0 = a; r.c1 = b; return r; }
FFMPC_DEV T2 tadd(const T2 &x, const T2 &y) { return mkt(x.c0 + y.c0, x.c1 + y.c1); }
FFMPC_DEV T2 tscale(double s, const T2 &x) { return mkt(s * x.c0, s * x.c1); }
FFMPC_DEV T2 taxpy(double s, const T2 &x, const T2 &y) { return mkt(fma(s, x.c0, y.c0), fma(s, x.c1, y.c1)); }  // s x + y

// D (8x8) += A (8x4) B (4x8): lane holds A[g][t], B[t][g], D[g][2t], D[g][2t+1]
FFMPC_DEV void dmma(T2 &d, double a, double b) {
#ifdef FFMPC_EMU
  const int lane = (int)(threadIdx.x & 31), g = lane >> 2, t = lane & 3;
  double A[32], B[32];
  emu::allgather2(a, b, A, B);
  for (int k = 0; k < 4; ++k) {
    d.c0 = fma(A[4 * g + k], B[4 * (2 * t) + k], d.c0);
    d.c1 = fma(A[4 * g + k], B[4 * (2 * t + 1) + k], d.c1);
  }
#else
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d.c0), "+d"(d.c1) : "d"(a), "d"(b));
#endif
}
// D += X Y^T for 8x8 tiles given as fragments (two k-steps)
FFMPC_DEV void mma88(T2 &d, const F2 &x, const F2 &y) {
  dmma(d, x.k0, y.k0);
  dmma(d, x.k1, y.k1);
}
// C layout -> fragment, inside each group of four lanes: two 64-bit shuffles.  Element j of row g lives on lane
// t' = j >> 1, register j & 1.  Shuffle 1 serves (receiver t even, k0) and (t odd, k1), shuffle 2 the other two, so
// that every source lane exposes exactly one register per shuffle.
FFMPC_DEV F2 frag_of(const T2 &m, int lane) {
  const int t = lane & 3, base = lane & ~3, odd = t & 1, h = t >> 1;
  const double x1 = (t < 2) ? m.c0 : m.c1;
  const double x2 = (t < 2) ? m.c1 : m.c0;
  const double r1 = __shfl_sync(0xffffffffu, x1, base + (odd ? 2 + h : h));
  const double r2 = __shfl_sync(0xffffffffu, x2, base + (odd ? h : 2 + h));
  F2 f;
  f.k0 = odd ? r2 : r1;
  f.k1 = odd ? r1 : r2;
  return f;
}
// transpose in C layout: slot e of lane (g, t) = M[2t+e][g] = register g & 1 of lane (2t+e, g >> 1); two shuffles
FFMPC_DEV T2 transpose_t(const T2 &m, int lane) {
  const int g = lane >> 2, t = lane & 3, odd = g & 1, h = g >> 1;
  const double x1 = odd ? m.c1 : m.c0;
  const double x2 = odd ? m.c0 : m.c1;
  const double r1 = __shfl_sync(0xffffffffu, x1, ((2 * t + odd) << 2) | h);
  const double r2 = __shfl_sync(0xffffffffu, x2, ((2 * t + 1 - odd) << 2) | h);
  return odd ? mkt(r2, r1) : mkt(r1, r2);
}

struct alignas(16) dbl2 { double x, y; };
FFMPC_DEV T2 lds_t(const double *p) { const dbl2 v = *reinterpret_cast<const dbl2 *>(p); return mkt(v.x, v.y); }
FFMPC_DEV F2 lds_f(const double *p) { const dbl2 v = *reinterpret_cast<const dbl2 *>(p); F2 f; f.k0 = v.x; f.k1 = v.y; return f; }
FFMPC_DEV void sts_2(double *p, double a, double b) { dbl2 v; v.x = a; v.y = b; *reinterpret_cast<dbl2 *>(p) = v; }

// ------------------------------------------------------------------------------------------------
// knot-record ring: one cp.async.bulk (TMA, 1-D) per knot, completion counted in bytes on an mbarrier
// ------------------------------------------------------------------------------------------------
FFMPC_DEV void ring_init(unsigned long long *mb, int lane) {
#if !defined(FFMPC_EMU)
  if (lane == 0) {
    const unsigned a = (unsigned)__cvta_generic_to_shared(mb);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(a) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(a + 8u) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
#else
  (void)mb; (void)lane;
#endif
  __syncwarp();
}
// called by the whole warp after a __syncwarp that ends every read of `dst`
FFMPC_DEV void ring_issue(double *dst, const double *src, unsigned long long *mb, int lane) {
#if defined(FFMPC_EMU)
  (void)mb;
  for (int i = lane; i < REC_SIZE; i += 32) dst[i] = src[i];
#else
  if (lane == 0) {
    const unsigned m = (unsigned)__cvta_generic_to_shared(mb), d = (unsigned)__cvta_generic_to_shared(dst);
    const unsigned bytes = (unsigned)(REC_SIZE * sizeof(double));
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(m), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d), "l"(src), "r"(bytes), "r"(m)
                 : "memory");
  }
#endif
}
FFMPC_DEV void ring_wait(unsigned long long *mb, unsigned parity) {
#if defined(FFMPC_EMU)
  (void)mb; (void)parity;
  __syncwarp();
#else
  const unsigned m = (unsigned)__cvta_generic_to_shared(mb);
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "LAB_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE;\n\t"
      "bra LAB_WAIT;\n\t"
      "DONE:\n\t"
      "}" ::"r"(m), "r"(parity)
      : "memory");
#endif
}

// per-warp shared memory (doubles)
struct BwdSmem {
  static constexpr int oST = 0;                    // two record stages
  static constexpr int oQS = 2 * REC_SIZE;         // 9 x 8: Quu tile rows 0..7, row 8 = Qu;  reused for the inverse (HI)
  static constexpr int oLL = oQS + 72;             // Cholesky factor rows
  static constexpr int oDI = oLL + 64;             // reciprocal pivots
  static constexpr int oMB = oDI + 8;              // two mbarriers
  static constexpr int TOTAL = oMB + 2;            // 1330 doubles = 10640 bytes (a multiple of 16)
  static_assert(TOTAL % 2 == 0 && oQS % 2 == 0 && oMB % 2 == 0, "16-byte alignment");
};

// ------------------------------------------------------------------------------------------------
// one backward sweep at regularisation `preg`; false = Cholesky / BoxQP failure or non-finite value function
// (SolverFDDP::backwardPass raising), the caller escalates the regularisation
// ------------------------------------------------------------------------------------------------
template <bool FF>
FFMPC_DEV bool riccati_sweep(const Params &P, int b, double *sm, double preg, bool feasible, unsigned &phase, double &dg_out,
                             double &dq_out) {
  constexpr int NBX = FF ? 3 : 2;   // state blocks
  constexpr int n = FF ? 21 : 14;
  const Workspace &ws = P.ws;
  const ffmpc_config &cfg = P.cfg;
  const int N = ws.N;
  const int lane = (int)(threadIdx.x & 31), g = lane >> 2, t4 = lane & 3;
  const int r7 = lane < NJ ? lane : NJ - 1;
  const bool act7 = lane < NJ;
  const bool row7 = (g == 7);                        // homogeneous row of tile row 0, padding row elsewhere
  const bool real_row = (g < NJ);
  const bool dg0 = real_row && (g == 2 * t4), dg1 = real_row && (g == 2 * t4 + 1);  // c0 / c1 on a tile diagonal
  const double dt = cfg.dt_ocp;
  const double alpha = fmin(fmax(cfg.ff_alpha, 0.0), 0.999999), beta = 1.0 - alpha;
  const bool box = cfg.use_box_fddp && feasible;
  double *ST0 = sm + BwdSmem::oST, *ST1 = ST0 + REC_SIZE, *QS = sm + BwdSmem::oQS, *LL = sm + BwdSmem::oLL, *DI = sm + BwdSmem::oDI,
         *HI = QS;
  unsigned long long *MB = reinterpret_cast<unsigned long long *>(sm + BwdSmem::oMB);
  const double *recb = ws.rec + (size_t)b * (N + 1) * REC_SIZE;
  const double *fsb = ws.fs + (size_t)b * (N + 1) * n;
  const double *usb = ws.us + ((size_t)ws.cur[b] * ws.B + b) * N * NJ;
  double *Kb = ws.K + (size_t)b * N * NJ * n;
  double *kb = ws.k + (size_t)b * N * NJ;
  double *Vfsb = ws.Vfs + (size_t)b * (N + 1) * n;
  const int toff = REC_LZZ + lane * 2, foff = REC_GH + lane * 2;
  // value function: upper tiles V[id], id = (0,0) (0,1) (0,2) (1,1) (1,2) (2,2) -> 0..5, and the transposed lower tiles
  T2 V[6], L10, L20, L21;
#pragma unroll
  for (int i = 0; i < 6; ++i) V[i] = mkt(0.0, 0.0);
  L10 = L20 = L21 = mkt(0.0, 0.0);
  double dgp = 0.0, dqp = 0.0;   // per-lane partial sums of Qu.k and -k.Quu k (lanes 0..6)
  double acc_g = 0.0, acc_q = 0.0;  // per-lane partial sums of Vx_old.fs (lane 28) and fs.Vxx fs

  // Vx += Vxx fs, Vxx fs -> HBM (the rollout's expected-improvement term), dg / dq contributions: only while gaps are open
  auto gap_update = [&](int t) {
    const double *f = fsb + (size_t)t * n;
    T2 fc[NBX];
    double fr[NBX];
#pragma unroll
    for (int bb = 0; bb < NBX; ++bb) {
      fc[bb] = mkt(f[7 * bb + 2 * t4], (t4 < 3) ? f[7 * bb + 2 * t4 + 1] : 0.0);
      fr[bb] = real_row ? f[7 * bb + g] : 0.0;
    }
    double y[NBX];
    y[0] = V[0].c0 * fc[0].c0 + V[0].c1 * fc[0].c1 + V[1].c0 * fc[1].c0 + V[1].c1 * fc[1].c1;
    y[1] = L10.c0 * fc[0].c0 + L10.c1 * fc[0].c1 + V[3].c0 * fc[1].c0 + V[3].c1 * fc[1].c1;
    if (FF) {
      y[0] += V[2].c0 * fc[NBX - 1].c0 + V[2].c1 * fc[NBX - 1].c1;
      y[1] += V[4].c0 * fc[NBX - 1].c0 + V[4].c1 * fc[NBX - 1].c1;
      y[NBX - 1] = L20.c0 * fc[0].c0 + L20.c1 * fc[0].c1 + L21.c0 * fc[1].c0 + L21.c1 * fc[1].c1 + V[5].c0 * fc[NBX - 1].c0 +
                   V[5].c1 * fc[NBX - 1].c1;
    }
#pragma unroll
    for (int a = 0; a < NBX; ++a) {
      y[a] += __shfl_xor_sync(0xffffffffu, y[a], 1);
      y[a] += __shfl_xor_sync(0xffffffffu, y[a], 2);
      // lanes of group g hold (Vxx fs)[7a + g];  group 7 of a = 0 holds Vx . fs (the homogeneous row)
      if (real_row && t4 == 0) Vfsb[(size_t)t * n + 7 * a + g] = y[a];
      acc_q += (real_row && t4 == 0) ? y[a] * fr[a] : 0.0;
      const double v0 = shfl_d(y[a], (2 * t4) << 2), v1 = shfl_d(y[a], ((2 * t4 + 1) & 7) << 2);
      const int id = a;  // tiles (0, a)
      if (row7) {
        V[id].c0 += v0;
        if (t4 < 3) V[id].c1 += v1;
      }
    }
    acc_g += (lane == 28) ? y[0] : 0.0;
  };
  auto all_finite = [&]() {
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i) s += V[i].c0 + V[i].c1;
    return !__any_sync(0xffffffffu, !isfinite(s));
  };

  // ---- terminal knot: Vxx = Lxx + preg, Vx = Lx (+ Vxx fs) ----
  __syncwarp();
  ring_issue(ST0, recb + (size_t)N * REC_SIZE, MB, lane);
  ring_issue(ST1, recb + (size_t)(N - 1) * REC_SIZE, MB + 1, lane);
  ring_wait(MB, phase & 1u);
  phase ^= 1u;
  V[0] = lds_t(ST0 + toff);
  V[1] = lds_t(ST0 + toff + 64);
  V[3] = lds_t(ST0 + toff + 3 * 64);
  if (FF) {
    V[2] = lds_t(ST0 + toff + 2 * 64);
    V[4] = lds_t(ST0 + toff + 4 * 64);
    V[5] = lds_t(ST0 + toff + 5 * 64);
  }
  {
    const double p0 = dg0 ? preg : 0.0, p1 = dg1 ? preg : 0.0;
    V[0].c0 += p0; V[0].c1 += p1;
    V[3].c0 += p0; V[3].c1 += p1;
    if (FF) { V[5].c0 += p0; V[5].c1 += p1; }
  }
  L10 = transpose_t(V[1], lane);
  if (FF) {
    L20 = transpose_t(V[2], lane);
    L21 = transpose_t(V[4], lane);
  }
  if (!feasible) gap_update(N);
  __syncwarp();

  for (int t = N - 1; t >= 0; --t) {
    const int s = (N - t) & 1;
    double *R = s ? ST1 : ST0;
    // the other stage held knot t+1, fully consumed before the barrier that ended the previous step
    if (t > 0) ring_issue(s ? ST0 : ST1, recb + (size_t)(t - 1) * REC_SIZE, MB + (1 - s), lane);
    // BoxQP operands of this knot (nominal control, warm start = stored k_t): requested early
    const double u0_pre = box ? usb[t * NJ + r7] : 0.0;
    const double kq_pre = box ? kb[t * NJ + r7] : 0.0;
    ring_wait(MB + s, (phase >> s) & 1u);
    phase ^= (1u << s);
    auto drain = [&]() {  // leave no copy in flight (and the phase bits consistent) when the sweep is abandoned
      if (t > 0) {
        ring_wait(MB + (1 - s), (phase >> (1 - s)) & 1u);
        phase ^= (1u << (1 - s));
      }
      __syncwarp();
    };
    F2 GT[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) GT[a] = lds_f(R + foff + 64 * a);

    // ---- structured products, tile-wise: VE = V Et, S = Et^T V Et = Tt_1 = W_11, Tt_0 = W_01 = VE_0 ----
    const T2 VE0 = taxpy(dt, V[0], V[1]);
    const T2 VE1 = taxpy(dt, L10, V[3]);
    const T2 S = taxpy(dt, VE0, VE1);
    T2 VE2 = mkt(0.0, 0.0), PV1 = mkt(0.0, 0.0);
    if (FF) {
      VE2 = taxpy(dt, L20, L21);
      PV1 = taxpy(dt, V[2], V[4]);  // (Pt^T V)[v, tau]
    }
    // ---- Z = Tt + 1/2 Gh^T S ----
    T2 Z[3];
    F2 Zf[3];
    {
      const F2 Sf = frag_of(tscale(0.5, S), lane);
      Z[0] = VE0;
      Z[1] = S;
      Z[2] = FF ? tscale(alpha, VE2) : mkt(0.0, 0.0);
#pragma unroll
      for (int a = 0; a < 3; ++a) mma88(Z[a], GT[a], Sf);
#pragma unroll
      for (int a = 0; a < 3; ++a) Zf[a] = frag_of(Z[a], lane);
    }
    // ---- Q = Lzz + W + Z Gh + Gh^T Z^T (upper tiles; row 7 of the tiles (0, b) = Qz) ----
    T2 Q[6];
    Q[0] = tadd(lds_t(R + toff), V[0]);
    Q[1] = tadd(lds_t(R + toff + 64), VE0);
    Q[3] = tadd(lds_t(R + toff + 3 * 64), S);
    if (FF) {
      Q[2] = taxpy(alpha, V[2], lds_t(R + toff + 2 * 64));
      Q[4] = taxpy(alpha, PV1, lds_t(R + toff + 4 * 64));
      Q[5] = taxpy(alpha * alpha, V[5], lds_t(R + toff + 5 * 64));
    } else {
      Q[2] = lds_t(R + toff + 2 * 64);
      Q[4] = lds_t(R + toff + 4 * 64);
      Q[5] = lds_t(R + toff + 5 * 64);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int bb = a; bb < 3; ++bb) {
        const int id = (a == 0) ? bb : (a == 1 ? 2 + bb : 5);
        mma88(Q[id], Zf[a], GT[bb]);
        mma88(Q[id], GT[a], Zf[bb]);
      }
    // ---- Qxu (row 7 of tile 0 = Qu), Quu + preg ----
    T2 Qxu[NBX], Quu;
    if (FF) {
      const F2 Rf = frag_of(tscale(beta, VE2), lane);  // (Et^T V Fu)^T
      Qxu[0] = tscale(beta, V[2]);
      Qxu[1] = tscale(beta, PV1);
      Qxu[NBX - 1] = tscale(alpha * beta, V[5]);
#pragma unroll
      for (int a = 0; a < NBX; ++a) mma88(Qxu[a], GT[a], Rf);
      if (row7) {
        const T2 lw = lds_t(R + REC_LW + 2 * t4);
        Qxu[0].c0 += lw.c0;
        Qxu[0].c1 += lw.c1;
      }
      Quu = tscale(beta * beta, V[5]);
      if (dg0) Quu.c0 += R[REC_LWW + g] + preg;
      if (dg1) Quu.c1 += R[REC_LWW + g] + preg;
    } else {
      Qxu[0] = Q[2];
      Qxu[1] = Q[4];
      Quu = Q[5];
      if (dg0) Quu.c0 += preg;
      if (dg1) Quu.c1 += preg;
    }
    // ---- gains, row-distributed over lanes 0..6: stage Quu (rows 0..7) and Qu (row 8), read back one row per lane ----
    sts_2(QS + g * 8 + 2 * t4, Quu.c0, Quu.c1);
    if (row7) sts_2(QS + 64 + 2 * t4, Qxu[0].c0, Qxu[0].c1);
    __syncwarp();
    double qrow[NJ], qu;
    {
      const dbl2 a0 = *reinterpret_cast<const dbl2 *>(QS + r7 * 8), a1 = *reinterpret_cast<const dbl2 *>(QS + r7 * 8 + 2),
                 a2 = *reinterpret_cast<const dbl2 *>(QS + r7 * 8 + 4), a3 = *reinterpret_cast<const dbl2 *>(QS + r7 * 8 + 6);
      qrow[0] = a0.x; qrow[1] = a0.y; qrow[2] = a1.x; qrow[3] = a1.y; qrow[4] = a2.x; qrow[5] = a2.y; qrow[6] = a3.x;
      qu = QS[64 + r7];
    }
    __syncwarp();  // QS is rewritten with L^-1 below
    // Factor Quu = L L^T (BoxQP: of the free block, clamped rows / columns replaced by identity); the factor stays in LL / DI.
    unsigned mask = 0x7fu;
    double kv = 0.0, wv = 0.0;
    if (!box) {
      // crocoddyl computeGains: Quu_llt.solve(Qxu^T), Quu_llt.solve(Qu)
      double a[NJ], dinv;
#pragma unroll
      for (int j = 0; j < NJ; ++j) a[j] = qrow[j];
      const bool okc = dist_chol7(a, 0x7fu, lane, dinv);
      if (act7) {
        sts_2(LL + lane * 8, a[0], a[1]); sts_2(LL + lane * 8 + 2, a[2], a[3]);
        sts_2(LL + lane * 8 + 4, a[4], a[5]); sts_2(LL + lane * 8 + 6, a[6], 0.0);
        DI[lane] = dinv;
      }
      __syncwarp();
      if (!okc) { drain(); return false; }
    } else {
      const double lb = -cfg.tau_limits[r7] - u0_pre, ub = cfg.tau_limits[r7] - u0_pre;
      double xq = kq_pre;  // warm start = stored k_t (crocoddyl passes k_[t])
      double hi[NJ];
      if (!dist_boxqp7(cfg, qrow, qu, lb, ub, xq, lane, LL, DI, mask, hi)) { drain(); return false; }
      kv = -xq;
      if (!((mask >> r7) & 1u)) qu = 0.0;  // clamped directions do not count for the stopping measure
      double kvv[NJ], qk = 0.0;
      gather7(kv, kvv);
#pragma unroll
      for (int bb = 0; bb < NJ; ++bb) qk += qrow[bb] * kvv[bb];
      dgp += act7 ? qu * kv : 0.0;
      dqp -= act7 ? kv * qk : 0.0;
      wv = act7 ? qk - 2.0 * qu : 0.0;  // Quu k - 2 Qu
      if (act7) kb[t * NJ + lane] = kv;
    }
    // Column `lane` of L^-1 by forward substitution (zero for a clamped index), stored transposed: HI[c * 8 + i] = Linv[i][c].
    // The gains and the value update go through W = Qxu L^-T (K^T = W L^-1, Qxu K = W W^T): no product with the explicit
    // inverse of an ill-conditioned Quu, and an exactly symmetric downdate.
    {
      double y[NJ];
#pragma unroll
      for (int i = 0; i < NJ; ++i) {
        double sacc = (i == r7) ? 1.0 : 0.0;
#pragma unroll
        for (int k = 0; k < i; ++k) sacc -= LL[i * 8 + k] * y[k];
        y[i] = sacc * DI[i];
      }
      const bool keep = act7 && ((mask >> r7) & 1u);
      if (lane < 8) {
        double *h = HI + lane * 8;
        sts_2(h, keep ? y[0] : 0.0, keep ? y[1] : 0.0);
        sts_2(h + 2, keep ? y[2] : 0.0, keep ? y[3] : 0.0);
        sts_2(h + 4, keep ? y[4] : 0.0, keep ? y[5] : 0.0);
        sts_2(h + 6, keep ? y[6] : 0.0, 0.0);
      }
    }
    __syncwarp();
    F2 Lf, LTf;  // fragments of L^-1 and of its transpose
    Lf.k0 = HI[t4 * 8 + g];
    Lf.k1 = HI[(4 + t4) * 8 + g];
    LTf.k0 = HI[g * 8 + t4];
    LTf.k1 = HI[g * 8 + 4 + t4];
    // ---- W = Qxu L^-T, K^T = W L^-1 -> HBM ----
    F2 Wf[NBX], An[NBX];
    double a7 = 0.0;
    if (box) {
      // row 7 of the A operand of tile row 0 := L^-1 (Quu k - 2 Qu), so that row 7 of the downdate is K^T (Quu k - 2 Qu)
      double wvv[NJ];
      gather7(wv, wvv);
#pragma unroll
      for (int c = 0; c < NJ; ++c) a7 += HI[c * 8 + r7] * wvv[c];
      if (!act7) a7 = 0.0;
    }
#pragma unroll
    for (int a = 0; a < NBX; ++a) {
      const F2 Qf = frag_of(Qxu[a], lane);
      T2 W = mkt(0.0, 0.0);
      mma88(W, Qf, Lf);   // W[g][j] = sum_k Qxu[g][k] Linv[j][k]
      Wf[a] = frag_of(W, lane);
      T2 KT = mkt(0.0, 0.0);
      mma88(KT, Wf[a], LTf);  // KT[g][j] = sum_k W[g][k] Linv[k][j] = K[j][7a + g]
      if (real_row) {
        double *Kt = Kb + (size_t)t * NJ * n + 7 * a + g;
        Kt[(2 * t4) * n] = KT.c0;
        if (t4 < 3) Kt[(2 * t4 + 1) * n] = KT.c1;
      }
      if (a == 0 && !box) {
        // unconstrained step: row 7 of W is (L^-1 Qu)^T, of K^T it is k^T = (Quu^-1 Qu)^T;  Qu.k = k.Quu k = |L^-1 Qu|^2
        if (row7) {
          kb[t * NJ + 2 * t4] = KT.c0;
          if (t4 < 3) kb[t * NJ + 2 * t4 + 1] = KT.c1;
          const double s2 = W.c0 * W.c0 + W.c1 * W.c1;
          dgp += s2;
          dqp -= s2;
        }
      }
      An[a].k0 = -Wf[a].k0;
      An[a].k1 = -Wf[a].k1;
    }
    if (box) {
      const double w0 = shfl_d(a7, t4), w1 = shfl_d(a7, 4 + t4);
      if (row7) {
        An[0].k0 = w0;
        An[0].k1 = w1;
      }
    }
    // ---- Vxx = Qxx - W W^T (+ preg), upper tiles (row 7 of tile row 0: Vx = Qx + K^T (Quu k - 2 Qu)); padding cleaned ----
    {
      const double p0 = dg0 ? preg : 0.0, p1 = dg1 ? preg : 0.0;
      const bool col7 = (t4 == 3);
#pragma unroll
      for (int a = 0; a < NBX; ++a)
#pragma unroll
        for (int bb = a; bb < NBX; ++bb) {
          const int id = (a == 0) ? bb : (a == 1 ? 2 + bb : 5);
          T2 v = Q[id];
          mma88(v, An[a], Wf[bb]);
          if (a == bb) { v.c0 += p0; v.c1 += p1; }
          if (col7) v.c1 = 0.0;                     // column 7: padding
          if (a > 0 && row7) v = mkt(0.0, 0.0);     // row 7 of the tile rows 1, 2: padding (tile row 0: Vx)
          V[id] = v;
        }
    }
    L10 = transpose_t(V[1], lane);
    if (FF) {
      L20 = transpose_t(V[2], lane);
      L21 = transpose_t(V[4], lane);
    }
    if (!feasible) gap_update(t);
    if (!all_finite()) { drain(); return false; }
    __syncwarp();  // every lane is done with this knot's record stage and the gain scratch
  }
  dg_out = warp_sum(dgp - (acc_g + acc_q));
  dq_out = warp_sum(dqp + acc_q);
  return true;
}

// Residency: FFMPC_BWD_WARPS warps (= instances) per block, FFMPC_BWD_MINBLOCKS blocks per SM (10.4 KB of shared memory per warp;
// 12 one-warp blocks = 168 registers per thread).
#ifndef FFMPC_BWD_WARPS
#define FFMPC_BWD_WARPS 1
#endif
#ifndef FFMPC_BWD_MINBLOCKS
#define FFMPC_BWD_MINBLOCKS 12
#endif
template <bool FF>
__global__ void __launch_bounds__(FFMPC_BWD_WARPS * 32, FFMPC_BWD_MINBLOCKS) k_backward(const __grid_constant__ Params P) {
#ifdef FFMPC_EMU
  double *smem = (double *)emu::S().smem;
#else
  extern __shared__ __align__(16) double smem[];
#endif
  const Workspace &ws = P.ws;
  // one warp per block by default: the instance index is then a function of blockIdx alone, the early exits below are
  // provably warp-uniform and the compiler drops the WARPSYNC / ENDCOLLECTIVE pair it otherwise wraps around every shuffle
  const int warp = (FFMPC_BWD_WARPS == 1) ? 0 : (int)(threadIdx.x >> 5), lane = (int)(threadIdx.x & 31);
  const int b = (int)blockIdx.x * FFMPC_BWD_WARPS + warp;
  if (b >= ws.B) return;
  if (ws.done[b]) return;
  double *sm = smem + (size_t)warp * BwdSmem::TOTAL;
  ring_init(reinterpret_cast<unsigned long long *>(sm + BwdSmem::oMB), lane);
  unsigned phase = 0u;
  double preg = ws.preg[b];
  const bool feasible = ws.feas[b] != 0;
  const int it = ws.iter[b];
  // computeDirection with the regularisation-escalation loop of SolverFDDP::solve
  while (true) {
    double dg, dq;
    const bool okb = riccati_sweep<FF>(P, b, sm, preg, feasible, phase, dg, dq);
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

}  // namespace ffmpc
