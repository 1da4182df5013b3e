// ffmpc_coop.cuh -- lane-cooperative evaluation of ONE action-model knot: 8 lanes ("octet") per knot,
// 4 knots per warp.  Lane j < 7 owns joint j (its transform, Jacobian column, velocity, inertia, and the
// columns j / 7+j / 14+j of every 21-wide derivative); lane 7 carries zeros.  The world-frame recursions of
// ffmpc_math.cuh, distributed so that
//   * serial-chain recursions become width-8 shuffle scans (prefix: velocities / accelerations,
//     suffix: composite inertias / forces / B operators),
//   * all-to-all products (CRBA, RNEA derivative blocks, KKT products, Gauss-Newton Hessian) read the
//     other joints' operands from a per-octet shared-memory scratch (CS_* below),
//   * a knot needs ~90 live doubles per lane instead of ~550 per thread: no local-memory spills.
// Replaces for the reference: crocoddyl_classical.py:558-728 (_make_dam), :542 (Euler),
// crocoddyl_force_feedback.py:211-290 (_AugmentedLPFActionModel).  Math: SURVEY.md App. A.1-A.6.
//
// Control flow is WARP-uniform by construction (every shuffle / __syncwarp is executed by all 32 lanes);
// octets without work run on dummy data with their writes disabled.
#pragma once

#include "ffmpc_math.cuh"

namespace ffmpc {

constexpr unsigned kFull = 0xffffffffu;

// ---- per-block table of the per-joint constants (model + config), one row per joint ----
constexpr int CT_ROT = 0, CT_POS = 9, CT_COM = 12, CT_INR = 15, CT_MASS = 21, CT_VDAMP = 22, CT_QSREF = 23, CT_QSLB = 24,
              CT_QSUB = 25, CT_TAULIM = 26, CT_YWQ = 27, CT_YWV = 28, CT_YWT = 29;
constexpr int CT_STRIDE = 30;
constexpr int CT_SIZE = NJ * CT_STRIDE + 2;  // +2: keeps the octet scratch 16-byte aligned and bank-staggered

// ---- per-octet scratch (doubles) ----
constexpr int CS_T = 0;     // 7 x 12  joint-local transforms (Rl row-major, pl)            [FK only]
constexpr int CS_J = 84;    // 7 x 6   J_i
constexpr int CS_DF = 126;  // 7 x 6   dFda_i = Ycrb_i J_i
constexpr int CS_HS = 0;    // 7 x 22  Gauss-Newton exchange (aliases T / J / DF, all dead by then)
constexpr int CS_HF = 154;  // 7 x 6   d lambda_x, d lambda_y / d (q_i, v_i, u_i): friction cone only (aliases DF tail / BT, dead by then)
constexpr int CS_BT = 168;  // 7 x 4   angular part of Bcrb_i^T J_i
constexpr int CS_JF = 196;  // 7 x 4   Jf_i (frame linear Jacobian column)
constexpr int CS_FR = 224;  // 32      EE frame block: pf vf wf a_cl Rf v7
constexpr int CS_LL = 256;  // 7 x 8   Cholesky factor rows of M, [56..62] reciprocal pivots
constexpr int CS_TL = 320;  // 7 x 8   KKT inverse top-left block (free: M^-1)
constexpr int CS_SIZE = 388;  // == 4 (mod 16): the four octets of a warp start 8 banks apart

FFMPC_DEV double oct_bcast(double v, int src) { return __shfl_sync(kFull, v, src, 8); }
FFMPC_DEV double oct_sum(double v) {
  v += __shfl_xor_sync(kFull, v, 1, 8);
  v += __shfl_xor_sync(kFull, v, 2, 8);
  v += __shfl_xor_sync(kFull, v, 4, 8);
  return v;
}
// inclusive prefix sum over the octet (lane j gets sum of lanes 0..j)
FFMPC_DEV double oct_scan_up(double v, int j) {
  // masked add as one DFMA (fma(1, t, v) == v + t exactly; a lane without a source receives its own, finite, value)
  double t = __shfl_up_sync(kFull, v, 1, 8);
  v = fma(j >= 1 ? 1.0 : 0.0, t, v);
  t = __shfl_up_sync(kFull, v, 2, 8);
  v = fma(j >= 2 ? 1.0 : 0.0, t, v);
  t = __shfl_up_sync(kFull, v, 4, 8);
  v = fma(j >= 4 ? 1.0 : 0.0, t, v);
  return v;
}
// inclusive suffix sum (lane j gets sum of lanes j..7)
FFMPC_DEV double oct_scan_down(double v, int j) {
  double t = __shfl_down_sync(kFull, v, 1, 8);
  v = fma(j + 1 < 8 ? 1.0 : 0.0, t, v);
  t = __shfl_down_sync(kFull, v, 2, 8);
  v = fma(j + 2 < 8 ? 1.0 : 0.0, t, v);
  t = __shfl_down_sync(kFull, v, 4, 8);
  v = fma(j + 4 < 8 ? 1.0 : 0.0, t, v);
  return v;
}
FFMPC_DEV v3 scan_up3(const v3 &a, int j) { return mk3(oct_scan_up(a.x, j), oct_scan_up(a.y, j), oct_scan_up(a.z, j)); }
FFMPC_DEV sv scan_up6(const sv &a, int j) { return mksv(scan_up3(a.l, j), scan_up3(a.a, j)); }
FFMPC_DEV v3 scan_down3(const v3 &a, int j) { return mk3(oct_scan_down(a.x, j), oct_scan_down(a.y, j), oct_scan_down(a.z, j)); }
FFMPC_DEV sv scan_down6(const sv &a, int j) { return mksv(scan_down3(a.l, j), scan_down3(a.a, j)); }
// value of lane j-1 (zero for lane 0)
FFMPC_DEV double oct_prev(double v, int j) {
  const double t = __shfl_up_sync(kFull, v, 1, 8);
  return j >= 1 ? t : 0.0;
}
FFMPC_DEV sv prev6(const sv &a, int j) {
  return mksv(mk3(oct_prev(a.l.x, j), oct_prev(a.l.y, j), oct_prev(a.l.z, j)), mk3(oct_prev(a.a.x, j), oct_prev(a.a.y, j), oct_prev(a.a.z, j)));
}
FFMPC_DEV sv bcast6(const sv &a, int src) {
  return mksv(mk3(oct_bcast(a.l.x, src), oct_bcast(a.l.y, src), oct_bcast(a.l.z, src)),
              mk3(oct_bcast(a.a.x, src), oct_bcast(a.a.y, src), oct_bcast(a.a.z, src)));
}
// p[0..7] per lane -> returns sum over the octet of p[j] on lane j (reduce-scatter, 7 exchanges)
FFMPC_DEV double oct_reduce_scatter8(const double *p, int j) {
  double q4[4], q2[2];
  const bool b4 = (j & 4) != 0, b2 = (j & 2) != 0, b1 = (j & 1) != 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const double send = b4 ? p[i] : p[i + 4];
    const double keep = b4 ? p[i + 4] : p[i];
    q4[i] = keep + __shfl_xor_sync(kFull, send, 4, 8);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const double send = b2 ? q4[i] : q4[i + 2];
    const double keep = b2 ? q4[i + 2] : q4[i];
    q2[i] = keep + __shfl_xor_sync(kFull, send, 2, 8);
  }
  const double send = b1 ? q2[0] : q2[1];
  const double keep = b1 ? q2[1] : q2[0];
  return keep + __shfl_xor_sync(kFull, send, 1, 8);
}

struct alignas(16) cdbl2 { double x, y; };
FFMPC_DEV void lds2(const double *p, double &a, double &b) {
  const cdbl2 t = *reinterpret_cast<const cdbl2 *>(p);
  a = t.x; b = t.y;
}
FFMPC_DEV void sts2(double *p, double a, double b) {
  cdbl2 t; t.x = a; t.y = b;
  *reinterpret_cast<cdbl2 *>(p) = t;
}
FFMPC_DEV sv lds_sv(const double *p) {
  sv r;
  lds2(p, r.l.x, r.l.y);
  lds2(p + 2, r.l.z, r.a.x);
  lds2(p + 4, r.a.y, r.a.z);
  return r;
}
FFMPC_DEV void sts_sv(double *p, const sv &a) {
  sts2(p, a.l.x, a.l.y);
  sts2(p + 2, a.l.z, a.a.x);
  sts2(p + 4, a.a.y, a.a.z);
}
FFMPC_DEV v3 lds_v3p(const double *p) {  // 4-double padded row
  v3 r; double pad;
  lds2(p, r.x, r.y);
  lds2(p + 2, r.z, pad);
  return r;
}
FFMPC_DEV void sts_v3p(double *p, const v3 &a) {
  sts2(p, a.x, a.y);
  sts2(p + 2, a.z, 0.0);
}
FFMPC_DEV sv sel6(bool c, const sv &a, const sv &b) {
  return mksv(mk3(c ? a.l.x : b.l.x, c ? a.l.y : b.l.y, c ? a.l.z : b.l.z), mk3(c ? a.a.x : b.a.x, c ? a.a.y : b.a.y, c ? a.a.z : b.a.z));
}

// Entry j of a 7-wide row of a record tile (8 doubles, the last one padding), one lane per entry: lane 6 writes the
// padding zero with its entry, so that the warp's store covers whole 32-byte sectors (no partial-sector write-backs).
FFMPC_DEV void row_store7(double *row, int j, double v) {
  if (j == NJ - 1) sts2(row + NJ - 1, v, 0.0);
  else row[j] = v;
}

// fills the per-block constant table; every thread of the block calls it, followed by __syncthreads()
FFMPC_DEV void coop_fill_table(const ffmpc_model &mdl, const ffmpc_config &cfg, double *ct) {
  for (int e = (int)threadIdx.x; e < NJ * CT_STRIDE; e += (int)blockDim.x) {
    const int i = e / CT_STRIDE, c = e - i * CT_STRIDE;
    double v;
    if (c < CT_POS) v = mdl.joint_rot[i][c];
    else if (c < CT_COM) v = mdl.joint_pos[i][c - CT_POS];
    else if (c < CT_INR) v = mdl.com[i][c - CT_COM];
    else if (c < CT_MASS) v = mdl.inertia[i][c - CT_INR];
    else if (c == CT_MASS) v = mdl.mass[i];
    else if (c == CT_VDAMP) v = cfg.v_damp_weights[i];
    else if (c == CT_QSREF) v = cfg.q_soft_ref[i];
    else if (c == CT_QSLB) v = cfg.q_soft_lb[i];
    else if (c == CT_QSUB) v = cfg.q_soft_ub[i];
    else if (c == CT_TAULIM) v = cfg.tau_limits[i];
    else if (c == CT_YWQ) v = cfg.y_reg_weights[i];
    else if (c == CT_YWV) v = cfg.y_reg_weights[NJ + i];
    else v = cfg.y_reg_weights[2 * NJ + i];
    ct[e] = v;
  }
}

// Per-lane inputs / outputs of one cooperative knot evaluation.
struct CoopIn {
  double q, qd, u;             // own joint: position, velocity, inner control (torque)
  double xr_q, xr_v, tau_ref;  // own components of x_reg_ref and tau_ref
  const double *p_ref, *v_ref; // [3] each (octet-uniform)
  bool terminal_costs, with_control;
};

// Differential action model, cooperative.  Outputs: a (own joint acceleration), lam[3] (uniform), cost split
// into cpart (per-lane partial, to be summed over the octet) and cuni (octet-uniform part); when DIFF the record
// blocks G / Lz / Lzz are written to `rec` (global memory) by the lanes with `wr` set.
//   Lz = sc * dL/dz + eg,  Lzz = sc * H + diag(eh)   (eg3 / eh3: own components q, v, u)
template <int NC, bool DIFF>
FFMPC_DEV void coop_dam(const ffmpc_model &mdl, const ffmpc_config &cfg, const double *ct, double *sm, int j, const CoopIn &in,
                        double &a_out, double *lam_out, double &cpart, double &cuni, double *rec, bool wr, double sc,
                        const double *eg3, const double *eh3) {
  const int jj = j < NJ ? j : NJ - 1;
  const bool act = j < NJ;
  const double *cj = ct + jj * CT_STRIDE;
  const double qd = in.qd;
  __syncwarp();  // the previous knot's readers of the scratch are done

  // ---------------- forward kinematics: local transform -> chain product ----------------
  {
    double s, c;
    sincos(in.q, &s, &c);
    double Rl[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      Rl[3 * r + 0] = cj[CT_ROT + 3 * r] * c + cj[CT_ROT + 3 * r + 1] * s;
      Rl[3 * r + 1] = -cj[CT_ROT + 3 * r] * s + cj[CT_ROT + 3 * r + 1] * c;
      Rl[3 * r + 2] = cj[CT_ROT + 3 * r + 2];
    }
    if (act) {
      double *T = sm + CS_T + 12 * j;
      sts2(T, Rl[0], Rl[1]); sts2(T + 2, Rl[2], Rl[3]); sts2(T + 4, Rl[4], Rl[5]); sts2(T + 6, Rl[6], Rl[7]);
      sts2(T + 8, Rl[8], cj[CT_POS]); sts2(T + 10, cj[CT_POS + 1], cj[CT_POS + 2]);
    }
  }
  __syncwarp();
  m3 R;
  v3 p;
  {
    const double *T = sm + CS_T;
    lds2(T, R.m[0], R.m[1]); lds2(T + 2, R.m[2], R.m[3]); lds2(T + 4, R.m[4], R.m[5]); lds2(T + 6, R.m[6], R.m[7]);
    lds2(T + 8, R.m[8], p.x); lds2(T + 10, p.y, p.z);
#pragma unroll
    for (int k = 1; k < NJ; ++k) {
      const double *Tk = sm + CS_T + 12 * k;
      m3 Rk;
      v3 pk;
      lds2(Tk, Rk.m[0], Rk.m[1]); lds2(Tk + 2, Rk.m[2], Rk.m[3]); lds2(Tk + 4, Rk.m[4], Rk.m[5]); lds2(Tk + 6, Rk.m[6], Rk.m[7]);
      lds2(Tk + 8, Rk.m[8], pk.x); lds2(Tk + 10, pk.y, pk.z);
      if (k <= jj) {
        p = p + mulv(R, pk);
        R = mulm(R, Rk);
      }
    }
  }
  const v3 z = mk3(R.m[2], R.m[5], R.m[8]);
  const sv J = mksv(cross(p, z), z);
  const double zero_or_one = act ? 1.0 : 0.0;
  const sv v = scan_up6(J * (act ? qd : 0.0), j);
  const sv dVdq = mxm(v, J);  // = v_j x J_j ; note v_{j-1} x J_j qd_j == v_j x J_j qd_j
  const sv ab = scan_up6(dVdq * (act ? qd : 0.0), j);

  // world inertia about the origin (zero on lane 7)
  si Y;
  {
    const v3 cw = p + mulv(R, mk3(cj[CT_COM], cj[CT_COM + 1], cj[CT_COM + 2]));
    m3 Ib;
    Ib.m[0] = cj[CT_INR + 0]; Ib.m[4] = cj[CT_INR + 1]; Ib.m[8] = cj[CT_INR + 2];
    Ib.m[1] = Ib.m[3] = cj[CT_INR + 3];
    Ib.m[2] = Ib.m[6] = cj[CT_INR + 4];
    Ib.m[5] = Ib.m[7] = cj[CT_INR + 5];
    const m3 RI = mulm(R, Ib);
    const double m = cj[CT_MASS] * zero_or_one;
    const double c2 = dot(cw, cw);
    Y.I[0] = zero_or_one * (RI.m[0] * R.m[0] + RI.m[1] * R.m[1] + RI.m[2] * R.m[2]) + m * (c2 - cw.x * cw.x);
    Y.I[1] = zero_or_one * (RI.m[0] * R.m[3] + RI.m[1] * R.m[4] + RI.m[2] * R.m[5]) - m * cw.x * cw.y;
    Y.I[2] = zero_or_one * (RI.m[0] * R.m[6] + RI.m[1] * R.m[7] + RI.m[2] * R.m[8]) - m * cw.x * cw.z;
    Y.I[3] = zero_or_one * (RI.m[3] * R.m[3] + RI.m[4] * R.m[4] + RI.m[5] * R.m[5]) + m * (c2 - cw.y * cw.y);
    Y.I[4] = zero_or_one * (RI.m[3] * R.m[6] + RI.m[4] * R.m[7] + RI.m[5] * R.m[8]) - m * cw.y * cw.z;
    Y.I[5] = zero_or_one * (RI.m[6] * R.m[6] + RI.m[7] * R.m[7] + RI.m[8] * R.m[8]) + m * (c2 - cw.z * cw.z);
    Y.m = m;
    Y.h = cw * m;
  }
  const sv ag = mksv(mk3(-mdl.gravity[0], -mdl.gravity[1], -mdl.gravity[2]), mk3(0, 0, 0));
  const sv fb = si_mul(Y, ab + ag) + mxf(v, si_mul(Y, v));
  si Yc;
  Yc.m = oct_scan_down(Y.m, j);
  Yc.h = scan_down3(Y.h, j);
#pragma unroll
  for (int i = 0; i < 6; ++i) Yc.I[i] = oct_scan_down(Y.I[i], j);
  const sv Fs = scan_down6(fb, j);
  const double nle = dot6(J, Fs);
  const sv dFda = si_mul(Yc, J);

  // EE frame (computed from joint 7's data; lanes 6 / 7 hold it, lane 6 publishes)
  {
    m3 eeR;
#pragma unroll
    for (int i = 0; i < 9; ++i) eeR.m[i] = mdl.ee_rot[i];
    const m3 Rf_ = mulm(R, eeR);
    const v3 pf_ = p + mulv(R, mk3(mdl.ee_pos[0], mdl.ee_pos[1], mdl.ee_pos[2]));
    const v3 wf_ = v.a;
    const v3 vf_ = v.l + cross(wf_, pf_);
    const v3 acl_ = ab.l + cross(ab.a, pf_) + cross(wf_, vf_);
    if (act) {
      sts_sv(sm + CS_J + 6 * j, J);
      sts_sv(sm + CS_DF + 6 * j, dFda);
    }
    if (j == NJ - 1) {
      double *F = sm + CS_FR;
      sts2(F, pf_.x, pf_.y); sts2(F + 2, pf_.z, vf_.x); sts2(F + 4, vf_.y, vf_.z);
      sts2(F + 6, wf_.x, wf_.y); sts2(F + 8, wf_.z, acl_.x); sts2(F + 10, acl_.y, acl_.z);
      sts2(F + 12, Rf_.m[0], Rf_.m[1]); sts2(F + 14, Rf_.m[2], Rf_.m[3]); sts2(F + 16, Rf_.m[4], Rf_.m[5]);
      sts2(F + 18, Rf_.m[6], Rf_.m[7]); sts2(F + 20, Rf_.m[8], v.l.x); sts2(F + 22, v.l.y, v.l.z);
      sts2(F + 24, v.a.x, v.a.y); sts2(F + 26, v.a.z, 0.0);
    }
  }
  __syncwarp();
  v3 pf, vf, wf, a_cl;
  m3 Rf;
  sv v7;
  {
    const double *F = sm + CS_FR;
    lds2(F, pf.x, pf.y); lds2(F + 2, pf.z, vf.x); lds2(F + 4, vf.y, vf.z);
    lds2(F + 6, wf.x, wf.y); lds2(F + 8, wf.z, a_cl.x); lds2(F + 10, a_cl.y, a_cl.z);
    lds2(F + 12, Rf.m[0], Rf.m[1]); lds2(F + 14, Rf.m[2], Rf.m[3]); lds2(F + 16, Rf.m[4], Rf.m[5]);
    lds2(F + 18, Rf.m[6], Rf.m[7]); lds2(F + 20, Rf.m[8], v7.l.x); lds2(F + 22, v7.l.y, v7.l.z);
    double pad;
    lds2(F + 24, v7.a.x, v7.a.y); lds2(F + 26, v7.a.z, pad);
  }
  const v3 Jf = J.l + cross(J.a, pf);  // d pf / d q_j
  if (act) sts_v3p(sm + CS_JF + 4 * j, Jf);

  // ---------------- CRBA column j, distributed Cholesky, M^-1 column j ----------------
  double mi[NJ];  // M^-1[:, j] (== row j)
  {
    double a[NJ];
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      const bool up = i <= jj;
      const sv o = lds_sv(sm + (up ? CS_J : CS_DF) + 6 * i);
      a[i] = dot6(o, sel6(up, dFda, J));
    }
    double dinv = 1.0;
#pragma unroll
    for (int c = 0; c < NJ; ++c) {
      const double d = oct_bcast(a[c], c);
      const double r = rsqrt(d);
      const double lij = a[c] * r;
      if (jj == c) dinv = r;
      a[c] = lij;
#pragma unroll
      for (int k = c + 1; k < NJ; ++k) a[k] -= lij * oct_bcast(lij, k);
    }
    if (act) {
      double *L = sm + CS_LL + 8 * j;
      sts2(L, a[0], a[1]); sts2(L + 2, a[2], a[3]); sts2(L + 4, a[4], a[5]); sts2(L + 6, a[6], 0.0);
      sm[CS_LL + 56 + j] = dinv;
    }
    __syncwarp();
    const double *LL = sm + CS_LL, *DI = sm + CS_LL + 56;
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      double s = (i == jj) ? 1.0 : 0.0;
#pragma unroll
      for (int k = 0; k < i; ++k) s -= LL[i * 8 + k] * mi[k];
      mi[i] = s * DI[i];
    }
#pragma unroll
    for (int i = NJ - 1; i >= 0; --i) {
      double s = mi[i];
#pragma unroll
      for (int k = i + 1; k < NJ; ++k) s -= LL[k * 8 + i] * mi[k];
      mi[i] = s * DI[i];
    }
  }
  double rhs0 = 0.0;
  {
    const double d = in.u - nle;
#pragma unroll
    for (int k = 0; k < NJ; ++k) rhs0 += mi[k] * oct_bcast(d, k);
  }

  // ---------------- contact KKT ----------------
  double a = rhs0;
  double lamv[3] = {0.0, 0.0, 0.0};
  double X[3] = {0.0, 0.0, 0.0};  // X[r] = (Minv Jc^T)[j][r]
  double Sinv[3][3];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int s2 = 0; s2 < 3; ++s2) Sinv[r][s2] = 0.0;
  const double z_target = in.p_ref[2] - cfg.z_press;
  if (NC != 0) {
    const double kp = cfg.contact_gains[0], kd = cfg.contact_gains[1];
    const v3 pc = mk3(in.p_ref[0], in.p_ref[1], z_target);
    const v3 a0v = a_cl + (pf - pc) * kp + vf * kd;
#pragma unroll
    for (int k = 0; k < NJ; ++k) {
      const v3 Jfk = lds_v3p(sm + CS_JF + 4 * k);
#pragma unroll
      for (int r = 0; r < NC; ++r) X[r] += mi[k] * comp(Jfk, (NC == 1) ? 2 : r);
    }
    const double m1 = zero_or_one;
    double S[3][3], b[3];
#pragma unroll
    for (int r = 0; r < NC; ++r) {
      const int ax = (NC == 1) ? 2 : r;
#pragma unroll
      for (int s2 = r; s2 < NC; ++s2) {
        S[r][s2] = oct_sum(m1 * comp(Jf, ax) * X[s2]);
        S[s2][r] = S[r][s2];
      }
      S[r][r] += cfg.contact_inv_damping;
      b[r] = -(comp(a0v, ax) + oct_sum(m1 * comp(Jf, ax) * rhs0));
    }
    if (NC == 1) {
      Sinv[0][0] = 1.0 / S[0][0];
    } else {
      const double c00 = S[1][1] * S[2][2] - S[1][2] * S[2][1], c01 = S[1][2] * S[2][0] - S[1][0] * S[2][2],
                   c02 = S[1][0] * S[2][1] - S[1][1] * S[2][0];
      const double idet = 1.0 / (S[0][0] * c00 + S[0][1] * c01 + S[0][2] * c02);
      Sinv[0][0] = c00 * idet;
      Sinv[0][1] = Sinv[1][0] = c01 * idet;
      Sinv[0][2] = Sinv[2][0] = c02 * idet;
      Sinv[1][1] = (S[0][0] * S[2][2] - S[0][2] * S[2][0]) * idet;
      Sinv[1][2] = Sinv[2][1] = (S[0][2] * S[1][0] - S[0][0] * S[1][2]) * idet;
      Sinv[2][2] = (S[0][0] * S[1][1] - S[0][1] * S[1][0]) * idet;
    }
    double lam[3];
#pragma unroll
    for (int r = 0; r < NC; ++r) {
      double s = 0.0;
#pragma unroll
      for (int s2 = 0; s2 < NC; ++s2) s += Sinv[r][s2] * b[s2];
      lam[r] = s;
      a += X[r] * s;
    }
    if (NC == 1) lamv[2] = lam[0];
    else { lamv[0] = lam[0]; lamv[1] = lam[1]; lamv[2] = lam[2]; }
  }
  if (!in.with_control) {  // classical terminal knot: kinematics + costs only
    a = 0.0;
    lamv[0] = lamv[1] = lamv[2] = 0.0;
  }
  a_out = a;
  lam_out[0] = lamv[0]; lam_out[1] = lamv[1]; lam_out[2] = lamv[2];

  // ---------------- dynamics derivatives: own columns q_j, v_j, u_j ----------------
  double dl[3] = {0.0, 0.0, 0.0};  // d lambda_normal / d (q_j, v_j, u_j)
  // friction cone (point3d, w_friction_cone > 0): the tangential force components enter the cost too
  const bool fric = (NC == 3) && cfg.w_friction_cone > 0.0;
  double dlt[2][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};  // d lambda_x, d lambda_y / d (q_j, v_j, u_j)
  if (DIFF) {
    const sv acc = scan_up6(J * (act ? a : 0.0), j);
    const sv at = ab + acc;
    const sv F = fb + si_mul(Y, acc);
    const sv at_prev = prev6(at, j), v_prev = prev6(v, j);
    const sv dAdq = mxm(at_prev + ag, J) + mxm(v_prev, dVdq);
    // body B operator, then composite (suffix sums)
    bop Bc;
    {
      const sv h = si_mul(Y, v);
      const v3 w = v.a, vl = v.l, hm = Y.h;
      const double *I = Y.I;
      const double I9[9] = {I[0], I[1], I[2], I[1], I[3], I[4], I[2], I[4], I[5]};
      const double wx[9] = {0.0, -w.z, w.y, w.z, 0.0, -w.x, -w.y, w.x, 0.0};
      const double hv[3] = {hm.x, hm.y, hm.z}, vv[3] = {vl.x, vl.y, vl.z};
      const double hax[9] = {0.0, -h.a.z, h.a.y, h.a.z, 0.0, -h.a.x, -h.a.y, h.a.x, 0.0};
      const double vh = dot(vl, hm);
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          double s = 0.0;
#pragma unroll
          for (int k = 0; k < 3; ++k) s += wx[3 * r + k] * I9[3 * k + c] - I9[3 * r + k] * wx[3 * k + c];
          s -= hv[r] * vv[c] + vv[r] * hv[c];
          if (r == c) s += 2.0 * vh;
          s -= hax[3 * r + c];
          Bc.A[3 * r + c] = oct_scan_down(s, j);
        }
      Bc.hl = scan_down3(h.l, j);
    }
    const sv Fc = scan_down6(F, j);
    const sv dFdq = si_mul(Yc, dAdq) + bop_mul(Bc, dVdq) + mxf(J, Fc);
    const sv dVdq2 = dVdq * 2.0;
    const sv dFdv = si_mul(Yc, dVdq2) + bop_mul(Bc, J);
    const sv BTJ = bop_mulT(Bc, J);
    if (act) sts_v3p(sm + CS_BT + 4 * j, BTJ.a);
    __syncwarp();
    double tq[NJ], tv[NJ];
    const v3 lw = mk3(lamv[0], lamv[1], lamv[2]);
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      const bool up = i <= jj;
      const sv o = lds_sv(sm + (up ? CS_J : CS_DF) + 6 * i);
      const v3 bt = lds_v3p(sm + CS_BT + 4 * i);
      double sq = dot6(o, sel6(up, dFdq, dAdq));
      double sv_ = dot6(o, sel6(up, dFdv, dVdq2));
      if (!up) {
        sq += dot(bt, dVdq.a);
        sv_ += dot(bt, J.a);
      }
      if (NC != 0) {
        // d(Jc^T lam)/dq : lam_w . (z_min x Jf_max), symmetric
        const v3 Jfi = lds_v3p(sm + CS_JF + 4 * i);
        const v3 zmin = up ? o.a : J.a;
        const v3 jmax = up ? Jf : Jfi;
        sq -= dot(lw, cross(zmin, jmax));
      }
      tq[i] = sq;
      tv[i] = sv_;
    }
    double Gq[NJ], Gv[NJ], Gu[NJ];
    if (NC == 0) {
      if (act) {
        double *T = sm + CS_TL + 8 * j;
        sts2(T, mi[0], mi[1]); sts2(T + 2, mi[2], mi[3]); sts2(T + 4, mi[4], mi[5]); sts2(T + 6, mi[6], 0.0);
      }
      __syncwarp();
#pragma unroll
      for (int i = 0; i < NJ; ++i) {
        const double *T = sm + CS_TL + 8 * i;
        double t[8];
        lds2(T, t[0], t[1]); lds2(T + 2, t[2], t[3]); lds2(T + 4, t[4], t[5]); lds2(T + 6, t[6], t[7]);
        double sq = 0.0, sv_ = 0.0;
#pragma unroll
        for (int k = 0; k < NJ; ++k) {
          sq += t[k] * tq[k];
          sv_ += t[k] * tv[k];
        }
        Gq[i] = -sq;
        Gv[i] = -sv_;
        Gu[i] = mi[i];
      }
    } else {
      // constraint-acceleration derivatives for the own columns
      const double kp = cfg.contact_gains[0], kd = cfg.contact_gains[1];
      const sv a7 = bcast6(at, NJ - 1);
      v3 dcq, dcv;
      {
        const sv dv = mxm(J, v7 - v);
        const sv da = mxm(J, a7 - at) + mxm(dVdq, v7 - v);
        const v3 dvp = dv.l + cross(dv.a, pf) + cross(wf, Jf);
        dcq = da.l + cross(da.a, pf) + cross(a7.a, Jf) + cross(dv.a, vf) + cross(wf, dvp) + Jf * kp + dvp * kd;
        const sv da2 = mxm(J, v7 - v * 2.0);
        const v3 dvp2 = J.l + cross(J.a, pf);
        dcv = da2.l + cross(da2.a, pf) + cross(J.a, vf) + cross(wf, dvp2) + dvp2 * kd;
      }
      // tr = X Sinv (own row), gathered over the octet; tl = Minv - tr X^T (own column == row)
      double tr_own[3] = {0.0, 0.0, 0.0};
#pragma unroll
      for (int r = 0; r < NC; ++r)
#pragma unroll
        for (int s2 = 0; s2 < NC; ++s2) tr_own[r] += X[s2] * Sinv[s2][r];
      constexpr int NCm = NC > 0 ? NC : 1;
      double trall[NJ][NCm];
#pragma unroll
      for (int i = 0; i < NJ; ++i)
#pragma unroll
        for (int r = 0; r < NC; ++r) trall[i][r] = oct_bcast(tr_own[r], i);
      double tl[NJ];
#pragma unroll
      for (int i = 0; i < NJ; ++i) {
        double s = mi[i];
#pragma unroll
        for (int r = 0; r < NC; ++r) s -= trall[i][r] * X[r];
        tl[i] = s;
      }
      if (act) {
        double *T = sm + CS_TL + 8 * j;
        sts2(T, tl[0], tl[1]); sts2(T + 2, tl[2], tl[3]); sts2(T + 4, tl[4], tl[5]); sts2(T + 6, tl[6], 0.0);
      }
      __syncwarp();
#pragma unroll
      for (int i = 0; i < NJ; ++i) {
        const double *T = sm + CS_TL + 8 * i;
        double t[8];
        lds2(T, t[0], t[1]); lds2(T + 2, t[2], t[3]); lds2(T + 4, t[4], t[5]); lds2(T + 6, t[6], t[7]);
        double sq = 0.0, sv_ = 0.0;
#pragma unroll
        for (int k = 0; k < NJ; ++k) {
          sq += t[k] * tq[k];
          sv_ += t[k] * tv[k];
        }
#pragma unroll
        for (int r = 0; r < NC; ++r) {
          const int ax = (NC == 1) ? 2 : r;
          sq += trall[i][r] * comp(dcq, ax);
          sv_ += trall[i][r] * comp(dcv, ax);
        }
        Gq[i] = -sq;
        Gv[i] = -sv_;
        Gu[i] = tl[i];
      }
      // d lam_normal / dz for the own columns: tr^T dtau_dx - Sinv da0_dx ; d/du = -tr^T
      {
        constexpr int rr = (NC == 1) ? 0 : 2;
        double sq = 0.0, sv_ = 0.0;
#pragma unroll
        for (int k = 0; k < NJ; ++k) {
          sq += trall[k][rr] * tq[k];
          sv_ += trall[k][rr] * tv[k];
        }
#pragma unroll
        for (int s2 = 0; s2 < NC; ++s2) {
          const int ax2 = (NC == 1) ? 2 : s2;
          sq -= Sinv[rr][s2] * comp(dcq, ax2);
          sv_ -= Sinv[rr][s2] * comp(dcv, ax2);
        }
        dl[0] = sq;
        dl[1] = sv_;
        dl[2] = -tr_own[rr];
      }
      if constexpr (NC == 3) {
        if (fric) {
#pragma unroll
          for (int rr = 0; rr < 2; ++rr) {
            double sq = 0.0, sv_ = 0.0;
#pragma unroll
            for (int k = 0; k < NJ; ++k) {
              sq += trall[k][rr] * tq[k];
              sv_ += trall[k][rr] * tv[k];
            }
#pragma unroll
            for (int s2 = 0; s2 < NCm; ++s2) {
              sq -= Sinv[rr][s2] * comp(dcq, s2);
              sv_ -= Sinv[rr][s2] * comp(dcv, s2);
            }
            dlt[rr][0] = sq;
            dlt[rr][1] = sv_;
            dlt[rr][2] = -tr_own[rr];
          }
        }
      }
    }
    if (!in.with_control) {
#pragma unroll
      for (int i = 0; i < NJ; ++i) Gq[i] = Gv[i] = Gu[i] = 0.0;
      dl[0] = dl[1] = dl[2] = 0.0;
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) dlt[rr][0] = dlt[rr][1] = dlt[rr][2] = 0.0;
    }
    if (wr && act) {
      // the record holds Gh = dt * d a/d z in the operand layout of the Riccati MMAs: the eight entries (k = 0..7) of
      // column 7a + j are one 64-byte chunk [k & 3][k >> 2] (rec_gh), written as four 128-bit stores per tile
      const double dtg = cfg.dt_ocp;
      double *gh = rec + REC_GH + 8 * j;
      sts2(gh + 0, dtg * Gq[0], dtg * Gq[4]); sts2(gh + 2, dtg * Gq[1], dtg * Gq[5]);
      sts2(gh + 4, dtg * Gq[2], dtg * Gq[6]); sts2(gh + 6, dtg * Gq[3], 0.0);
      sts2(gh + 64, dtg * Gv[0], dtg * Gv[4]); sts2(gh + 66, dtg * Gv[1], dtg * Gv[5]);
      sts2(gh + 68, dtg * Gv[2], dtg * Gv[6]); sts2(gh + 70, dtg * Gv[3], 0.0);
      sts2(gh + 128, dtg * Gu[0], dtg * Gu[4]); sts2(gh + 130, dtg * Gu[1], dtg * Gu[5]);
      sts2(gh + 132, dtg * Gu[2], dtg * Gu[6]); sts2(gh + 134, dtg * Gu[3], 0.0);
    }
  }

  // ---------------- cost stack ----------------
  const bool contact = (NC != 0);
  const bool state_reg = (cfg.variant == FFMPC_VARIANT_CLASSICAL) || cfg.use_inner_state_reg;
  const bool tau_reg = ((cfg.variant == FFMPC_VARIANT_CLASSICAL) || cfg.use_inner_tau_reg) && !in.terminal_costs;
  double cp = 0.0;                       // own-joint cost terms
  double gq = 0.0, gv = 0.0, gu = 0.0;   // gradient accumulators of the own components
  double hq = 0.0, hv_ = 0.0, hu = 0.0;  // diagonal Hessian accumulators
  if (state_reg) {
    const double rq = in.q - in.xr_q, rv = qd - in.xr_v;
    cp += 0.5 * cfg.w_posture * (rq * rq + rv * rv);
    const double wv = cfg.w_v * cj[CT_VDAMP];
    cp += 0.5 * wv * qd * qd;
    gq += cfg.w_posture * rq;
    gv += cfg.w_posture * rv + wv * qd;
    hq += cfg.w_posture;
    hv_ += cfg.w_posture + wv;
  }
  if (cfg.w_q_soft_limits > 0.0) {
    double av, Ar, Arr;
    barrier_act(in.q - cj[CT_QSREF], cj[CT_QSLB], cj[CT_QSUB], av, Ar, Arr);
    cp += cfg.w_q_soft_limits * av;
    gq += cfg.w_q_soft_limits * Ar;
    hq += cfg.w_q_soft_limits * Arr;
  }
  if (tau_reg) {
    const double ui = in.with_control ? in.u : 0.0;
    const double r = in.with_control ? ui - in.tau_ref : 0.0;
    cp += 0.5 * cfg.w_tau * r * r;
    if (in.with_control) {
      gu += cfg.w_tau * r;
      hu += cfg.w_tau;
    }
    if (cfg.w_tau_soft_limits > 0.0) {
      double mn = cfg.tau_limits[0];
#pragma unroll
      for (int k = 1; k < NJ; ++k) mn = fmin(mn, cfg.tau_limits[k]);
      const double margin = fmin(fmax(cfg.tau_soft_limit_margin, 0.0), mn - 1.0e-6);
      double av, Ar, Arr;
      barrier_act(ui, -cj[CT_TAULIM] + margin, cj[CT_TAULIM] - margin, av, Ar, Arr);
      cp += cfg.w_tau_soft_limits * av;
      if (in.with_control) {
        gu += cfg.w_tau_soft_limits * Ar;
        hu += cfg.w_tau_soft_limits * Arr;
      }
    }
  }
  cpart = act ? cp : 0.0;

  // frame terms (octet-uniform): rows 0-2 orientation log, 3-5 translation, 6-8 linear velocity, 9-11 angular velocity
  double cu = 0.0;
  double rw[12], rg[12];
  v3 rori;
  {
    m3 Re;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int k = 0; k < 3; ++k) Re.m[3 * i + k] = cfg.R_des[i] * Rf.m[k] + cfg.R_des[3 + i] * Rf.m[3 + k] + cfg.R_des[6 + i] * Rf.m[6 + k];
    const double tr = Re.m[0] + Re.m[4] + Re.m[8];
    double t;
    if (tr >= 3.0 - 1e-14) {
      t = 0.5;
    } else {
      const double ca = fmax((tr - 1.0) * 0.5, -1.0);
      const double th = acos(ca);
      t = (th > 1e-7) ? th / (2.0 * sin(th)) : 0.5;
    }
    rori = mk3((Re.m[7] - Re.m[5]) * t, (Re.m[2] - Re.m[6]) * t, (Re.m[3] - Re.m[1]) * t);
  }
  {
    const v3 dpr = pf - mk3(in.p_ref[0], in.p_ref[1], in.p_ref[2]);
    const double r3[3] = {rori.x, rori.y, rori.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      rw[k] = cfg.w_ee_ori * cfg.ori_weights[k];
      rg[k] = rw[k] * r3[k];
      cu += 0.5 * rw[k] * r3[k] * r3[k];
    }
    double rt[3], wt[3], rv[3], wvv[3];
    if (!contact) {
      rt[0] = dpr.x; rt[1] = dpr.y; rt[2] = dpr.z;
      wt[0] = cfg.w_ee_pos; wt[1] = cfg.w_ee_pos; wt[2] = 2.5 * cfg.w_ee_pos;
      rv[0] = rv[1] = rv[2] = 0.0;
      wvv[0] = wvv[1] = wvv[2] = 0.0;
    } else {
      rt[0] = dpr.x; rt[1] = dpr.y; rt[2] = pf.z - z_target;
      wt[0] = cfg.w_tangent_pos; wt[1] = cfg.w_tangent_pos; wt[2] = cfg.w_plane_z > 0.0 ? cfg.w_plane_z : 0.0;
      rv[0] = vf.x - in.v_ref[0]; rv[1] = vf.y - in.v_ref[1]; rv[2] = vf.z;
      wvv[0] = cfg.w_tangent_vel; wvv[1] = cfg.w_tangent_vel; wvv[2] = cfg.w_vz > 0.0 ? cfg.w_vz : 0.0;
    }
    const double wa[3] = {wf.x, wf.y, wf.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      rw[3 + k] = wt[k];
      rg[3 + k] = wt[k] * rt[k];
      cu += 0.5 * wt[k] * rt[k] * rt[k];
      rw[6 + k] = wvv[k];
      rg[6 + k] = wvv[k] * rv[k];
      cu += 0.5 * wvv[k] * rv[k] * rv[k];
      rw[9 + k] = cfg.w_wdamp * cfg.w_wdamp_weights[k];
      rg[9 + k] = rw[9 + k] * wa[k];
      cu += 0.5 * rw[9 + k] * wa[k] * wa[k];
    }
  }
  double fw = 0.0, fg = 0.0;                  // force-space Hessian zz / gradient z
  double fgt[2] = {0.0, 0.0};                 // gradient x, y            (friction cone only)
  double fwt[4] = {0.0, 0.0, 0.0, 0.0};       // Hessian xx, yy, xz, yz   (friction cone only)
  if (contact) {
    const double lz = in.with_control ? lamv[2] : 0.0;
    if (cfg.w_unilateral > 0.0) {
      double av, Ar, Arr;
      barrier_act(lz, cfg.friction_margin, INFINITY, av, Ar, Arr);
      cu += cfg.w_unilateral * av;
      fg += cfg.w_unilateral * Ar;
      fw += cfg.w_unilateral * Arr;
    }
    if (cfg.w_fn > 0.0) {
      const double r = in.with_control ? lz - cfg.fn_des : 0.0;
      cu += 0.5 * cfg.w_fn * r * r;
      fg += cfg.w_fn * r;
      fw += cfg.w_fn;
    }
    // "friction_cone" (crocoddyl_classical.py:678-686): r = A lambda over FrictionCone(I, mu, 4, false), barrier with
    // ub = -margin on the facet rows (+-1, 0, -mu), (0, +-1, -mu) and lb = +margin on the normal row (:891-904)
    if constexpr (NC == 3) {
      if (fric) {
        const double eps = fmax(cfg.friction_margin, 0.0), wfc = cfg.w_friction_cone, mu = cfg.mu;
        const double lx = in.with_control ? lamv[0] : 0.0, ly = in.with_control ? lamv[1] : 0.0;
#pragma unroll
        for (int f = 0; f < 4; ++f) {
          const double cx = (f == 0) ? 1.0 : (f == 2) ? -1.0 : 0.0, cy = (f == 1) ? 1.0 : (f == 3) ? -1.0 : 0.0;
          double av, Ar, Arr;
          barrier_act(cx * lx + cy * ly - mu * lz, -INFINITY, -eps, av, Ar, Arr);
          cu += wfc * av;
          const double g = wfc * Ar, h = wfc * Arr;
          fgt[0] += g * cx; fgt[1] += g * cy; fg -= g * mu;
          fwt[0] += h * cx * cx; fwt[1] += h * cy * cy;          // xx, yy (xy is identically 0: cx cy = 0 on every facet)
          fwt[2] -= h * cx * mu; fwt[3] -= h * cy * mu;          // xz, yz
          fw += h * mu * mu;
        }
        double av, Ar, Arr;
        barrier_act(lz, eps, INFINITY, av, Ar, Arr);
        cu += wfc * av;
        fg += wfc * Ar;
        fw += wfc * Arr;
      }
    }
    if (!in.with_control) { fg = 0.0; fw = 0.0; fgt[0] = fgt[1] = 0.0; fwt[0] = fwt[1] = fwt[2] = fwt[3] = 0.0; }
  }
  cuni = cu;

  if (DIFF) {
    // own Jacobian columns of the frame residuals: Jq[r] (q_j), Jv[r] rows 6..11 (v_j; rows 0-5 are zero)
    double Jq[12], Jv[6];
    {
      const double th2 = dot(rori, rori);
      double cc;
      if (th2 < 1e-10) {
        cc = 1.0 / 12.0;
      } else {
        const double th = sqrt(th2);
        cc = 1.0 / th2 - (1.0 + cos(th)) / (2.0 * th * sin(th));
      }
      const double rx[9] = {0.0, -rori.z, rori.y, rori.z, 0.0, -rori.x, -rori.y, rori.x, 0.0};
      double Jl[9];
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          double s2 = 0.0;
#pragma unroll
          for (int k = 0; k < 3; ++k) s2 += rx[3 * r + k] * rx[3 * k + c];
          Jl[3 * r + c] = (r == c ? 1.0 : 0.0) + 0.5 * rx[3 * r + c] + cc * s2;
        }
      // o = Jlog3 * (Rf^T z_j)
      const v3 rz = mulTv(Rf, J.a);
      Jq[0] = Jl[0] * rz.x + Jl[1] * rz.y + Jl[2] * rz.z;
      Jq[1] = Jl[3] * rz.x + Jl[4] * rz.y + Jl[5] * rz.z;
      Jq[2] = Jl[6] * rz.x + Jl[7] * rz.y + Jl[8] * rz.z;
      Jq[3] = Jf.x; Jq[4] = Jf.y; Jq[5] = Jf.z;
      const sv dv = mxm(J, v7 - v);
      const v3 dvp = dv.l + cross(dv.a, pf) + cross(wf, Jf);
      Jq[6] = dvp.x; Jq[7] = dvp.y; Jq[8] = dvp.z;
      Jq[9] = dv.a.x; Jq[10] = dv.a.y; Jq[11] = dv.a.z;
      Jv[0] = Jf.x; Jv[1] = Jf.y; Jv[2] = Jf.z;
      Jv[3] = J.a.x; Jv[4] = J.a.y; Jv[5] = J.a.z;
    }
    // gradient, own components
    {
      double sq = gq, sv_ = gv;
#pragma unroll
      for (int r = 0; r < 12; ++r) sq += rg[r] * Jq[r];
#pragma unroll
      for (int r = 0; r < 6; ++r) sv_ += rg[6 + r] * Jv[r];
      if (wr && act) {
        double fgq = fg * dl[0], fgv = fg * dl[1], fgu = fg * dl[2];
        if constexpr (NC == 3) {
          fgq += fgt[0] * dlt[0][0] + fgt[1] * dlt[1][0];
          fgv += fgt[0] * dlt[0][1] + fgt[1] * dlt[1][1];
          fgu += fgt[0] * dlt[0][2] + fgt[1] * dlt[1][2];
        }
        row_store7(rec + rec_lz(0, 0), j, sc * (sq + fgq) + (eg3 ? eg3[0] : 0.0));
        row_store7(rec + rec_lz(1, 0), j, sc * (sv_ + fgv) + (eg3 ? eg3[1] : 0.0));
        row_store7(rec + rec_lz(2, 0), j, sc * (gu + fgu) + (eg3 ? eg3[2] : 0.0));
      }
    }
    // Gauss-Newton Hessian: publish own columns, then every lane forms the entries (row joint i, own columns)
    __syncwarp();  // T / J / DF are dead from here on (HS aliases them)
    if (act) {
      double *H = sm + CS_HS + 22 * j;
#pragma unroll
      for (int r = 0; r < 12; r += 2) sts2(H + r, Jq[r], Jq[r + 1]);
#pragma unroll
      for (int r = 0; r < 6; r += 2) sts2(H + 12 + r, Jv[r], Jv[r + 1]);
      sts2(H + 18, dl[0], dl[1]);
      sts2(H + 20, dl[2], 0.0);
      if constexpr (NC == 3) {
        if (fric) {
          double *HF = sm + CS_HF + 6 * j;
          sts2(HF, dlt[0][0], dlt[0][1]);
          sts2(HF + 2, dlt[0][2], dlt[1][0]);
          sts2(HF + 4, dlt[1][1], dlt[1][2]);
        }
      }
    }
    __syncwarp();
    double wq[12], wv[6];
#pragma unroll
    for (int r = 0; r < 12; ++r) wq[r] = rw[r] * Jq[r];
#pragma unroll
    for (int r = 0; r < 6; ++r) wv[r] = rw[6 + r] * Jv[r];
    double fq = fw * dl[0], fv = fw * dl[1], fu = fw * dl[2];
    // friction cone: (force-space Hessian) x (own columns of d lambda), rows x and y; row z gains the xz / yz couplings
    double ftq[2] = {0.0, 0.0}, ftv[2] = {0.0, 0.0}, ftu[2] = {0.0, 0.0};
    if constexpr (NC == 3) {
      if (fric) {
#pragma unroll
        for (int a2 = 0; a2 < 2; ++a2) {
          ftq[a2] = fwt[a2] * dlt[a2][0] + fwt[2 + a2] * dl[0];
          ftv[a2] = fwt[a2] * dlt[a2][1] + fwt[2 + a2] * dl[1];
          ftu[a2] = fwt[a2] * dlt[a2][2] + fwt[2 + a2] * dl[2];
          fq += fwt[2 + a2] * dlt[a2][0];
          fv += fwt[2 + a2] * dlt[a2][1];
          fu += fwt[2 + a2] * dlt[a2][2];
        }
      }
    }
    const double ehq = eh3 ? eh3[0] : 0.0, ehv = eh3 ? eh3[1] : 0.0, ehu = eh3 ? eh3[2] : 0.0;
    double pend_q = 0.0, pend_v = 0.0, pend_u = 0.0;  // even-row entries of the diagonal tiles waiting for their pair
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      const double *H = sm + CS_HS + 22 * i;
      double hqv[12], hvv[6], di[4];
#pragma unroll
      for (int r = 0; r < 12; r += 2) lds2(H + r, hqv[r], hqv[r + 1]);
#pragma unroll
      for (int r = 0; r < 6; r += 2) lds2(H + 12 + r, hvv[r], hvv[r + 1]);
      lds2(H + 18, di[0], di[1]);
      lds2(H + 20, di[2], di[3]);
      double sqq = di[0] * fq, sqv = di[0] * fv, svv = di[1] * fv;
#pragma unroll
      for (int r = 0; r < 12; ++r) sqq += hqv[r] * wq[r];
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        sqv += hqv[6 + r] * wv[r];
        svv += hvv[r] * wv[r];
      }
      double squ = di[0] * fu, svu = di[1] * fu, suu = di[2] * fu;
      if constexpr (NC == 3) {
        if (fric) {
          const double *HF = sm + CS_HF + 6 * i;
          double dt_[6];
          lds2(HF, dt_[0], dt_[1]); lds2(HF + 2, dt_[2], dt_[3]); lds2(HF + 4, dt_[4], dt_[5]);
#pragma unroll
          for (int a2 = 0; a2 < 2; ++a2) {
            sqq += dt_[3 * a2] * ftq[a2];
            sqv += dt_[3 * a2] * ftv[a2];
            svv += dt_[3 * a2 + 1] * ftv[a2];
            squ += dt_[3 * a2] * ftu[a2];
            svu += dt_[3 * a2 + 1] * ftu[a2];
            suu += dt_[3 * a2 + 2] * ftu[a2];
          }
        }
      }
      if (wr && act) {
        const bool diag = (i == j);
        // Tiles are row-major 8x8 (the MMA accumulator layout).  Off-diagonal tiles: entry (i, j), lanes contiguous.
        // Diagonal tiles are symmetric: the lane writes ROW j (= its column) as 128-bit pairs (i, i + 1).
        const double hqq = sc * (sqq + (diag ? hq : 0.0)) + (diag ? ehq : 0.0);
        const double hvv = sc * (svv + (diag ? hv_ : 0.0)) + (diag ? ehv : 0.0);
        const double huu = sc * (suu + (diag ? hu : 0.0)) + (diag ? ehu : 0.0);
        row_store7(rec + rec_lzz(0, i, 1, 0), j, sc * sqv);
        row_store7(rec + rec_lzz(0, i, 2, 0), j, sc * squ);
        row_store7(rec + rec_lzz(1, i, 2, 0), j, sc * svu);
        if (i & 1) {
          sts2(rec + rec_lzz(0, j, 0, i - 1), pend_q, hqq);
          sts2(rec + rec_lzz(1, j, 1, i - 1), pend_v, hvv);
          sts2(rec + rec_lzz(2, j, 2, i - 1), pend_u, huu);
        } else if (i == NJ - 1) {
          sts2(rec + rec_lzz(0, j, 0, i), hqq, 0.0);
          sts2(rec + rec_lzz(1, j, 1, i), hvv, 0.0);
          sts2(rec + rec_lzz(2, j, 2, i), huu, 0.0);
        }
        pend_q = hqq; pend_v = hvv; pend_u = huu;
      }
    }
  }
}

// Per-lane state of a knot for the integrated (Euler [+ LPF]) action model.
struct CoopX {
  double q, v, t;  // own components of x: q_j, v_j, tau_hat_j (FF only)
};
struct CoopCtx {
  const double *p_ref, *v_ref;  // [3] of this knot (octet-uniform pointers)
  double xr_q, xr_v, tau_ref;   // own components of x_reg_ref / tau_ref
  double yr_q, yr_v, yr_t;      // own components of y_ref = y0 (FF)
  bool terminal;
};

// ActionModel.calc (+ calcDiff when DIFF), cooperative.  u = own component of the control (classical: torque, FF: w).
// Returns the own components of xnext, the octet-uniform knot cost and contact force.
template <int NC, bool FF, bool DIFF>
FFMPC_DEV void coop_action_node(const ffmpc_model &mdl, const ffmpc_config &cfg, const double *ct, double *sm, int j, const CoopCtx &ctx,
                                const CoopX &x, double u, CoopX &xn, double &cost_out, double *lam, double *rec, bool wr) {
  const int jj = j < NJ ? j : NJ - 1;
  const bool act = j < NJ;
  const double *cj = ct + jj * CT_STRIDE;
  const double dt = cfg.dt_ocp;
  CoopIn in;
  in.q = x.q;
  in.qd = x.v;
  in.u = FF ? x.t : u;
  in.xr_q = ctx.xr_q;
  in.xr_v = ctx.xr_v;
  in.tau_ref = ctx.tau_ref;
  in.p_ref = ctx.p_ref;
  in.v_ref = ctx.v_ref;
  in.terminal_costs = ctx.terminal;
  in.with_control = FF ? true : !ctx.terminal;
  const bool running = in.with_control;
  const double w_y = FF ? fmax(cfg.w_y, 0.0) : 0.0;
  double eg3[3] = {0.0, 0.0, 0.0}, eh3[3] = {0.0, 0.0, 0.0};
  double extra = 0.0;  // own part of the LPF wrapper's cost terms
  if (FF) {
    const double wq2 = cj[CT_YWQ] * cj[CT_YWQ], wv2 = cj[CT_YWV] * cj[CT_YWV], wt2 = cj[CT_YWT] * cj[CT_YWT];
    const double dq = x.q - ctx.yr_q, dv = x.v - ctx.yr_v, dtau = x.t - ctx.yr_t;
    extra += 0.5 * w_y * (wq2 * dq * dq + wv2 * dv * dv + wt2 * dtau * dtau);
    eg3[0] = w_y * wq2 * dq; eg3[1] = w_y * wv2 * dv; eg3[2] = w_y * wt2 * dtau;
    eh3[0] = w_y * wq2; eh3[1] = w_y * wv2; eh3[2] = w_y * wt2;
  }
  double a, cpart, cuni;
  const bool have_e = FF && w_y > 0.0;
  coop_dam<NC, DIFF>(mdl, cfg, ct, sm, j, in, a, lam, cpart, cuni, rec, wr, running ? dt : 1.0, have_e ? eg3 : nullptr,
                     have_e ? eh3 : nullptr);
  if (!(FF && w_y > 0.0)) extra = 0.0;
  double scale = 1.0;
  if (running) {
    xn.q = x.q + (dt * x.v + dt * dt * a);
    xn.v = x.v + dt * a;
    scale = dt;
  } else {
    xn.q = x.q;
    xn.v = x.v;
  }
  xn.t = 0.0;
  if (FF) {
    const double alpha = fmin(fmax(cfg.ff_alpha, 0.0), 0.999999), beta = 1.0 - alpha;
    const double w_w = fmax(cfg.w_w, 0.0), w_soft = fmax(cfg.w_w_soft_limits, 0.0);
    const double w = ctx.terminal ? 0.0 : u;
    xn.t = alpha * x.t + beta * w;
    double c = 0.0, g = 0.0, h = 0.0;
    if (w_soft > 0.0) {
      // lpf_soft_term with the own joint's limit
      const double margin = fmax(cfg.tau_soft_limit_margin, 0.0);
      const double lim = fmax(cj[CT_TAULIM] - margin, 1.0e-9);
      const double over = fmax(fabs(w) - lim, 0.0);
      const bool active = over > 0.0;
      c = 0.5 * over * over;
      g = active ? over * (w > 0.0 ? 1.0 : (w < 0.0 ? -1.0 : 0.0)) : 0.0;
      h = active ? 1.0 : 0.0;
    }
    if (w_w > 0.0) extra += 0.5 * w_w * w * w;
    if (w_soft > 0.0) extra += w_soft * c;
    if (DIFF && wr && act) {
      row_store7(rec + REC_LW, j, w_w * w + w_soft * g);
      row_store7(rec + REC_LWW, j, w_w + w_soft * h);
    }
  }
  cost_out = oct_sum(act ? (scale * cpart + extra) : 0.0) + scale * cuni;
}

}  // namespace ffmpc
