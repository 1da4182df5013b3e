// ffmpc_node.cuh -- device math of ONE action-model knot (calc + analytic calcDiff).
//
// Replaces, for the 7-DoF Panda chain, what the reference gets from Crocoddyl/Pinocchio at
//   crocoddyl_classical.py:558-728 (_make_dam: DifferentialActionModelFree/ContactFwdDynamics + cost stack)
//   crocoddyl_classical.py:542     (IntegratedActionModelEuler)
//   crocoddyl_force_feedback.py:211-290 (_AugmentedLPFActionModel.calc / calcDiff)
// Math spec: SURVEY.md App. A.1-A.6.
//
// One CUDA thread evaluates one knot.  Everything is expressed in the WORLD frame at the world
// origin (Carpentier & Mansard-style recursions), spatial vectors are [linear; angular]:
//   v_i = v_{i-1} + J_i qd_i                       J_i = [p_i x z_i ; z_i]
//   a_i = a_{i-1} + J_i qdd_i + v_i x J_i qd_i     a_0 = -gravity
//   f_i = Y_i a_i + v_i x* (Y_i v_i)               tau_i = J_i . sum_{k>=i} f_k
// and the analytic derivatives
//   dtau_i/dq_j = J_i . (Ycrb_j dAdq_j + Bcrb_j dVdq_j + J_j x* F_j)      (i <= j)
//               = (Ycrb_i J_i) . dAdq_j + (Bcrb_i^T J_i) . dVdq_j         (i >  j)
//   dtau_i/dv_j = J_i . (Ycrb_j dAdv_j + Bcrb_j J_j)                      (i <= j)
//               = (Ycrb_i J_i) . dAdv_j + (Bcrb_i^T J_i) . J_j            (i >  j)
// with dVdq_j = v_j x J_j, dAdq_j = a_{j-1} x J_j + v_{j-1} x dVdq_j, dAdv_j = 2 dVdq_j and
// B_k u = v_k x* (Y_k u) - Y_k (v_k x u) + u x* (Y_k v_k), which collapses to
//   B_k = [[0, -2 [hl]x], [0, Baa]],  Baa = [w]x I_O - I_O [w]x - (hm vl^T + vl hm^T - 2 (vl.hm) 1) - [ha]x.
// The contact-constrained derivatives follow SURVEY App. A.3 (KKT inverse blocks).
// The code is plain scalar C++ so that tests/host_emul can compile it for the host and compare it
// against the oracle without a GPU; the product only ever runs it inside the CUDA kernels.
#pragma once

#include <math.h>

#include "../../include/ffmpc.h"

#ifndef FFMPC_DEV
#ifdef __CUDACC__
#define FFMPC_DEV __device__ __forceinline__
#else
#define FFMPC_DEV inline
#endif
#endif

namespace ffmpc {

constexpr int NJ = 7;
constexpr int NZ = 21;
constexpr int NZZ = NZ * (NZ + 1) / 2;  // packed upper triangle of the 21x21 joint Hessian

// packed upper-triangular index, i <= j
FFMPC_DEV int sym_idx(int i, int j) { return i * NZ - (i * (i - 1)) / 2 + (j - i); }
FFMPC_DEV int sym_idx_any(int i, int j) { return i <= j ? sym_idx(i, j) : sym_idx(j, i); }

struct v3 {
  double x, y, z;
};
FFMPC_DEV v3 mk3(double x, double y, double z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
FFMPC_DEV v3 operator+(const v3 &a, const v3 &b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
FFMPC_DEV v3 operator-(const v3 &a, const v3 &b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
FFMPC_DEV v3 operator*(const v3 &a, double s) { return mk3(a.x * s, a.y * s, a.z * s); }
FFMPC_DEV v3 cross(const v3 &a, const v3 &b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
FFMPC_DEV double dot(const v3 &a, const v3 &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
FFMPC_DEV double comp(const v3 &a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

struct m3 {
  double m[9];  // row-major
};
FFMPC_DEV v3 mulv(const m3 &A, const v3 &b) {
  return mk3(A.m[0] * b.x + A.m[1] * b.y + A.m[2] * b.z, A.m[3] * b.x + A.m[4] * b.y + A.m[5] * b.z,
             A.m[6] * b.x + A.m[7] * b.y + A.m[8] * b.z);
}
FFMPC_DEV v3 mulTv(const m3 &A, const v3 &b) {
  return mk3(A.m[0] * b.x + A.m[3] * b.y + A.m[6] * b.z, A.m[1] * b.x + A.m[4] * b.y + A.m[7] * b.z,
             A.m[2] * b.x + A.m[5] * b.y + A.m[8] * b.z);
}
FFMPC_DEV m3 mulm(const m3 &A, const m3 &B) {
  m3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.m[3 * i + j] = A.m[3 * i] * B.m[j] + A.m[3 * i + 1] * B.m[3 + j] + A.m[3 * i + 2] * B.m[6 + j];
  return C;
}

struct sv {  // spatial vector [linear; angular]
  v3 l, a;
};
FFMPC_DEV sv mksv(const v3 &l, const v3 &a) { sv r; r.l = l; r.a = a; return r; }
FFMPC_DEV sv operator+(const sv &a, const sv &b) { return mksv(a.l + b.l, a.a + b.a); }
FFMPC_DEV sv operator-(const sv &a, const sv &b) { return mksv(a.l - b.l, a.a - b.a); }
FFMPC_DEV sv operator*(const sv &a, double s) { return mksv(a.l * s, a.a * s); }
FFMPC_DEV sv mxm(const sv &a, const sv &b) { return mksv(cross(a.a, b.l) + cross(a.l, b.a), cross(a.a, b.a)); }   // motion x motion
FFMPC_DEV sv mxf(const sv &m, const sv &f) { return mksv(cross(m.a, f.l), cross(m.a, f.a) + cross(m.l, f.l)); }   // motion x* force
FFMPC_DEV double dot6(const sv &a, const sv &b) { return dot(a.l, b.l) + dot(a.a, b.a); }

struct si {        // spatial inertia about the world origin
  double m;        // mass
  v3 h;            // first moment m*c
  double I[6];     // rotational inertia about the origin: xx xy xz yy yz zz
};
FFMPC_DEV v3 symv(const double *I, const v3 &w) {
  return mk3(I[0] * w.x + I[1] * w.y + I[2] * w.z, I[1] * w.x + I[3] * w.y + I[4] * w.z, I[2] * w.x + I[4] * w.y + I[5] * w.z);
}
FFMPC_DEV sv si_mul(const si &Y, const sv &v) { return mksv(v.l * Y.m + cross(v.a, Y.h), symv(Y.I, v.a) + cross(Y.h, v.l)); }
FFMPC_DEV void si_add(si &A, const si &B) {
  A.m += B.m;
  A.h = A.h + B.h;
#pragma unroll
  for (int i = 0; i < 6; ++i) A.I[i] += B.I[i];
}

// composite "B" operator: B u = (-2 hl x u.a ; Baa u.a),  B^T u = (0 ; 2 hl x u.l + Baa^T u.a)
struct bop {
  v3 hl;
  double A[9];
};
FFMPC_DEV sv bop_mul(const bop &B, const sv &u) {
  return mksv(cross(B.hl, u.a) * (-2.0),
              mk3(B.A[0] * u.a.x + B.A[1] * u.a.y + B.A[2] * u.a.z, B.A[3] * u.a.x + B.A[4] * u.a.y + B.A[5] * u.a.z,
                  B.A[6] * u.a.x + B.A[7] * u.a.y + B.A[8] * u.a.z));
}
FFMPC_DEV sv bop_mulT(const bop &B, const sv &u) {
  const v3 t = cross(B.hl, u.l) * 2.0;
  return mksv(mk3(0.0, 0.0, 0.0),
              mk3(t.x + B.A[0] * u.a.x + B.A[3] * u.a.y + B.A[6] * u.a.z, t.y + B.A[1] * u.a.x + B.A[4] * u.a.y + B.A[7] * u.a.z,
                  t.z + B.A[2] * u.a.x + B.A[5] * u.a.y + B.A[8] * u.a.z));
}

// Activation helpers. value / first / second derivative of the scalar activation.
FFMPC_DEV void barrier_act(double r, double lb, double ub, double &a, double &Ar, double &Arr) {
  const double lo = fmin(r - lb, 0.0), hi = fmax(r - ub, 0.0);
  a = 0.5 * (lo * lo + hi * hi);
  Ar = lo + hi;
  Arr = ((r - lb) <= 0.0 ? 1.0 : 0.0) + ((r - ub) >= 0.0 ? 1.0 : 0.0);
}

struct NodeArgs {
  const double *x;          // [14] (q, v)
  const double *u;          // [7]  inner control (torque)
  const double *p_ref;      // [3]
  const double *v_ref;      // [3]
  const double *x_reg_ref;  // [14]
  const double *tau_ref;    // [7]
  bool terminal_costs;      // drop tau_reg / tau_soft_limits
  bool with_control;        // false: classical terminal node, kinematics + costs only
};

// Differential action model: forward dynamics (free: NC=0, ContactModel1D: NC=1, ContactModel3D: NC=3)
// and cost stack.  Outputs: a[7], lam[3] (world-aligned contact force), cost, and when DIFF
// dadz[7*21] (d a / d(q,v,u)), Lz[21], Lzz[231] (Gauss-Newton, packed upper).  The derivative outputs
// are written exactly once (they may point to global memory): Lz = sc * dL/dz + eg, Lzz = sc * H + diag(eh),
// where sc is the Euler time step (running knots) and eg / eh carry the LPF state regularisation.
template <int NC, bool DIFF>
FFMPC_DEV void dam_node(const ffmpc_model &mdl, const ffmpc_config &cfg, const NodeArgs &in, double *a_out, double *lam_out,
                        double *cost_out, double *dadz, double *Lz, double *Lzz, double sc, const double *eg,
                        const double *eh) {
  sv J[NJ], v[NJ], ab[NJ];
  si Y[NJ];
  v3 p[NJ];
  m3 R;
  const double *q = in.x, *qd = in.x + NJ;

  // ---------------- forward kinematics, velocities, bias accelerations, world inertias -------------
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
    if (i == 0) {
      p[0] = pl;
      R = Rl;
    } else {
      p[i] = p[i - 1] + mulv(R, pl);
      R = mulm(R, Rl);
    }
    const v3 z = mk3(R.m[2], R.m[5], R.m[8]);
    J[i] = mksv(cross(p[i], z), z);
    const sv vj = J[i] * qd[i];
    if (i == 0) {
      v[0] = vj;
      ab[0] = mksv(mk3(0, 0, 0), mk3(0, 0, 0));
    } else {
      v[i] = v[i - 1] + vj;
      ab[i] = ab[i - 1] + mxm(v[i - 1], vj);
    }
    // inertia about the world origin
    const v3 cw = p[i] + mulv(R, mk3(mdl.com[i][0], mdl.com[i][1], mdl.com[i][2]));
    const double *in6 = mdl.inertia[i];  // xx yy zz xy xz yz
    m3 Ib;
    Ib.m[0] = in6[0]; Ib.m[4] = in6[1]; Ib.m[8] = in6[2];
    Ib.m[1] = Ib.m[3] = in6[3];
    Ib.m[2] = Ib.m[6] = in6[4];
    Ib.m[5] = Ib.m[7] = in6[5];
    const m3 RI = mulm(R, Ib);
    const double m = mdl.mass[i];
    const double c2 = dot(cw, cw);
    // Iw = RI R^T (symmetric), I_O = Iw + m (|c|^2 1 - c c^T)
    Y[i].I[0] = RI.m[0] * R.m[0] + RI.m[1] * R.m[1] + RI.m[2] * R.m[2] + m * (c2 - cw.x * cw.x);
    Y[i].I[1] = RI.m[0] * R.m[3] + RI.m[1] * R.m[4] + RI.m[2] * R.m[5] - m * cw.x * cw.y;
    Y[i].I[2] = RI.m[0] * R.m[6] + RI.m[1] * R.m[7] + RI.m[2] * R.m[8] - m * cw.x * cw.z;
    Y[i].I[3] = RI.m[3] * R.m[3] + RI.m[4] * R.m[4] + RI.m[5] * R.m[5] + m * (c2 - cw.y * cw.y);
    Y[i].I[4] = RI.m[3] * R.m[6] + RI.m[4] * R.m[7] + RI.m[5] * R.m[8] - m * cw.y * cw.z;
    Y[i].I[5] = RI.m[6] * R.m[6] + RI.m[7] * R.m[7] + RI.m[8] * R.m[8] + m * (c2 - cw.z * cw.z);
    Y[i].m = m;
    Y[i].h = cw * m;
  }
  // EE frame
  m3 eeR;
#pragma unroll
  for (int i = 0; i < 9; ++i) eeR.m[i] = mdl.ee_rot[i];
  const m3 Rf = mulm(R, eeR);
  const v3 pf = p[NJ - 1] + mulv(R, mk3(mdl.ee_pos[0], mdl.ee_pos[1], mdl.ee_pos[2]));
  const sv v7 = v[NJ - 1];
  const v3 wf = v7.a;
  const v3 vf = v7.l + cross(wf, pf);
  v3 Jf[NJ];  // d pf / d q_j  (LOCAL_WORLD_ALIGNED linear Jacobian column)
#pragma unroll
  for (int j = 0; j < NJ; ++j) Jf[j] = J[j].l + cross(J[j].a, pf);

  const double z_target = in.p_ref[2] - cfg.z_press;
  const sv ag = mksv(mk3(-mdl.gravity[0], -mdl.gravity[1], -mdl.gravity[2]), mk3(0, 0, 0));

  double a[NJ];
  double lamv[3] = {0.0, 0.0, 0.0};
  double Minv[NJ][NJ];
  sv fb[NJ], dFda[NJ];
  double X[3][NJ];     // Minv Jc^T
  double Sinv[3][3];   // (Jc Minv Jc^T + mu)^-1
#pragma unroll
  for (int i = 0; i < NJ; ++i) a[i] = 0.0;

  if (in.with_control) {
    // ---------------- CRBA + bias forces ----------------
    double M[NJ][NJ], nle[NJ];
    {
      si Yc = Y[NJ - 1];
      sv Fs = mksv(mk3(0, 0, 0), mk3(0, 0, 0));
#pragma unroll
      for (int i = NJ - 1; i >= 0; --i) {
        if (i < NJ - 1) si_add(Yc, Y[i]);
        fb[i] = si_mul(Y[i], ab[i] + ag) + mxf(v[i], si_mul(Y[i], v[i]));
        Fs = Fs + fb[i];
        nle[i] = dot6(J[i], Fs);
        dFda[i] = si_mul(Yc, J[i]);
#pragma unroll
        for (int j = 0; j <= i; ++j) {
          M[j][i] = dot6(J[j], dFda[i]);
          M[i][j] = M[j][i];
        }
      }
    }
    // ---------------- M^-1 via Cholesky ----------------
    {
      double L[NJ][NJ], Li[NJ][NJ];
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        double s = M[j][j];
#pragma unroll
        for (int k = 0; k < j; ++k) s -= L[j][k] * L[j][k];
        const double l = sqrt(s);
        const double il = 1.0 / l;
        L[j][j] = l;
        Li[j][j] = il;
#pragma unroll
        for (int i = j + 1; i < NJ; ++i) {
          double t = M[i][j];
#pragma unroll
          for (int k = 0; k < j; ++k) t -= L[i][k] * L[j][k];
          L[i][j] = t * il;
        }
      }
      // Li = L^-1 (lower)
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
          Minv[i][j] = s;
          Minv[j][i] = s;
        }
    }
    double rhs0[NJ];
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      double s = 0.0;
#pragma unroll
      for (int j = 0; j < NJ; ++j) s += Minv[i][j] * (in.u[j] - nle[j]);
      rhs0[i] = s;
    }
    if (NC == 0) {
#pragma unroll
      for (int i = 0; i < NJ; ++i) a[i] = rhs0[i];
    } else {
      // ---------------- contact: [M Jc^T; Jc -mu] [a; -lam] = [tau - nle; -a0] ----------------
      const sv ab7 = ab[NJ - 1];
      const v3 a_cl = ab7.l + cross(ab7.a, pf) + cross(wf, vf);
      const double kp = cfg.contact_gains[0], kd = cfg.contact_gains[1];
      const v3 pc = mk3(in.p_ref[0], in.p_ref[1], z_target);
      const v3 a0v = a_cl + (pf - pc) * kp + vf * kd;
      double b[3];
#pragma unroll
      for (int r = 0; r < NC; ++r) {
        const int ax = (NC == 1) ? 2 : r;
#pragma unroll
        for (int i = 0; i < NJ; ++i) {
          double s = 0.0;
#pragma unroll
          for (int j = 0; j < NJ; ++j) s += Minv[i][j] * comp(Jf[j], ax);
          X[r][i] = s;
        }
      }
      double S[3][3];
#pragma unroll
      for (int r = 0; r < NC; ++r) {
        const int ax = (NC == 1) ? 2 : r;
#pragma unroll
        for (int s2 = 0; s2 < NC; ++s2) {
          double s = 0.0;
#pragma unroll
          for (int i = 0; i < NJ; ++i) s += comp(Jf[i], ax) * X[s2][i];
          S[r][s2] = s;
        }
        S[r][r] += cfg.contact_inv_damping;
        double s = comp(a0v, ax);
#pragma unroll
        for (int i = 0; i < NJ; ++i) s += comp(Jf[i], ax) * rhs0[i];
        b[r] = -s;
      }
      if (NC == 1) {
        Sinv[0][0] = 1.0 / S[0][0];
      } else {
        // symmetric 3x3 inverse by cofactors
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
      }
#pragma unroll
      for (int i = 0; i < NJ; ++i) {
        double s = rhs0[i];
#pragma unroll
        for (int r = 0; r < NC; ++r) s += X[r][i] * lam[r];
        a[i] = s;
      }
      if (NC == 1) lamv[2] = lam[0];
      else { lamv[0] = lam[0]; lamv[1] = lam[1]; lamv[2] = lam[2]; }
    }
  }
#pragma unroll
  for (int i = 0; i < NJ; ++i) a_out[i] = a[i];
  lam_out[0] = lamv[0]; lam_out[1] = lamv[1]; lam_out[2] = lamv[2];

  // ---------------- dynamics derivatives ----------------
  double dlam[3][NZ];  // d lambda / dz rows (only rows used by the costs are consumed)
  if (DIFF) {
    if (!in.with_control) {
#pragma unroll
      for (int i = 0; i < NJ * NZ; ++i) dadz[i] = 0.0;
    }
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int j = 0; j < NZ; ++j) dlam[r][j] = 0.0;
  }
  if (DIFF && in.with_control) {
    // accelerations / forces with qdd = a
    sv at[NJ];   // true spatial acceleration (no gravity)
    sv F[NJ];    // body force, later composite
    {
      sv acc = mksv(mk3(0, 0, 0), mk3(0, 0, 0));
#pragma unroll
      for (int i = 0; i < NJ; ++i) {
        acc = acc + J[i] * a[i];
        at[i] = ab[i] + acc;
        F[i] = fb[i] + si_mul(Y[i], acc);
      }
    }
    sv dVdq[NJ], dAdq[NJ];
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      dVdq[i] = mxm(v[i], J[i]);
      if (i == 0) dAdq[i] = mxm(ag, J[i]);
      else dAdq[i] = mxm(at[i - 1] + ag, J[i]) + mxm(v[i - 1], dVdq[i]);
    }
    double tq[NJ][NJ], tv[NJ][NJ];
    {
      si Yc = Y[NJ - 1];
      bop Bc;
      sv Fc = mksv(mk3(0, 0, 0), mk3(0, 0, 0));
#pragma unroll
      for (int m = NJ - 1; m >= 0; --m) {
        if (m < NJ - 1) si_add(Yc, Y[m]);
        // body B operator
        const sv h = si_mul(Y[m], v[m]);
        const v3 w = v[m].a, vl = v[m].l, hm = Y[m].h;
        const double *I = Y[m].I;
        const double I9[9] = {I[0], I[1], I[2], I[1], I[3], I[4], I[2], I[4], I[5]};
        const double wx[9] = {0.0, -w.z, w.y, w.z, 0.0, -w.x, -w.y, w.x, 0.0};
        const double hv[3] = {hm.x, hm.y, hm.z}, vv[3] = {vl.x, vl.y, vl.z};
        const double hax[9] = {0.0, -h.a.z, h.a.y, h.a.z, 0.0, -h.a.x, -h.a.y, h.a.x, 0.0};
        const double vh = dot(vl, hm);
        double Baa[9];
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
            Baa[3 * r + c] = s;
          }
        if (m == NJ - 1) {
          Bc.hl = h.l;
#pragma unroll
          for (int k = 0; k < 9; ++k) Bc.A[k] = Baa[k];
        } else {
          Bc.hl = Bc.hl + h.l;
#pragma unroll
          for (int k = 0; k < 9; ++k) Bc.A[k] += Baa[k];
        }
        Fc = Fc + F[m];
        const sv dFdq = si_mul(Yc, dAdq[m]) + bop_mul(Bc, dVdq[m]) + mxf(J[m], Fc);
        const sv dFdv = si_mul(Yc, dVdq[m] * 2.0) + bop_mul(Bc, J[m]);
        const sv BTJ = bop_mulT(Bc, J[m]);
#pragma unroll
        for (int i = 0; i <= m; ++i) {
          tq[i][m] = dot6(J[i], dFdq);
          tv[i][m] = dot6(J[i], dFdv);
        }
#pragma unroll
        for (int j = 0; j < m; ++j) {
          tq[m][j] = dot6(dFda[m], dAdq[j]) + dot6(BTJ, dVdq[j]);
          tv[m][j] = dot6(dFda[m], dVdq[j] * 2.0) + dot6(BTJ, J[j]);
        }
      }
    }
    if (NC == 0) {
#pragma unroll
      for (int i = 0; i < NJ; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          double sq = 0.0, sv_ = 0.0;
#pragma unroll
          for (int k = 0; k < NJ; ++k) {
            sq += Minv[i][k] * tq[k][j];
            sv_ += Minv[i][k] * tv[k][j];
          }
          dadz[i * NZ + j] = -sq;
          dadz[i * NZ + NJ + j] = -sv_;
          dadz[i * NZ + 2 * NJ + j] = Minv[i][j];
        }
    } else {
      // d(Jc^T lam)/dq : E[i][j] = lam_w . (z_min x Jf_max)   (kinematic Hessian, symmetric)
      const v3 lw = mk3(lamv[0], lamv[1], lamv[2]);
#pragma unroll
      for (int i = 0; i < NJ; ++i)
#pragma unroll
        for (int j = i; j < NJ; ++j) {
          const double e = dot(lw, cross(J[i].a, Jf[j]));
          tq[i][j] -= e;
          if (j != i) tq[j][i] -= e;
        }
      // da0/dx: derivative of the constraint acceleration c = a_O + al x p + w x vp + kp (p - pc) + kd vp, qdd fixed
      const double kp = cfg.contact_gains[0], kd = cfg.contact_gains[1];
      const sv a7 = at[NJ - 1];
      double dc[3][2 * NJ];
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        // wrt q_j
        const sv dv = mxm(J[j], v7 - v[j]);
        const sv da = mxm(J[j], a7 - at[j]) + mxm(dVdq[j], v7 - v[j]);
        const v3 dp = Jf[j];
        const v3 dvp = dv.l + cross(dv.a, pf) + cross(wf, dp);
        v3 d = da.l + cross(da.a, pf) + cross(a7.a, dp) + cross(dv.a, vf) + cross(wf, dvp) + dp * kp + dvp * kd;
        dc[0][j] = d.x; dc[1][j] = d.y; dc[2][j] = d.z;
        // wrt qd_j
        const sv dv2 = J[j];
        const sv da2 = mxm(J[j], v7 - v[j] * 2.0);
        const v3 dvp2 = dv2.l + cross(dv2.a, pf);
        d = da2.l + cross(da2.a, pf) + cross(dv2.a, vf) + cross(wf, dvp2) + dvp2 * kd;
        dc[0][NJ + j] = d.x; dc[1][NJ + j] = d.y; dc[2][NJ + j] = d.z;
      }
      // KKT inverse blocks: tr = X Sinv (7 x NC), tl = Minv - tr X^T
      double tr[NJ][3];
#pragma unroll
      for (int i = 0; i < NJ; ++i)
#pragma unroll
        for (int r = 0; r < NC; ++r) {
          double s = 0.0;
#pragma unroll
          for (int s2 = 0; s2 < NC; ++s2) s += X[s2][i] * Sinv[s2][r];
          tr[i][r] = s;
        }
      double tl[NJ][NJ];
#pragma unroll
      for (int i = 0; i < NJ; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          double s = Minv[i][j];
#pragma unroll
          for (int r = 0; r < NC; ++r) s -= tr[i][r] * X[r][j];
          tl[i][j] = s;
        }
#pragma unroll
      for (int i = 0; i < NJ; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          double sq = 0.0, sv_ = 0.0;
#pragma unroll
          for (int k = 0; k < NJ; ++k) {
            sq += tl[i][k] * tq[k][j];
            sv_ += tl[i][k] * tv[k][j];
          }
#pragma unroll
          for (int r = 0; r < NC; ++r) {
            const int ax = (NC == 1) ? 2 : r;
            sq += tr[i][r] * dc[ax][j];
            sv_ += tr[i][r] * dc[ax][NJ + j];
          }
          dadz[i * NZ + j] = -sq;
          dadz[i * NZ + NJ + j] = -sv_;
          dadz[i * NZ + 2 * NJ + j] = tl[i][j];
        }
      // d lam / dz = tr^T dtau_dx - Sinv da0_dx ; d lam / du = -tr^T
#pragma unroll
      for (int r = 0; r < NC; ++r) {
        const int ax = (NC == 1) ? 2 : r;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          double sq = 0.0, sv_ = 0.0;
#pragma unroll
          for (int k = 0; k < NJ; ++k) {
            sq += tr[k][r] * tq[k][j];
            sv_ += tr[k][r] * tv[k][j];
          }
#pragma unroll
          for (int s2 = 0; s2 < NC; ++s2) {
            const int ax2 = (NC == 1) ? 2 : s2;
            sq -= Sinv[r][s2] * dc[ax2][j];
            sv_ -= Sinv[r][s2] * dc[ax2][NJ + j];
          }
          dlam[ax][j] = sq;
          dlam[ax][NJ + j] = sv_;
          dlam[ax][2 * NJ + j] = -tr[j][r];
        }
      }
    }
  }

  // ---------------- cost stack ----------------
  const bool contact = (NC != 0);
  const bool state_reg = (cfg.variant == FFMPC_VARIANT_CLASSICAL) || cfg.use_inner_state_reg;
  const bool tau_reg = ((cfg.variant == FFMPC_VARIANT_CLASSICAL) || cfg.use_inner_tau_reg) && !in.terminal_costs;
  double cost = 0.0;
  double gz[NZ];   // gradient accumulator
  double hd[NZ];   // diagonal Hessian accumulator
  if (DIFF) {
#pragma unroll
    for (int i = 0; i < NZ; ++i) { gz[i] = 0.0; hd[i] = 0.0; }
  }
  // state-space terms -------------------------------------------------------
#pragma unroll
  for (int i = 0; i < NJ; ++i) {
    if (state_reg) {
      const double rq = q[i] - in.x_reg_ref[i], rv = qd[i] - in.x_reg_ref[NJ + i];
      cost += 0.5 * cfg.w_posture * (rq * rq + rv * rv);
      const double wv = cfg.w_v * cfg.v_damp_weights[i];
      cost += 0.5 * wv * qd[i] * qd[i];
      if (DIFF) {
        gz[i] += cfg.w_posture * rq;
        gz[NJ + i] += cfg.w_posture * rv + wv * qd[i];
        hd[i] += cfg.w_posture;
        hd[NJ + i] += cfg.w_posture + wv;
      }
    }
    if (cfg.w_q_soft_limits > 0.0) {
      double av, Ar, Arr;
      barrier_act(q[i] - cfg.q_soft_ref[i], cfg.q_soft_lb[i], cfg.q_soft_ub[i], av, Ar, Arr);
      cost += cfg.w_q_soft_limits * av;
      if (DIFF) {
        gz[i] += cfg.w_q_soft_limits * Ar;
        hd[i] += cfg.w_q_soft_limits * Arr;
      }
    }
    if (tau_reg) {
      const double ui = in.with_control ? in.u[i] : 0.0;
      const double r = in.with_control ? ui - in.tau_ref[i] : 0.0;
      cost += 0.5 * cfg.w_tau * r * r;
      if (DIFF && in.with_control) {
        gz[2 * NJ + i] += cfg.w_tau * r;
        hd[2 * NJ + i] += cfg.w_tau;
      }
      if (cfg.w_tau_soft_limits > 0.0) {
        double mn = cfg.tau_limits[0];
#pragma unroll
        for (int k = 1; k < NJ; ++k) mn = fmin(mn, cfg.tau_limits[k]);
        const double margin = fmin(fmax(cfg.tau_soft_limit_margin, 0.0), mn - 1.0e-6);
        double av, Ar, Arr;
        barrier_act(ui, -cfg.tau_limits[i] + margin, cfg.tau_limits[i] - margin, av, Ar, Arr);
        cost += cfg.w_tau_soft_limits * av;
        if (DIFF && in.with_control) {
          gz[2 * NJ + i] += cfg.w_tau_soft_limits * Ar;
          hd[2 * NJ + i] += cfg.w_tau_soft_limits * Arr;
        }
      }
    }
  }
  // frame terms ------------------------------------------------------------------
  // rows: 0-2 orientation log (q), 3-5 translation (q), 6-8 linear velocity (q,v), 9-11 angular velocity (q,v)
  double rw[12];    // activation second derivative * cost weight per row
  double rg[12];    // activation first derivative * cost weight per row
  v3 rori;
  {
    // r = log3(R_des^T Rf)
    m3 Rd;
#pragma unroll
    for (int i = 0; i < 9; ++i) Rd.m[i] = cfg.R_des[i];
    m3 Re;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) Re.m[3 * i + j] = Rd.m[i] * Rf.m[j] + Rd.m[3 + i] * Rf.m[3 + j] + Rd.m[6 + i] * Rf.m[6 + j];
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
  const v3 dpr = pf - mk3(in.p_ref[0], in.p_ref[1], in.p_ref[2]);
  {
    const double wo = cfg.w_ee_ori;
    const double r3[3] = {rori.x, rori.y, rori.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      rw[k] = wo * cfg.ori_weights[k];
      rg[k] = rw[k] * r3[k];
      cost += 0.5 * rw[k] * r3[k] * r3[k];
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
      cost += 0.5 * wt[k] * rt[k] * rt[k];
      rw[6 + k] = wvv[k];
      rg[6 + k] = wvv[k] * rv[k];
      cost += 0.5 * wvv[k] * rv[k] * rv[k];
      rw[9 + k] = cfg.w_wdamp * cfg.w_wdamp_weights[k];
      rg[9 + k] = rw[9 + k] * wa[k];
      cost += 0.5 * rw[9 + k] * wa[k] * wa[k];
    }
  }
  // contact-force terms ------------------------------------------------------------
  double fw = 0.0, fg = 0.0;  // second / first derivative weights on the normal-force row
  if (contact) {
    const double lz = in.with_control ? lamv[2] : 0.0;
    if (cfg.w_unilateral > 0.0) {
      double av, Ar, Arr;
      barrier_act(lz, cfg.friction_margin, INFINITY, av, Ar, Arr);
      cost += cfg.w_unilateral * av;
      fg += cfg.w_unilateral * Ar;
      fw += cfg.w_unilateral * Arr;
    }
    if (cfg.w_fn > 0.0) {
      const double r = in.with_control ? lz - cfg.fn_des : 0.0;
      cost += 0.5 * cfg.w_fn * r * r;
      fg += cfg.w_fn * r;
      fw += cfg.w_fn;
    }
    if (!in.with_control) { fg = 0.0; fw = 0.0; }
  }
  *cost_out = cost;

  if (DIFF) {
    // Jacobian rows of the frame residuals
    double Jr[12][2 * NJ];
    {
      // Jlog3(r) = 1 + 0.5 [r]x + c [r]x^2
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
      m3 JlRt;  // Jlog3 * Rf^T
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) JlRt.m[3 * r + c] = Jl[3 * r] * Rf.m[3 * c] + Jl[3 * r + 1] * Rf.m[3 * c + 1] + Jl[3 * r + 2] * Rf.m[3 * c + 2];
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const v3 o = mulv(JlRt, J[j].a);
        Jr[0][j] = o.x; Jr[1][j] = o.y; Jr[2][j] = o.z;
        Jr[0][NJ + j] = Jr[1][NJ + j] = Jr[2][NJ + j] = 0.0;
        Jr[3][j] = Jf[j].x; Jr[4][j] = Jf[j].y; Jr[5][j] = Jf[j].z;
        Jr[3][NJ + j] = Jr[4][NJ + j] = Jr[5][NJ + j] = 0.0;
        const sv dv = mxm(J[j], v7 - v[j]);
        const v3 dvp = dv.l + cross(dv.a, pf) + cross(wf, Jf[j]);
        Jr[6][j] = dvp.x; Jr[7][j] = dvp.y; Jr[8][j] = dvp.z;
        Jr[9][j] = dv.a.x; Jr[10][j] = dv.a.y; Jr[11][j] = dv.a.z;
        Jr[6][NJ + j] = Jf[j].x; Jr[7][NJ + j] = Jf[j].y; Jr[8][NJ + j] = Jf[j].z;
        Jr[9][NJ + j] = J[j].a.x; Jr[10][NJ + j] = J[j].a.y; Jr[11][NJ + j] = J[j].a.z;
      }
    }
    // gradient
#pragma unroll
    for (int c = 0; c < 2 * NJ; ++c) {
      double s = gz[c];
#pragma unroll
      for (int r = 0; r < 12; ++r) s += rg[r] * Jr[r][c];
      Lz[c] = sc * (s + fg * dlam[2][c]) + (eg ? eg[c] : 0.0);
    }
#pragma unroll
    for (int c = 2 * NJ; c < NZ; ++c) Lz[c] = sc * (gz[c] + fg * dlam[2][c]) + (eg ? eg[c] : 0.0);
    // Gauss-Newton Hessian, packed upper
#pragma unroll
    for (int i = 0; i < NZ; ++i)
#pragma unroll
      for (int j = i; j < NZ; ++j) {
        double s = (i == j) ? hd[i] : 0.0;
        if (j < 2 * NJ) {
          // rows 0-5 have zero v-columns
          const int r0 = (i < NJ && j < NJ) ? 0 : 6;
#pragma unroll
          for (int r = 0; r < 12; ++r)
            if (r >= r0) s += rw[r] * Jr[r][i] * Jr[r][j];
        }
        s += fw * dlam[2][i] * dlam[2][j];
        Lzz[sym_idx(i, j)] = sc * s + ((i == j && eh) ? eh[i] : 0.0);
      }
  }
}

// ---- _AugmentedLPFActionModel soft limit on |w| (crocoddyl_force_feedback.py:195-209) ----
FFMPC_DEV void lpf_soft_term(const ffmpc_config &cfg, int i, double w, double &c, double &g, double &h) {
  const double margin = fmax(cfg.tau_soft_limit_margin, 0.0);
  const double lim = fmax(cfg.tau_limits[i] - margin, 1.0e-9);
  const double over = fmax(fabs(w) - lim, 0.0);
  const bool active = over > 0.0;
  c = 0.5 * over * over;
  g = active ? over * (w > 0.0 ? 1.0 : (w < 0.0 ? -1.0 : 0.0)) : 0.0;
  h = active ? 1.0 : 0.0;
}

// Per-knot problem data shared by the action-model wrappers.
struct KnotCtx {
  const double *p_ref;      // [3]
  const double *v_ref;      // [3]
  const double *x_reg_ref;  // [14]
  const double *tau_ref;    // [7]
  const double *y_ref;      // [21] (FF)
  bool terminal;
};

// Node record written by the calcDiff kernel and consumed by the backward (Riccati) kernel.
//   [0,147)    G    = d a / d(q,v,u_inner)            (7 x 21, row-major, NOT scaled by dt)
//   [147,168)  Lz   = gradient over (x_mb, u_inner)   (already scaled: dt for running knots)
//   [168,399)  Lzz  = Gauss-Newton Hessian, packed upper 21x21 (scaled)
//   [399,406)  Lw   = FF only: gradient wrt the unfiltered control w
//   [406,413)  Lww  = FF only: diagonal of the Hessian wrt w
//   [413,416)  pad
constexpr int REC_G = 0;
constexpr int REC_LZ = 147;
constexpr int REC_LZZ = 168;
constexpr int REC_LW = 399;
constexpr int REC_LWW = 406;
constexpr int REC_SIZE = 416;

// ActionModel.calc (+ calcDiff when DIFF) of knot `ctx` at state x (nx = 14 classical / 21 FF) and control u.
//   xnext[nx], cost, lam[3];  rec[REC_SIZE] when DIFF.
template <int NC, bool FF, bool DIFF>
FFMPC_DEV void action_node(const ffmpc_model &mdl, const ffmpc_config &cfg, const KnotCtx &ctx, const double *x, const double *u,
                           double *xnext, double *cost_out, double *lam, double *rec) {
  const double dt = cfg.dt_ocp;
  NodeArgs in;
  in.x = x;
  in.p_ref = ctx.p_ref;
  in.v_ref = ctx.v_ref;
  in.x_reg_ref = ctx.x_reg_ref;
  in.tau_ref = ctx.tau_ref;
  in.terminal_costs = ctx.terminal;
  // FF: the inner control is the torque state and the wrapper always evaluates the RUNNING form of
  // the inner Euler model, terminal knot included (crocoddyl_force_feedback.py:211-220, quirk a16).
  in.with_control = FF ? true : !ctx.terminal;
  in.u = FF ? (x + 14) : u;
  const bool running = in.with_control;
  const double w_y = FF ? fmax(cfg.w_y, 0.0) : 0.0;
  double a[NJ], cost;
  double eg[NZ], eh[NZ];
  double sy = 0.0;
  if (FF) {
#pragma unroll
    for (int i = 0; i < NZ; ++i) {
      const double wy2 = cfg.y_reg_weights[i] * cfg.y_reg_weights[i];
      const double dy = x[i] - ctx.y_ref[i];
      sy += wy2 * dy * dy;
      eg[i] = w_y * wy2 * dy;
      eh[i] = w_y * wy2;
    }
  }
  dam_node<NC, DIFF>(mdl, cfg, in, a, lam, &cost, DIFF ? rec + REC_G : nullptr, DIFF ? rec + REC_LZ : nullptr,
                     DIFF ? rec + REC_LZZ : nullptr, running ? dt : 1.0, (FF && w_y > 0.0) ? eg : nullptr,
                     (FF && w_y > 0.0) ? eh : nullptr);
  // IntegratedActionModelEuler
  if (running) {
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      xnext[i] = x[i] + (dt * x[NJ + i] + dt * dt * a[i]);
      xnext[NJ + i] = x[NJ + i] + dt * a[i];
    }
    cost *= dt;
  } else {
#pragma unroll
    for (int i = 0; i < 2 * NJ; ++i) xnext[i] = x[i];
  }
  if (FF) {
    const double alpha = fmin(fmax(cfg.ff_alpha, 0.0), 0.999999), beta = 1.0 - alpha;
    const double w_w = fmax(cfg.w_w, 0.0), w_soft = fmax(cfg.w_w_soft_limits, 0.0);
    double sw = 0.0, ss = 0.0;
#pragma unroll
    for (int i = 0; i < NJ; ++i) {
      const double w = ctx.terminal ? 0.0 : u[i];
      xnext[2 * NJ + i] = alpha * x[2 * NJ + i] + beta * w;
      sw += w * w;
      double c = 0.0, g = 0.0, h = 0.0;
      if (w_soft > 0.0) lpf_soft_term(cfg, i, w, c, g, h);
      ss += c;
      if (DIFF) {
        rec[REC_LW + i] = w_w * w + w_soft * g;
        rec[REC_LWW + i] = w_w + w_soft * h;
      }
    }
    if (w_y > 0.0) cost += 0.5 * w_y * sy;
    if (w_w > 0.0) cost += 0.5 * w_w * sw;
    if (w_soft > 0.0) cost += w_soft * ss;
  }
  *cost_out = cost;
}

}  // namespace ffmpc
