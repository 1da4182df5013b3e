// ffmpc_oracle.cpp -- CPU ORACLE (test infrastructure, NOT the product).
//
// A from-scratch CPU restatement, in plain C++17 / fp64, of the hot path of
// omar-a-aman/Franka-Force-Feedback-MPC: the per-tick OCP of
//   src/mpc/crocoddyl_classical.py:521-728, 944-997        (ClassicalCrocoddylMPC)
//   src/mpc/crocoddyl_force_feedback.py:149-290, 776-1009  (ForceFeedbackCrocoddylMPC)
// solved by Crocoddyl's SolverBoxFDDP / SolverFDDP (call site crocoddyl_classical.py:367).
//
// PARITY STATUS: "parity unpinned" against real Crocoddyl/Pinocchio.  Those libraries are
// third-party, un-pinned conda-forge dependencies (environment.yml:15-17), absent from
// /root/reference, not installed in this image and not installable (no network); the reference
// holds no tests, golden vectors or fixtures.  The oracle is therefore pinned by
//   (1) the pieces of the path that ARE reference source: _AugmentedLPFActionModel.calc/calcDiff
//       (crocoddyl_force_feedback.py:211-290) run here under stub modules -> tests/golden/,
//   (2) first-principles checks: central finite differences of every derivative, rigid-body
//       identities (RNEA(ABA)=id, M SPD, KKT residual), the geometric anchors of SURVEY App. C,
//   (3) the published algorithms (Mastalli 2020/2022 FDDP/BoxFDDP, Tassa 2014 BoxQP) as written
//       down in SURVEY.md App. A.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  The product (libffmpc_b200.so) never links or calls it.
//
// Design choice that keeps this file independent of the CUDA kernels: every dynamics / residual
// Jacobian here is obtained by forward-mode automatic differentiation (dual numbers with 21
// partials, z = (q, v, u)) of the *forward* functions, i.e. by construction it is the exact
// derivative of what calc() computes -- the property Crocoddyl's analytic derivatives have
// (SURVEY App. A.3/A.4).  The CUDA kernels use hand-derived analytic recursions instead.
// Cost Hessians are Gauss-Newton (Rx^T Arr Rx), exactly like CostModelResidual.

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <thread>
#include <atomic>
#include <type_traits>
#include <vector>

#include "../include/ffmpc.h"

namespace {

constexpr int NJ = 7;
constexpr int NZ = 21;  // (q, v, u) stacking of the inner DAM

// ----------------------------------------------------------------------------------------------
// forward-mode scalar
// ----------------------------------------------------------------------------------------------
// work counters (how many rollouts the line search really runs): feeds the FLOP accounting of bench.py / DESIGN.md
std::atomic<long long> g_rollouts{0}, g_iterations{0};
std::atomic<long long> g_trial_hist[32];

// The scalar of every computation below.  double: the oracle proper (bit-for-bit the round-1 oracle).  __float128
// (-DFFMPC_ORACLE_QUAD, libffmpc_oracle_quad.so): the same source as a higher-precision ARBITER -- when the fp64 GPU
// result and the fp64 oracle disagree at the 1e-8 level after ten nonlinear iterations, the quad build says which of
// the two moved (tests/test_gpu_parity.py).  Inputs, outputs, model and config stay fp64 at the C boundary.
#ifdef FFMPC_ORACLE_QUAD
#include <quadmath.h>
typedef __float128 real;
#define RPI (4 * atanq(1))
inline real rsqrt_(real x) { return sqrtq(x); }
inline real rsin(real x) { return sinq(x); }
inline real rcos(real x) { return cosq(x); }
inline real racos(real x) { return acosq(x); }
inline real rabs(real x) { return fabsq(x); }
inline bool rfinite(real x) { return finiteq(x) != 0; }
inline real rldexp(real x, int e) { return ldexpq(x, e); }
#elif defined(FFMPC_ORACLE_COUNT)
// Counting scalar (libffmpc_oracle_count.so): fp64 arithmetic that counts its own operations -- add / sub / mul / div /
// sqrt / sin / cos / acos = 1 each, so a multiply-add is 2 (SURVEY.md 8d "superseding figure").  What it counts is THIS
// restatement: forward-mode duals with 21 tangent lanes for every derivative, dense products in the Riccati sweep.
struct real {
  double v;
  real() : v(0.0) {}
  real(double x) : v(x) {}  // NOLINT implicit on purpose
  explicit operator double() const { return v; }
};
static thread_local long long g_flops_tl = 0;
std::atomic<long long> g_flops{0};
#define CNT_BIN(op)                                                                                   \
  inline real operator op(const real &a, const real &b) { ++g_flops_tl; return real(a.v op b.v); }     \
  inline real operator op(const real &a, double b) { ++g_flops_tl; return real(a.v op b); }            \
  inline real operator op(double a, const real &b) { ++g_flops_tl; return real(a op b.v); }
CNT_BIN(+) CNT_BIN(-) CNT_BIN(*) CNT_BIN(/)
#define CNT_CMP(op)                                                                  \
  inline bool operator op(const real &a, const real &b) { return a.v op b.v; }       \
  inline bool operator op(const real &a, double b) { return a.v op b; }              \
  inline bool operator op(double a, const real &b) { return a op b.v; }
CNT_CMP(<) CNT_CMP(>) CNT_CMP(<=) CNT_CMP(>=) CNT_CMP(==) CNT_CMP(!=)
inline real operator-(const real &a) { return real(-a.v); }
inline real &operator+=(real &a, const real &b) { ++g_flops_tl; a.v += b.v; return a; }
inline real &operator-=(real &a, const real &b) { ++g_flops_tl; a.v -= b.v; return a; }
inline real &operator*=(real &a, const real &b) { ++g_flops_tl; a.v *= b.v; return a; }
inline real &operator/=(real &a, const real &b) { ++g_flops_tl; a.v /= b.v; return a; }
#define RPI 3.14159265358979323846
inline real rsqrt_(real x) { ++g_flops_tl; return std::sqrt(x.v); }
inline real rsin(real x) { ++g_flops_tl; return std::sin(x.v); }
inline real rcos(real x) { ++g_flops_tl; return std::cos(x.v); }
inline real racos(real x) { ++g_flops_tl; return std::acos(x.v); }
inline real rabs(real x) { return std::fabs(x.v); }
inline bool rfinite(real x) { return std::isfinite(x.v); }
inline real rldexp(real x, int e) { return std::ldexp(x.v, e); }
#else
typedef double real;
#define RPI 3.14159265358979323846
inline real rsqrt_(real x) { return std::sqrt(x); }
inline real rsin(real x) { return std::sin(x); }
inline real rcos(real x) { return std::cos(x); }
inline real racos(real x) { return std::acos(x); }
inline real rabs(real x) { return std::fabs(x); }
inline bool rfinite(real x) { return std::isfinite(x); }
inline real rldexp(real x, int e) { return std::ldexp(x, e); }
#endif
inline real rmin(real a, real b) { return (b < a) ? b : a; }  // std::min / std::max semantics
inline real rmax(real a, real b) { return (a < b) ? b : a; }
inline real rinf() { return std::numeric_limits<double>::infinity(); }
inline real rnan() { return std::numeric_limits<double>::quiet_NaN(); }
#ifdef FFMPC_ORACLE_COUNT
inline double to_d(const real &x) { return x.v; }
#else
inline double to_d(real x) { return (double)x; }
#endif
template <class It> void copy_out(It first, It last, double *dst) { for (; first != last; ++first, ++dst) *dst = to_d(*first); }

struct Dual {
  real v;
  real d[NZ];
  Dual() : v(0.0) { std::memset(d, 0, sizeof(d)); }
  Dual(real x) : v(x) { std::memset(d, 0, sizeof(d)); }  // NOLINT implicit on purpose
};
inline Dual operator+(const Dual &a, const Dual &b) { Dual r; r.v = a.v + b.v; for (int i = 0; i < NZ; ++i) r.d[i] = a.d[i] + b.d[i]; return r; }
inline Dual operator-(const Dual &a, const Dual &b) { Dual r; r.v = a.v - b.v; for (int i = 0; i < NZ; ++i) r.d[i] = a.d[i] - b.d[i]; return r; }
inline Dual operator-(const Dual &a) { Dual r; r.v = -a.v; for (int i = 0; i < NZ; ++i) r.d[i] = -a.d[i]; return r; }
inline Dual operator*(const Dual &a, const Dual &b) { Dual r; r.v = a.v * b.v; for (int i = 0; i < NZ; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i]; return r; }
inline Dual operator/(const Dual &a, const Dual &b) { Dual r; const real ib = 1.0 / b.v; r.v = a.v * ib; for (int i = 0; i < NZ; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * ib; return r; }
inline Dual &operator+=(Dual &a, const Dual &b) { a = a + b; return a; }
inline Dual &operator-=(Dual &a, const Dual &b) { a = a - b; return a; }
inline Dual sqrt(const Dual &a) { Dual r; r.v = rsqrt_(a.v); const real s = 0.5 / r.v; for (int i = 0; i < NZ; ++i) r.d[i] = a.d[i] * s; return r; }
inline Dual sin(const Dual &a) { Dual r; r.v = rsin(a.v); const real c = rcos(a.v); for (int i = 0; i < NZ; ++i) r.d[i] = a.d[i] * c; return r; }
inline Dual cos(const Dual &a) { Dual r; r.v = rcos(a.v); const real s = -rsin(a.v); for (int i = 0; i < NZ; ++i) r.d[i] = a.d[i] * s; return r; }
inline Dual acos(const Dual &a) { Dual r; r.v = racos(a.v); const real s = -1.0 / rsqrt_(1.0 - a.v * a.v); for (int i = 0; i < NZ; ++i) r.d[i] = a.d[i] * s; return r; }
inline real val(real x) { return x; }
inline real val(const Dual &x) { return x.v; }
inline real sqrt(real x) { return rsqrt_(x); }
inline real sin(real x) { return rsin(x); }
inline real cos(real x) { return rcos(x); }
inline real acos(real x) { return racos(x); }

// ----------------------------------------------------------------------------------------------
// tiny 3-vector / 3x3 algebra, generic in the scalar
// ----------------------------------------------------------------------------------------------
template <class T> struct V3 { T x, y, z; };
template <class T> struct M3 { T m[3][3]; };
template <class T> struct S6 { V3<T> l, a; };  // spatial vector, Pinocchio order [linear; angular]

template <class T> V3<T> mk(const T &x, const T &y, const T &z) { return V3<T>{x, y, z}; }
template <class T> V3<T> operator+(const V3<T> &a, const V3<T> &b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <class T> V3<T> operator-(const V3<T> &a, const V3<T> &b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <class T> V3<T> operator*(const V3<T> &a, const T &s) { return {a.x * s, a.y * s, a.z * s}; }
template <class T> V3<T> cross(const V3<T> &a, const V3<T> &b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
template <class T> T dot(const V3<T> &a, const V3<T> &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <class T> V3<T> mul(const M3<T> &A, const V3<T> &b) {
  return {A.m[0][0] * b.x + A.m[0][1] * b.y + A.m[0][2] * b.z, A.m[1][0] * b.x + A.m[1][1] * b.y + A.m[1][2] * b.z,
          A.m[2][0] * b.x + A.m[2][1] * b.y + A.m[2][2] * b.z};
}
template <class T> M3<T> mul(const M3<T> &A, const M3<T> &B) {
  M3<T> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.m[i][j] = A.m[i][0] * B.m[0][j] + A.m[i][1] * B.m[1][j] + A.m[i][2] * B.m[2][j];
  return C;
}
template <class T> M3<T> transpose(const M3<T> &A) {
  M3<T> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.m[i][j] = A.m[j][i];
  return C;
}
template <class T> V3<T> col(const M3<T> &A, int j) { return {A.m[0][j], A.m[1][j], A.m[2][j]}; }
template <class T, class S> M3<T> m3_from(const S *r) {
  M3<T> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.m[i][j] = T(r[3 * i + j]);
  return C;
}
template <class T, class S> V3<T> v3_from(const S *r) { return {T(r[0]), T(r[1]), T(r[2])}; }
template <class T> S6<T> operator+(const S6<T> &a, const S6<T> &b) { return {a.l + b.l, a.a + b.a}; }
template <class T> S6<T> operator*(const S6<T> &a, const T &s) { return {a.l * s, a.a * s}; }
// motion x motion
template <class T> S6<T> mxm(const S6<T> &a, const S6<T> &b) { return {cross(a.a, b.l) + cross(a.l, b.a), cross(a.a, b.a)}; }
// motion x* force
template <class T> S6<T> mxf(const S6<T> &m, const S6<T> &f) { return {cross(m.a, f.l), cross(m.a, f.a) + cross(m.l, f.l)}; }
template <class T> T dot6(const S6<T> &a, const S6<T> &b) { return dot(a.l, b.l) + dot(a.a, b.a); }

// spatial inertia about the world origin: mass, first moment h = m c, rotational inertia about the origin
template <class T> struct SI {
  T m;
  V3<T> h;
  M3<T> I;
};
template <class T> S6<T> mul(const SI<T> &Y, const S6<T> &v) { return {v.l * Y.m + cross(v.a, Y.h), mul(Y.I, v.a) + cross(Y.h, v.l)}; }
template <class T> SI<T> add(const SI<T> &A, const SI<T> &B) {
  SI<T> C;
  C.m = A.m + B.m;
  C.h = A.h + B.h;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.I.m[i][j] = A.I.m[i][j] + B.I.m[i][j];
  return C;
}

// ----------------------------------------------------------------------------------------------
// rigid-body kinematics / dynamics of the 7-DoF revolute-z chain
// ----------------------------------------------------------------------------------------------
template <class T> struct Kin {
  M3<T> R[NJ];     // world rotation of joint frame i
  V3<T> p[NJ];     // world position
  S6<T> J[NJ];     // world spatial joint axis [p x z; z]
  S6<T> v[NJ];     // world spatial velocity
  S6<T> ab[NJ];    // world spatial bias acceleration (qdd = 0, no gravity)
  SI<T> Y[NJ];     // world spatial inertia of body i
  V3<T> pf;        // EE frame origin
  M3<T> Rf;        // EE frame rotation
};

template <class T> void kinematics(const ffmpc_model &mdl, const T *q, const T *qd, Kin<T> &k) {
  for (int i = 0; i < NJ; ++i) {
    const T c = cos(q[i]), s = sin(q[i]);
    M3<T> Rz;
    Rz.m[0][0] = c; Rz.m[0][1] = -s; Rz.m[0][2] = T(0.0);
    Rz.m[1][0] = s; Rz.m[1][1] = c; Rz.m[1][2] = T(0.0);
    Rz.m[2][0] = T(0.0); Rz.m[2][1] = T(0.0); Rz.m[2][2] = T(1.0);
    const M3<T> Rl = mul(m3_from<T>(mdl.joint_rot[i]), Rz);
    const V3<T> pl = v3_from<T>(mdl.joint_pos[i]);
    if (i == 0) {
      k.R[i] = Rl;
      k.p[i] = pl;
    } else {
      k.R[i] = mul(k.R[i - 1], Rl);
      k.p[i] = k.p[i - 1] + mul(k.R[i - 1], pl);
    }
    const V3<T> z = col(k.R[i], 2);
    k.J[i] = {cross(k.p[i], z), z};
    const S6<T> vj = k.J[i] * qd[i];
    if (i == 0) {
      k.v[i] = vj;
      k.ab[i] = {mk(T(0.0), T(0.0), T(0.0)), mk(T(0.0), T(0.0), T(0.0))};
    } else {
      k.v[i] = k.v[i - 1] + vj;
      k.ab[i] = k.ab[i - 1] + mxm(k.v[i - 1], vj);
    }
    // world inertia
    const V3<T> cw = k.p[i] + mul(k.R[i], v3_from<T>(mdl.com[i]));
    M3<T> Ib;
    const double *in = mdl.inertia[i];
    Ib.m[0][0] = T(in[0]); Ib.m[1][1] = T(in[1]); Ib.m[2][2] = T(in[2]);
    Ib.m[0][1] = Ib.m[1][0] = T(in[3]);
    Ib.m[0][2] = Ib.m[2][0] = T(in[4]);
    Ib.m[1][2] = Ib.m[2][1] = T(in[5]);
    M3<T> Iw = mul(mul(k.R[i], Ib), transpose(k.R[i]));
    const T m = T(mdl.mass[i]);
    // I_O = I_c - m [c]x[c]x = I_c + m (|c|^2 I - c c^T)
    const T c2 = dot(cw, cw);
    const T cv[3] = {cw.x, cw.y, cw.z};
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) Iw.m[a][b] = Iw.m[a][b] + m * ((a == b ? c2 : T(0.0)) - cv[a] * cv[b]);
    k.Y[i].m = m;
    k.Y[i].h = cw * m;
    k.Y[i].I = Iw;
  }
  k.pf = k.p[NJ - 1] + mul(k.R[NJ - 1], v3_from<T>(mdl.ee_pos));
  k.Rf = mul(k.R[NJ - 1], m3_from<T>(mdl.ee_rot));
}

// tau = RNEA(q, qd, qdd) with gravity, optional world force `fw` applied on body 7 at the EE origin
template <class T> void rnea(const ffmpc_model &mdl, const Kin<T> &k, const T *qdd, const V3<T> *fw, T *tau) {
  S6<T> a_prev = {mk(T(-mdl.gravity[0]), T(-mdl.gravity[1]), T(-mdl.gravity[2])), mk(T(0.0), T(0.0), T(0.0))};
  S6<T> f[NJ];
  S6<T> acc = {mk(T(0.0), T(0.0), T(0.0)), mk(T(0.0), T(0.0), T(0.0))};  // sum_k J_k qdd_k
  for (int i = 0; i < NJ; ++i) {
    acc = acc + k.J[i] * qdd[i];
    const S6<T> a = a_prev + k.ab[i] + acc;
    f[i] = mul(k.Y[i], a) + mxf(k.v[i], mul(k.Y[i], k.v[i]));
  }
  if (fw) {
    const S6<T> phi = {*fw, cross(k.pf, *fw)};
    f[NJ - 1] = f[NJ - 1] + phi * T(-1.0);
  }
  for (int i = NJ - 1; i >= 0; --i) {
    tau[i] = dot6(k.J[i], f[i]);
    if (i > 0) f[i - 1] = f[i - 1] + f[i];
  }
}

template <class T> void crba(const Kin<T> &k, T M[NJ][NJ]) {
  SI<T> Yc[NJ];
  Yc[NJ - 1] = k.Y[NJ - 1];
  for (int i = NJ - 2; i >= 0; --i) Yc[i] = add(k.Y[i], Yc[i + 1]);
  for (int j = 0; j < NJ; ++j) {
    const S6<T> F = mul(Yc[j], k.J[j]);
    for (int i = 0; i <= j; ++i) {
      M[i][j] = dot6(k.J[i], F);
      M[j][i] = M[i][j];
    }
  }
}

// dense Cholesky A = L L^T (lower), returns false on a non-positive pivot (Eigen::LLT NumericalIssue)
template <class T, int N> bool chol(T A[N][N], int n) {
  for (int j = 0; j < n; ++j) {
    T s = A[j][j];
    for (int k = 0; k < j; ++k) s = s - A[j][k] * A[j][k];
    if (!(val(s) > 0.0)) return false;
    const T l = sqrt(s);
    A[j][j] = l;
    for (int i = j + 1; i < n; ++i) {
      T t = A[i][j];
      for (int k = 0; k < j; ++k) t = t - A[i][k] * A[j][k];
      A[i][j] = t / l;
    }
  }
  return true;
}
template <class T, int N> void chol_solve(const T L[N][N], int n, T *b) {
  for (int i = 0; i < n; ++i) {
    T s = b[i];
    for (int k = 0; k < i; ++k) s = s - L[i][k] * b[k];
    b[i] = s / L[i][i];
  }
  for (int i = n - 1; i >= 0; --i) {
    T s = b[i];
    for (int k = i + 1; k < n; ++k) s = s - L[k][i] * b[k];
    b[i] = s / L[i][i];
  }
}

// log3 of a rotation matrix (Pinocchio convention, series guard near 0)
template <class T> V3<T> log3(const M3<T> &R) {
  const T tr = R.m[0][0] + R.m[1][1] + R.m[2][2];
  T t;
  const real trv = val(tr);
  if (trv >= 3.0 - 1e-14) {
    t = T(0.5);
  } else {
    T ca = (tr - T(1.0)) * T(0.5);
    if (val(ca) < -1.0) ca = T(-1.0);
    const T th = acos(ca);
    if (val(th) > 1e-7) t = th / (sin(th) * T(2.0)); else t = T(0.5);
  }
  return {(R.m[2][1] - R.m[1][2]) * t, (R.m[0][2] - R.m[2][0]) * t, (R.m[1][0] - R.m[0][1]) * t};
}

// ----------------------------------------------------------------------------------------------
// cost accumulation: CostModelSum of CostModelResidual terms
// ----------------------------------------------------------------------------------------------
struct CostAcc {
  real cost = 0.0;
  real Lz[NZ];
  real Lzz[NZ][NZ];
  CostAcc() {
    std::memset(Lz, 0, sizeof(Lz));
    std::memset(Lzz, 0, sizeof(Lzz));
  }
};
inline void acc_jac(CostAcc &, real, const real &, real, real) {}
inline void acc_jac(CostAcc &c, real w, const Dual &r, real Ar, real Arr) {
  if (Ar != 0.0)
    for (int a = 0; a < NZ; ++a) c.Lz[a] += w * Ar * r.d[a];
  if (Arr != 0.0)
    for (int a = 0; a < NZ; ++a) {
      const real wa = w * Arr * r.d[a];
      if (wa != 0.0)
        for (int b = 0; b < NZ; ++b) c.Lzz[a][b] += wa * r.d[b];
    }
}
// ActivationModelWeightedQuad (w == nullptr -> plain quad)
template <class T, class W> void add_quad(CostAcc &c, real weight, int nr, const T *r, const W *aw) {
  real a = 0.0;
  for (int i = 0; i < nr; ++i) {
    const real wi = aw ? aw[i] : 1.0;
    const real ri = val(r[i]);
    a += wi * ri * ri;
    acc_jac(c, weight, r[i], wi * ri, wi);
  }
  c.cost += weight * 0.5 * a;
}
// ActivationModelQuadraticBarrier(ActivationBounds(lb, ub, 1.0))
template <class T> void add_barrier(CostAcc &c, real weight, int nr, const T *r, const real *lb, const real *ub) {
  real a = 0.0;
  for (int i = 0; i < nr; ++i) {
    const real ri = val(r[i]);
    const real lo = rmin(ri - lb[i], 0.0);
    const real hi = rmax(ri - ub[i], 0.0);
    a += lo * lo + hi * hi;
    const real Arr = ((ri - lb[i]) <= 0.0 ? 1.0 : 0.0) + ((ri - ub[i]) >= 0.0 ? 1.0 : 0.0);
    acc_jac(c, weight, r[i], lo + hi, Arr);
  }
  c.cost += weight * 0.5 * a;
}

struct NodeRefs {
  const real *p_ref;      // [3]
  const real *v_ref;      // [3]
  const real *x_reg_ref;  // [14]
  const real *tau_ref;    // [7]
};

template <class T> struct DamOut {
  T a[NJ];
  T lam[3];
  CostAcc c;
  bool ok = true;
};

// One differential action model evaluation: Free / Contact forward dynamics + the full cost stack
// (crocoddyl_classical.py:558-728; FF :838-1009).  z = (q, v, u).
//   terminal_costs : drop tau_reg / tau_soft_limits (terminal=True in _make_dam)
//   with_control   : false reproduces DAM.calc(data, x) of a *classical* terminal node: no dynamics,
//                    control and contact-force residuals are zero with zero Jacobians (App. A.3)
template <class T>
void dam_eval(const ffmpc_model &mdl, const ffmpc_config &cfg, int mode, bool terminal_costs, bool with_control,
              const NodeRefs &ref, const T *q, const T *qd, const T *u, DamOut<T> &out) {
  Kin<T> k;
  kinematics(mdl, q, qd, k);
  const int nc = (mode == FFMPC_MODE_CONTACT) ? cfg.contact_nc : 0;

  // frame quantities, LOCAL_WORLD_ALIGNED
  const S6<T> &v7 = k.v[NJ - 1];
  const V3<T> vf = v7.l + cross(v7.a, k.pf);  // velocity of the frame origin
  const V3<T> wf = v7.a;
  const real z_target = ref.p_ref[2] - cfg.z_press;
  const V3<T> p_contact = mk(T(ref.p_ref[0]), T(ref.p_ref[1]), T(z_target));

  for (int i = 0; i < 3; ++i) out.lam[i] = T(0.0);
  for (int i = 0; i < NJ; ++i) out.a[i] = T(0.0);

  if (with_control) {
    T M[NJ][NJ], L[NJ][NJ];
    crba(k, M);
    T zero[NJ];
    for (int i = 0; i < NJ; ++i) zero[i] = T(0.0);
    T nle[NJ];
    rnea<T>(mdl, k, zero, nullptr, nle);
    for (int i = 0; i < NJ; ++i)
      for (int j = 0; j < NJ; ++j) L[i][j] = M[i][j];
    if (!chol<T, NJ>(L, NJ)) out.ok = false;
    T rhs0[NJ];
    for (int i = 0; i < NJ; ++i) rhs0[i] = u[i] - nle[i];
    chol_solve<T, NJ>(L, NJ, rhs0);
    if (nc == 0) {
      // DifferentialActionModelFreeFwdDynamics: a = ABA(q, v, tau) = M^-1 (tau - nle)
      for (int i = 0; i < NJ; ++i) out.a[i] = rhs0[i];
    } else {
      // ContactModel1D / ContactModel3D, LOCAL_WORLD_ALIGNED, R = I, Baumgarte gains (App. A.3)
      const S6<T> &ab7 = k.ab[NJ - 1];
      const V3<T> a_cl = ab7.l + cross(ab7.a, k.pf) + cross(wf, vf);  // classical drift acceleration
      const T kp = T(cfg.contact_gains[0]), kd = T(cfg.contact_gains[1]);
      const V3<T> a0v = a_cl + (k.pf - p_contact) * kp + vf * kd;
      T Jc[3][NJ], a0[3];
      for (int j = 0; j < NJ; ++j) {
        const V3<T> jl = k.J[j].l + cross(k.J[j].a, k.pf);
        if (nc == 1) {
          Jc[0][j] = jl.z;
        } else {
          Jc[0][j] = jl.x; Jc[1][j] = jl.y; Jc[2][j] = jl.z;
        }
      }
      if (nc == 1) a0[0] = a0v.z; else { a0[0] = a0v.x; a0[1] = a0v.y; a0[2] = a0v.z; }
      // X = M^-1 Jc^T ; S = Jc X + mu I ; lambda = -S^-1 (Jc M^-1 (tau - nle) + a0) ; a = M^-1(tau-nle) + X lambda
      T X[3][NJ];
      for (int r = 0; r < nc; ++r) {
        for (int i = 0; i < NJ; ++i) X[r][i] = Jc[r][i];
        chol_solve<T, NJ>(L, NJ, X[r]);
      }
      T S[3][3];
      T b[3];
      for (int r = 0; r < nc; ++r) {
        for (int s = 0; s < nc; ++s) {
          T acc = T(0.0);
          for (int i = 0; i < NJ; ++i) acc = acc + Jc[r][i] * X[s][i];
          S[r][s] = acc;
        }
        S[r][r] = S[r][r] + T(cfg.contact_inv_damping);
        T acc = a0[r];
        for (int i = 0; i < NJ; ++i) acc = acc + Jc[r][i] * rhs0[i];
        b[r] = -acc;
      }
      if (!chol<T, 3>(S, nc)) out.ok = false;
      chol_solve<T, 3>(S, nc, b);
      for (int i = 0; i < NJ; ++i) {
        T acc = rhs0[i];
        for (int r = 0; r < nc; ++r) acc = acc + X[r][i] * b[r];
        out.a[i] = acc;
      }
      if (nc == 1) out.lam[2] = b[0]; else { out.lam[0] = b[0]; out.lam[1] = b[1]; out.lam[2] = b[2]; }
    }
  }

  // ---- cost stack; CostModelSum iterates a std::map, i.e. alphabetical order of the names ----
  CostAcc &c = out.c;
  const bool contact = (nc > 0);
  const bool state_reg = (cfg.variant == FFMPC_VARIANT_CLASSICAL) || cfg.use_inner_state_reg;
  const bool tau_reg = (cfg.variant == FFMPC_VARIANT_CLASSICAL) || cfg.use_inner_tau_reg;
  // "ee_ori"
  {
    const M3<T> Rerr = mul(transpose(m3_from<T>(cfg.R_des)), k.Rf);
    const V3<T> r3 = log3(Rerr);
    const T r[3] = {r3.x, r3.y, r3.z};
    add_quad<T>(c, cfg.w_ee_ori, 3, r, cfg.ori_weights);
  }
  const V3<T> dp_ref = k.pf - v3_from<T>(ref.p_ref);
  if (!contact) {  // "ee_pos"
    const T r[3] = {dp_ref.x, dp_ref.y, dp_ref.z};
    const real aw[3] = {1.0, 1.0, 2.5};
    add_quad<T>(c, cfg.w_ee_pos, 3, r, aw);
  } else {
    {  // "ee_vxy"
      const T r[6] = {vf.x - T(ref.v_ref[0]), vf.y - T(ref.v_ref[1]), vf.z, wf.x, wf.y, wf.z};
      const real aw[6] = {1.0, 1.0, 0.0, 0.0, 0.0, 0.0};
      add_quad<T>(c, cfg.w_tangent_vel, 6, r, aw);
    }
    {  // "ee_xy"
      const T r[3] = {dp_ref.x, dp_ref.y, dp_ref.z};
      const real aw[3] = {1.0, 1.0, 0.0};
      add_quad<T>(c, cfg.w_tangent_pos, 3, r, aw);
    }
    if (cfg.w_fn > 0.0) {  // "fn_track"
      if (nc == 1) {
        const T r[1] = {with_control ? out.lam[2] - T(cfg.fn_des) : T(0.0)};
        const real aw[1] = {1.0};
        add_quad<T>(c, cfg.w_fn, 1, r, aw);
      } else {
        T r[3] = {out.lam[0], out.lam[1], out.lam[2] - T(cfg.fn_des)};
        if (!with_control) r[0] = r[1] = r[2] = T(0.0);
        const real aw[3] = {0.0, 0.0, 1.0};
        add_quad<T>(c, cfg.w_fn, 3, r, aw);
      }
    }
    if (nc == 3 && cfg.w_friction_cone > 0.0) {  // "friction_cone" (crocoddyl_classical.py:678-686)
      // FrictionCone(R = I, mu, nf = 4, inner_appr = false) [upstream]: facet rows (-mu n + t_i)^T, t_i = (cos, sin, 0) at
      // i 2pi/nf, bounds (-inf, 0]; last row n^T, bounds [min_nforce = 0, inf).  Margin: :891-904.
      const real inf = rinf();
      const real eps = rmax(cfg.friction_margin, 0.0);
      const real pi = RPI;
      T r[5];
      real lb[5], ub[5];
      for (int i = 0; i < 4; ++i) {
        const real th = 2.0 * pi / 4.0 * i;
        r[i] = with_control ? T(rcos(th)) * out.lam[0] + T(rsin(th)) * out.lam[1] - T(cfg.mu) * out.lam[2] : T(0.0);
        lb[i] = -inf;
        ub[i] = 0.0 - eps;
      }
      r[4] = with_control ? out.lam[2] : T(0.0);
      lb[4] = 0.0 + eps;
      ub[4] = inf;
      add_barrier<T>(c, cfg.w_friction_cone, 5, r, lb, ub);
    }
    if (cfg.w_plane_z > 0.0) {  // "plane_z"
      const V3<T> d = k.pf - p_contact;
      const T r[3] = {d.x, d.y, d.z};
      const real aw[3] = {0.0, 0.0, 1.0};
      add_quad<T>(c, cfg.w_plane_z, 3, r, aw);
    }
  }
  if (state_reg) {  // "posture"
    T r[14];
    for (int i = 0; i < NJ; ++i) {
      r[i] = q[i] - T(ref.x_reg_ref[i]);
      r[NJ + i] = qd[i] - T(ref.x_reg_ref[NJ + i]);
    }
    add_quad<T>(c, cfg.w_posture, 14, r, (const real *)nullptr);
  }
  if (cfg.w_q_soft_limits > 0.0) {  // "q_soft_limits"
    T r[14];
    real lb[14], ub[14];
    for (int i = 0; i < NJ; ++i) {
      r[i] = q[i] - T(cfg.q_soft_ref[i]);
      r[NJ + i] = qd[i];
      lb[i] = cfg.q_soft_lb[i];
      ub[i] = cfg.q_soft_ub[i];
      lb[NJ + i] = -rinf();
      ub[NJ + i] = rinf();
    }
    add_barrier<T>(c, cfg.w_q_soft_limits, 14, r, lb, ub);
  }
  if (!terminal_costs && tau_reg) {
    {  // "tau_reg"
      T r[NJ];
      for (int i = 0; i < NJ; ++i) r[i] = with_control ? u[i] - T(ref.tau_ref[i]) : T(0.0);
      add_quad<T>(c, cfg.w_tau, NJ, r, (const real *)nullptr);
    }
    if (cfg.w_tau_soft_limits > 0.0) {  // "tau_soft_limits"
      real mn = cfg.tau_limits[0];
      for (int i = 1; i < NJ; ++i) mn = rmin(mn, cfg.tau_limits[i]);
      real margin = rmax(cfg.tau_soft_limit_margin, 0.0);
      margin = rmin(margin, mn - 1.0e-6);
      T r[NJ];
      real lb[NJ], ub[NJ];
      for (int i = 0; i < NJ; ++i) {
        r[i] = with_control ? u[i] : T(0.0);
        lb[i] = -cfg.tau_limits[i] + margin;
        ub[i] = cfg.tau_limits[i] - margin;
      }
      add_barrier<T>(c, cfg.w_tau_soft_limits, NJ, r, lb, ub);
    }
  }
  if (contact && cfg.w_unilateral > 0.0) {  // "unilateral"
    const real inf = rinf();
    if (nc == 1) {
      const T r[1] = {with_control ? out.lam[2] : T(0.0)};
      const real lb[1] = {cfg.friction_margin}, ub[1] = {inf};
      add_barrier<T>(c, cfg.w_unilateral, 1, r, lb, ub);
    } else {
      T r[3] = {out.lam[0], out.lam[1], out.lam[2]};
      if (!with_control) r[0] = r[1] = r[2] = T(0.0);
      const real lb[3] = {-inf, -inf, cfg.friction_margin}, ub[3] = {inf, inf, inf};
      add_barrier<T>(c, cfg.w_unilateral, 3, r, lb, ub);
    }
  }
  if (state_reg) {  // "v_damp"
    T r[14];
    real aw[14];
    for (int i = 0; i < NJ; ++i) {
      r[i] = q[i];
      r[NJ + i] = qd[i];
      aw[i] = 0.0;
      aw[NJ + i] = cfg.v_damp_weights[i];
    }
    add_quad<T>(c, cfg.w_v, 14, r, aw);
  }
  if (contact && cfg.w_vz > 0.0) {  // "vz_damp"
    const T r[6] = {vf.x, vf.y, vf.z, wf.x, wf.y, wf.z};
    const real aw[6] = {0.0, 0.0, 1.0, 0.0, 0.0, 0.0};
    add_quad<T>(c, cfg.w_vz, 6, r, aw);
  }
  {  // "w_damp"
    const T r[6] = {vf.x, vf.y, vf.z, wf.x, wf.y, wf.z};
    const real aw[6] = {0.0, 0.0, 0.0, cfg.w_wdamp_weights[0], cfg.w_wdamp_weights[1], cfg.w_wdamp_weights[2]};
    add_quad<T>(c, cfg.w_wdamp, 6, r, aw);
  }
}

// ----------------------------------------------------------------------------------------------
// action models: IntegratedActionModelEuler (App. A.5) and _AugmentedLPFActionModel (App. A.6)
// ----------------------------------------------------------------------------------------------
struct NodeData {
  int nx = 0;
  real xnext[21];
  real cost = 0.0;
  real lam[3] = {0, 0, 0};
  bool ok = true;
  // derivatives (only after calcDiff)
  real Fx[21][21], Fu[21][7], Lx[21], Lu[7], Lxx[21][21], Lxu[21][7], Luu[7][7];
};

struct Problem {
  ffmpc_model mdl;
  ffmpc_config cfg;
  int mode;
  int N;
  int nx;
  std::vector<real> p_ref, v_ref;  // [(N+1)*3]
  real x_reg_ref[14];
  real tau_ref[7];
  real y_ref[21];
  real x0[21];
};

void lpf_soft(const ffmpc_config &cfg, const real *w, real &cost, real *grad, real *hess) {
  // _AugmentedLPFActionModel._soft_limit_terms (crocoddyl_force_feedback.py:195-209)
  cost = 0.0;
  const real margin = rmax(cfg.tau_soft_limit_margin, 0.0);
  for (int i = 0; i < NJ; ++i) {
    const real lim = rmax(cfg.tau_limits[i] - margin, 1.0e-9);
    const real over = rmax(rabs(w[i]) - lim, 0.0);
    const bool active = over > 0.0;
    cost += over * over;
    grad[i] = active ? over * (w[i] > 0.0 ? 1.0 : (w[i] < 0.0 ? -1.0 : 0.0)) : 0.0;
    hess[i] = active ? 1.0 : 0.0;
  }
  cost *= 0.5;
}

// evaluates knot t at (x, u); diff = false -> calc only.  t == N is the terminal model.
void node_eval(const Problem &P, int t, const real *x, const real *u_in, bool diff, NodeData &nd) {
  const ffmpc_config &cfg = P.cfg;
  const bool ff = cfg.variant == FFMPC_VARIANT_FF;
  const bool terminal = (t == P.N);
  const real dt = cfg.dt_ocp;
  const int nx = P.nx;
  nd.nx = nx;
  NodeRefs ref{&P.p_ref[3 * t], &P.v_ref[3 * t], P.x_reg_ref, P.tau_ref};
  real zero7[NJ] = {0, 0, 0, 0, 0, 0, 0};
  // FF: the inner control is the torque *state*; the wrapper always calls inner.calc(x_mb, tau),
  // so the FF terminal node uses the RUNNING form of the inner Euler model (quirk a16).
  const real *u_inner = ff ? (x + 14) : (terminal ? zero7 : u_in);
  const real *w = ff ? (terminal ? zero7 : u_in) : nullptr;
  const bool with_control = ff ? true : !terminal;
  const bool running_form = ff ? true : !terminal;

  real a[NJ];
  real da[NJ][NZ];
  CostAcc c;
  if (diff) {
    Dual q[NJ], qd[NJ], uu[NJ];
    for (int i = 0; i < NJ; ++i) {
      q[i] = Dual(x[i]); q[i].d[i] = 1.0;
      qd[i] = Dual(x[NJ + i]); qd[i].d[NJ + i] = 1.0;
      uu[i] = Dual(u_inner[i]); uu[i].d[2 * NJ + i] = 1.0;
    }
    DamOut<Dual> o;
    dam_eval<Dual>(P.mdl, cfg, P.mode, terminal, with_control, ref, q, qd, uu, o);
    for (int i = 0; i < NJ; ++i) {
      a[i] = o.a[i].v;
      for (int j = 0; j < NZ; ++j) da[i][j] = o.a[i].d[j];
    }
    for (int i = 0; i < 3; ++i) nd.lam[i] = o.lam[i].v;
    c = o.c;
    nd.ok = o.ok;
  } else {
    DamOut<real> o;
    dam_eval<real>(P.mdl, cfg, P.mode, terminal, with_control, ref, x, x + NJ, u_inner, o);
    for (int i = 0; i < NJ; ++i) a[i] = o.a[i];
    for (int i = 0; i < 3; ++i) nd.lam[i] = o.lam[i];
    c = o.c;
    nd.ok = o.ok;
  }

  // ---- IntegratedActionModelEuler ----
  real xn[14], cost;
  if (running_form) {
    for (int i = 0; i < NJ; ++i) {
      xn[i] = x[i] + (dt * x[NJ + i] + dt * dt * a[i]);
      xn[NJ + i] = x[NJ + i] + dt * a[i];
    }
    cost = dt * c.cost;
  } else {
    for (int i = 0; i < 14; ++i) xn[i] = x[i];
    cost = c.cost;
  }
  const real sc = running_form ? dt : 1.0;
  // inner derivatives over z = (x_mb, u_inner)
  real Fz[14][NZ];
  if (diff) {
    for (int i = 0; i < 14; ++i)
      for (int j = 0; j < NZ; ++j) Fz[i][j] = 0.0;
    if (running_form) {
      for (int i = 0; i < NJ; ++i)
        for (int j = 0; j < NZ; ++j) {
          Fz[i][j] = dt * dt * da[i][j];
          Fz[NJ + i][j] = dt * da[i][j];
        }
      for (int i = 0; i < NJ; ++i) Fz[i][NJ + i] += dt;
    }
    for (int i = 0; i < 14; ++i) Fz[i][i] += 1.0;
  }

  if (!ff) {
    for (int i = 0; i < 14; ++i) nd.xnext[i] = xn[i];
    nd.cost = cost;
    if (diff) {
      for (int i = 0; i < 14; ++i) {
        for (int j = 0; j < 14; ++j) nd.Fx[i][j] = Fz[i][j];
        for (int j = 0; j < NJ; ++j) nd.Fu[i][j] = Fz[i][14 + j];
        nd.Lx[i] = sc * c.Lz[i];
        for (int j = 0; j < 14; ++j) nd.Lxx[i][j] = sc * c.Lzz[i][j];
        for (int j = 0; j < NJ; ++j) nd.Lxu[i][j] = sc * c.Lzz[i][14 + j];
      }
      for (int i = 0; i < NJ; ++i) {
        nd.Lu[i] = sc * c.Lz[14 + i];
        for (int j = 0; j < NJ; ++j) nd.Luu[i][j] = sc * c.Lzz[14 + i][14 + j];
      }
    }
    return;
  }

  // ---- _AugmentedLPFActionModel (crocoddyl_force_feedback.py:211-290) ----
  const real alpha = rmin(rmax(cfg.ff_alpha, 0.0), 0.999999);
  const real beta = 1.0 - alpha;
  for (int i = 0; i < 14; ++i) nd.xnext[i] = xn[i];
  for (int i = 0; i < NJ; ++i) nd.xnext[14 + i] = alpha * x[14 + i] + beta * w[i];
  const real w_y = rmax(cfg.w_y, 0.0), w_w = rmax(cfg.w_w, 0.0), w_soft = rmax(cfg.w_w_soft_limits, 0.0);
  if (w_y > 0.0) {
    real s = 0.0;
    for (int i = 0; i < 21; ++i) {
      const real dy = x[i] - P.y_ref[i];
      s += cfg.y_reg_weights[i] * cfg.y_reg_weights[i] * dy * dy;
    }
    cost += 0.5 * w_y * s;
  }
  if (w_w > 0.0) {
    real s = 0.0;
    for (int i = 0; i < NJ; ++i) s += w[i] * w[i];
    cost += 0.5 * w_w * s;
  }
  real csoft = 0.0, gsoft[NJ], hsoft[NJ];
  if (w_soft > 0.0) {
    lpf_soft(cfg, w, csoft, gsoft, hsoft);
    cost += w_soft * csoft;
  }
  nd.cost = cost;
  if (!diff) return;
  for (int i = 0; i < 21; ++i) {
    for (int j = 0; j < 21; ++j) {
      nd.Fx[i][j] = 0.0;
      nd.Lxx[i][j] = sc * c.Lzz[i][j];
    }
    for (int j = 0; j < NJ; ++j) {
      nd.Fu[i][j] = 0.0;
      nd.Lxu[i][j] = 0.0;
    }
    nd.Lx[i] = sc * c.Lz[i];
  }
  for (int i = 0; i < 14; ++i)
    for (int j = 0; j < 21; ++j) nd.Fx[i][j] = Fz[i][j];
  for (int i = 0; i < NJ; ++i) {
    nd.Fx[14 + i][14 + i] = alpha;
    nd.Fu[14 + i][i] = beta;
    nd.Lu[i] = 0.0;
    for (int j = 0; j < NJ; ++j) nd.Luu[i][j] = 0.0;
  }
  if (w_y > 0.0)
    for (int i = 0; i < 21; ++i) {
      const real wy2 = cfg.y_reg_weights[i] * cfg.y_reg_weights[i];
      nd.Lx[i] += w_y * wy2 * (x[i] - P.y_ref[i]);
      nd.Lxx[i][i] += w_y * wy2;
    }
  if (w_w > 0.0)
    for (int i = 0; i < NJ; ++i) {
      nd.Lu[i] += w_w * w[i];
      nd.Luu[i][i] += w_w;
    }
  if (w_soft > 0.0)
    for (int i = 0; i < NJ; ++i) {
      nd.Lu[i] += w_soft * gsoft[i];
      nd.Luu[i][i] += w_soft * hsoft[i];
    }
}

// ----------------------------------------------------------------------------------------------
// BoxQP (Tassa 2014 projected Newton; crocoddyl::BoxQP) -- SURVEY App. A.8
// ----------------------------------------------------------------------------------------------
struct BoxQPSol {
  real x[NJ];
  int free_idx[NJ], clamped_idx[NJ];
  int nf = 0, ncl = 0;
  real Hff_inv[NJ][NJ];
  bool ok = true;
};

bool boxqp_free_inverse(const real H[NJ][NJ], real reg, BoxQPSol &s, real Lff[NJ][NJ]) {
  for (int i = 0; i < s.nf; ++i)
    for (int j = 0; j < s.nf; ++j) Lff[i][j] = H[s.free_idx[i]][s.free_idx[j]] + ((i == j) ? reg : 0.0);
  if (!chol<real, NJ>(Lff, s.nf)) return false;
  for (int j = 0; j < s.nf; ++j) {
    real e[NJ];
    for (int i = 0; i < s.nf; ++i) e[i] = (i == j) ? 1.0 : 0.0;
    chol_solve<real, NJ>(Lff, s.nf, e);
    for (int i = 0; i < s.nf; ++i) s.Hff_inv[i][j] = e[i];
  }
  return true;
}

void boxqp_solve(const ffmpc_config &cfg, const real H[NJ][NJ], const real *q, const real *lb, const real *ub,
                 const real *xinit, BoxQPSol &s) {
  const int n = NJ;
  real x[NJ], g[NJ], dx[NJ], xnew[NJ], Lff[NJ][NJ];
  for (int i = 0; i < n; ++i) x[i] = rmax(rmin(xinit[i], ub[i]), lb[i]);
  s.ok = true;
  unsigned inv_mask = 0xffffffffu;
  for (int k = 0; k < cfg.boxqp_maxiter; ++k) {
    s.nf = s.ncl = 0;
    for (int i = 0; i < n; ++i) {
      real gi = q[i];
      for (int j = 0; j < n; ++j) gi += H[i][j] * x[j];
      g[i] = gi;
    }
    real gnorm = 0.0;
    for (int j = 0; j < n; ++j) {
      gnorm = rmax(gnorm, rabs(g[j]));
      if ((x[j] == lb[j] && g[j] > 0.0) || (x[j] == ub[j] && g[j] < 0.0)) s.clamped_idx[s.ncl++] = j;
      else s.free_idx[s.nf++] = j;
    }
    unsigned mask = 0;
    for (int i = 0; i < s.nf; ++i) mask |= 1u << s.free_idx[i];
    // convergence check
    if (gnorm <= cfg.boxqp_th_grad || s.nf == 0) {
      // crocoddyl::BoxQP factorises here only when k == 0 and otherwise returns the Hff_inv of the
      // previous iteration; that is the same matrix unless the active set changed in the very last
      // step, in which case upstream would hand back a stale block.  We re-factorise in that
      // corner case so that Hff_inv always belongs to the returned free set.
      if (k == 0 || mask != inv_mask) {
        if (!boxqp_free_inverse(H, cfg.boxqp_reg, s, Lff)) s.ok = false;
        inv_mask = mask;
      }
      break;
    }
    // Newton step on the free subspace
    if (!boxqp_free_inverse(H, cfg.boxqp_reg, s, Lff)) {
      s.ok = false;
      break;
    }
    inv_mask = mask;
    real dxf[NJ];
    for (int i = 0; i < s.nf; ++i) {
      const int fi = s.free_idx[i];
      real r = -q[fi];
      for (int j = 0; j < s.ncl; ++j) r -= H[fi][s.clamped_idx[j]] * x[s.clamped_idx[j]];
      dxf[i] = r;
    }
    chol_solve<real, NJ>(Lff, s.nf, dxf);
    for (int i = 0; i < n; ++i) dx[i] = 0.0;
    for (int i = 0; i < s.nf; ++i) dx[s.free_idx[i]] = dxf[i] - x[s.free_idx[i]];
    // line search
    auto fval = [&](const real *y) {
      real f = 0.0;
      for (int i = 0; i < n; ++i) {
        real hy = 0.0;
        for (int j = 0; j < n; ++j) hy += H[i][j] * y[j];
        f += 0.5 * y[i] * hy + q[i] * y[i];
      }
      return f;
    };
    const real fold = fval(x);
    bool changed = false;
    real steplength = 1.0;
    for (int ia = 0; ia < cfg.n_alphas; ++ia, steplength *= 0.5) {
      for (int i = 0; i < n; ++i) xnew[i] = rmax(rmin(x[i] + steplength * dx[i], ub[i]), lb[i]);
      const real fnew = fval(xnew);
      real gd = 0.0;
      for (int i = 0; i < n; ++i) gd += g[i] * (x[i] - xnew[i]);
      if (fold - fnew > cfg.boxqp_th_acceptstep * gd) {
        for (int i = 0; i < n; ++i) {
          if (xnew[i] != x[i]) changed = true;
          x[i] = xnew[i];
        }
        break;
      }
    }
    // The iteration map is a deterministic function of x: once x stops changing every further
    // iteration (up to maxiter) reproduces the same active set and Hff_inv, so stopping here is
    // output-identical to running all `boxqp_maxiter` iterations.
    if (!changed) break;
  }
  for (int i = 0; i < n; ++i) s.x[i] = x[i];
}

// ----------------------------------------------------------------------------------------------
// SolverFDDP / SolverBoxFDDP -- SURVEY App. A.7 / A.8
// ----------------------------------------------------------------------------------------------
struct Solver {
  const Problem &P;
  int N, nx;
  std::vector<real> xs, us, xs_try, us_try, fs;             // [(N+1)*nx], [N*7]
  std::vector<NodeData> nd;                                    // N+1 knots
  std::vector<real> Vxx, Vx, K, k, Qu, Quuk;                  // [(N+1)*nx*nx], [(N+1)*nx], [N*7*nx], [N*7] ...
  real cost = 0, cost_try = 0, preg, dg = 0, dq = 0, d1 = 0, d2 = 0, dV = 0, dVexp = 0, stop = 0, steplength = 1;
  bool is_feasible = false, was_feasible = false;
  int iter = 0;
  real f_pred[2][3];

  explicit Solver(const Problem &p) : P(p), N(p.N), nx(p.nx) {
    xs.assign((N + 1) * nx, 0.0); xs_try = xs; fs = xs;
    us.assign(N * NJ, 0.0); us_try = us;
    nd.resize(N + 1);
    Vxx.assign((N + 1) * nx * nx, 0.0); Vx.assign((N + 1) * nx, 0.0);
    K.assign(N * NJ * nx, 0.0); k.assign(N * NJ, 0.0); Qu = k; Quuk = k;
    for (auto &r : f_pred) for (auto &v : r) v = rnan();
  }

  void record_force(int t, const NodeData &d) {
    if (t < 2 && P.mode == FFMPC_MODE_CONTACT)
      for (int i = 0; i < 3; ++i) f_pred[t][i] = d.lam[i];
  }

  void calc_diff() {
    // problem.calc (iter 0) + problem.calcDiff: every knot re-evaluated at (xs, us)
    cost = 0.0;
    for (int t = 0; t <= N; ++t) {
      node_eval(P, t, &xs[t * nx], t < N ? &us[t * NJ] : nullptr, true, nd[t]);
      cost += nd[t].cost;
      if (iter == 0) record_force(t, nd[t]);
    }
    for (int i = 0; i < nx; ++i) fs[i] = P.x0[i] - xs[i];
    for (int t = 0; t < N; ++t)
      for (int i = 0; i < nx; ++i) fs[(t + 1) * nx + i] = nd[t].xnext[i] - xs[(t + 1) * nx + i];
  }

  bool finite_vec(const real *v, int n) {
    for (int i = 0; i < n; ++i) if (!rfinite(v[i])) return false;
    return true;
  }

  bool backward_pass() {
    const int n = nx;
    real *VxxN = &Vxx[N * n * n], *VxN = &Vx[N * n];
    for (int i = 0; i < n; ++i) {
      for (int j = 0; j < n; ++j) VxxN[i * n + j] = nd[N].Lxx[i][j];
      VxxN[i * n + i] += preg;
      VxN[i] = nd[N].Lx[i];
    }
    if (!is_feasible)
      for (int i = 0; i < n; ++i) {
        real s = 0.0;
        for (int j = 0; j < n; ++j) s += VxxN[i * n + j] * fs[N * n + j];
        VxN[i] += s;
      }
    std::vector<real> FxTV(n * n), Qxx(n * n), Qxu(n * NJ), FuTV(NJ * n);
    for (int t = N - 1; t >= 0; --t) {
      const NodeData &d = nd[t];
      const real *Vp = &Vxx[(t + 1) * n * n], *vp = &Vx[(t + 1) * n];
      real Quu[NJ][NJ], Qx[21], *Qu_t = &Qu[t * NJ];
      for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
          real s = 0.0;
          for (int l = 0; l < n; ++l) s += d.Fx[l][i] * Vp[l * n + j];
          FxTV[i * n + j] = s;
        }
      for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) {
          real s = d.Lxx[i][j];
          for (int l = 0; l < n; ++l) s += FxTV[i * n + l] * d.Fx[l][j];
          Qxx[i * n + j] = s;
        }
        real s = d.Lx[i];
        for (int l = 0; l < n; ++l) s += d.Fx[l][i] * vp[l];
        Qx[i] = s;
      }
      for (int a = 0; a < NJ; ++a)
        for (int j = 0; j < n; ++j) {
          real s = 0.0;
          for (int l = 0; l < n; ++l) s += d.Fu[l][a] * Vp[l * n + j];
          FuTV[a * n + j] = s;
        }
      for (int i = 0; i < n; ++i)
        for (int a = 0; a < NJ; ++a) {
          real s = d.Lxu[i][a];
          for (int l = 0; l < n; ++l) s += FxTV[i * n + l] * d.Fu[l][a];
          Qxu[i * NJ + a] = s;
        }
      for (int a = 0; a < NJ; ++a) {
        for (int b = 0; b < NJ; ++b) {
          real s = d.Luu[a][b];
          for (int l = 0; l < n; ++l) s += FuTV[a * n + l] * d.Fu[l][b];
          Quu[a][b] = s;
        }
        Quu[a][a] += preg;
        real s = d.Lu[a];
        for (int l = 0; l < n; ++l) s += d.Fu[l][a] * vp[l];
        Qu_t[a] = s;
      }
      // computeGains
      real *Kt = &K[t * NJ * n], *kt = &k[t * NJ];
      const bool box = P.cfg.use_box_fddp && is_feasible;
      if (!box) {
        real L[NJ][NJ];
        for (int a = 0; a < NJ; ++a) for (int b = 0; b < NJ; ++b) L[a][b] = Quu[a][b];
        if (!chol<real, NJ>(L, NJ)) return false;
        for (int j = 0; j < n; ++j) {
          real col[NJ];
          for (int a = 0; a < NJ; ++a) col[a] = Qxu[j * NJ + a];
          chol_solve<real, NJ>(L, NJ, col);
          for (int a = 0; a < NJ; ++a) Kt[a * n + j] = col[a];
        }
        for (int a = 0; a < NJ; ++a) kt[a] = Qu_t[a];
        chol_solve<real, NJ>(L, NJ, kt);
      } else {
        real lb[NJ], ub[NJ];
        for (int a = 0; a < NJ; ++a) {
          lb[a] = -P.cfg.tau_limits[a] - us[t * NJ + a];
          ub[a] = P.cfg.tau_limits[a] - us[t * NJ + a];
        }
        BoxQPSol s;
        boxqp_solve(P.cfg, Quu, Qu_t, lb, ub, kt, s);
        if (!s.ok) return false;
        real Qinv[NJ][NJ];
        for (int a = 0; a < NJ; ++a) for (int b = 0; b < NJ; ++b) Qinv[a][b] = 0.0;
        for (int i = 0; i < s.nf; ++i)
          for (int j = 0; j < s.nf; ++j) Qinv[s.free_idx[i]][s.free_idx[j]] = s.Hff_inv[i][j];
        for (int a = 0; a < NJ; ++a)
          for (int j = 0; j < n; ++j) {
            real acc = 0.0;
            for (int b = 0; b < NJ; ++b) acc += Qinv[a][b] * Qxu[j * NJ + b];
            Kt[a * n + j] = acc;
          }
        for (int a = 0; a < NJ; ++a) kt[a] = -s.x[a];
        for (int i = 0; i < s.ncl; ++i) Qu_t[s.clamped_idx[i]] = 0.0;
      }
      // value function
      real *Vt = &Vxx[t * n * n], *vt = &Vx[t * n], *Quuk_t = &Quuk[t * NJ];
      for (int a = 0; a < NJ; ++a) {
        real s = 0.0;
        for (int b = 0; b < NJ; ++b) s += Quu[a][b] * kt[b];
        Quuk_t[a] = s;
      }
      for (int i = 0; i < n; ++i) {
        real s = Qx[i];
        for (int a = 0; a < NJ; ++a) s += Kt[a * n + i] * Quuk_t[a];
        for (int a = 0; a < NJ; ++a) s -= 2.0 * (Kt[a * n + i] * Qu_t[a]);
        vt[i] = s;
        for (int j = 0; j < n; ++j) {
          real v = Qxx[i * n + j];
          for (int a = 0; a < NJ; ++a) v -= Qxu[i * NJ + a] * Kt[a * n + j];
          Vt[i * n + j] = v;
        }
      }
      for (int i = 0; i < n; ++i)
        for (int j = i; j < n; ++j) {
          const real v = 0.5 * (Vt[i * n + j] + Vt[j * n + i]);
          Vt[i * n + j] = Vt[j * n + i] = v;
        }
      for (int i = 0; i < n; ++i) Vt[i * n + i] += preg;
      if (!is_feasible)
        for (int i = 0; i < n; ++i) {
          real s = 0.0;
          for (int j = 0; j < n; ++j) s += Vt[i * n + j] * fs[t * n + j];
          vt[i] += s;
        }
      if (!finite_vec(vt, n) || !finite_vec(Vt, n * n)) return false;
    }
    return true;
  }

  void update_expected_improvement() {
    const int n = nx;
    dg = 0.0; dq = 0.0;
    auto fVf = [&](int t) {
      real s = 0.0;
      for (int i = 0; i < n; ++i) {
        real r = 0.0;
        for (int j = 0; j < n; ++j) r += Vxx[t * n * n + i * n + j] * fs[t * n + j];
        s += fs[t * n + i] * r;
      }
      return s;
    };
    if (!is_feasible) {
      real s = 0.0;
      for (int i = 0; i < n; ++i) s += Vx[N * n + i] * fs[N * n + i];
      dg -= s;
      dq += fVf(N);
    }
    for (int t = 0; t < N; ++t) {
      real a = 0.0, b = 0.0;
      for (int i = 0; i < NJ; ++i) {
        a += Qu[t * NJ + i] * k[t * NJ + i];
        b += k[t * NJ + i] * Quuk[t * NJ + i];
      }
      dg += a;
      dq -= b;
      if (!is_feasible) {
        real s = 0.0;
        for (int i = 0; i < n; ++i) s += Vx[t * n + i] * fs[t * n + i];
        dg -= s;
        dq += fVf(t);
      }
    }
  }

  void expected_improvement() {
    const int n = nx;
    real dv = 0.0;
    if (!is_feasible) {
      auto term = [&](int t) {
        real s = 0.0;
        for (int i = 0; i < n; ++i) {
          real r = 0.0;
          for (int j = 0; j < n; ++j) r += Vxx[t * n * n + i * n + j] * (xs[t * n + j] - xs_try[t * n + j]);
          s += fs[t * n + i] * r;
        }
        return s;
      };
      dv -= term(N);
      for (int t = 0; t < N; ++t) dv -= term(t);
    }
    d1 = dg + dv;
    d2 = dq - 2.0 * dv;
  }

  // returns false on a NaN/inf rollout ("forward_error")
  bool forward_pass(real alpha) {
    const int n = nx;
    cost_try = 0.0;
    real xnext[21];
    for (int i = 0; i < n; ++i) xnext[i] = P.x0[i];
    const bool close = is_feasible || alpha == 1.0;
    NodeData d;
    for (int t = 0; t < N; ++t) {
      real *xt = &xs_try[t * n], *ut = &us_try[t * NJ];
      for (int i = 0; i < n; ++i) xt[i] = close ? xnext[i] : xnext[i] + fs[t * n + i] * (alpha - 1.0);
      for (int a = 0; a < NJ; ++a) {
        real s = us[t * NJ + a] - k[t * NJ + a] * alpha;
        real kd = 0.0;
        for (int j = 0; j < n; ++j) kd += K[t * NJ * n + a * n + j] * (xt[j] - xs[t * n + j]);
        s -= kd;
        if (P.cfg.use_box_fddp) s = rmin(rmax(s, -P.cfg.tau_limits[a]), P.cfg.tau_limits[a]);
        ut[a] = s;
      }
      node_eval(P, t, xt, ut, false, d);
      record_force(t, d);
      for (int i = 0; i < n; ++i) xnext[i] = d.xnext[i];
      cost_try += d.cost;
      if (!rfinite(cost_try) || !finite_vec(xnext, n)) return false;
    }
    real *xt = &xs_try[N * n];
    for (int i = 0; i < n; ++i) xt[i] = close ? xnext[i] : xnext[i] + fs[N * n + i] * (alpha - 1.0);
    node_eval(P, N, xt, nullptr, false, d);
    cost_try += d.cost;
    if (!rfinite(cost_try)) return false;
    return true;
  }

  void increase_reg() { preg *= P.cfg.reg_incfactor; if (preg > P.cfg.reg_max) preg = P.cfg.reg_max; }
  void decrease_reg() { preg /= P.cfg.reg_decfactor; if (preg < P.cfg.reg_min) preg = P.cfg.reg_min; }

  bool solve(const real *xs_init, const real *us_init, int maxiter, bool fixed_iters) {
    const ffmpc_config &c = P.cfg;
    std::copy(xs_init, xs_init + (N + 1) * nx, xs.begin());
    std::copy(us_init, us_init + N * NJ, us.begin());
    for (int i = 0; i < nx; ++i) xs_try[i] = P.x0[i];
    is_feasible = false;
    was_feasible = false;
    preg = c.reg_min;
    bool recalc = true;
    for (iter = 0; iter < maxiter; ++iter) {
      while (true) {
        if (recalc) calc_diff();
        if (backward_pass()) break;
        recalc = false;
        increase_reg();
        if (preg == c.reg_max) return false;
      }
      update_expected_improvement();
      recalc = false;
      steplength = 1.0;
      g_iterations.fetch_add(1, std::memory_order_relaxed);
      for (int ia = 0; ia < c.n_alphas; ++ia) {
        steplength = rldexp(1.0, -ia);
        g_rollouts.fetch_add(1, std::memory_order_relaxed);
        g_trial_hist[ia].fetch_add(1, std::memory_order_relaxed);
        if (!forward_pass(steplength)) continue;
        dV = cost - cost_try;
        expected_improvement();
        dVexp = steplength * (d1 + 0.5 * steplength * d2);
        bool accept = false;
        if (dVexp >= 0.0) {
          if (d1 < c.th_grad || dV > c.th_acceptstep * dVexp) accept = true;
        } else {
          if (!is_feasible && dV > c.th_acceptnegstep * dVexp) accept = true;
        }
        if (accept) {
          was_feasible = is_feasible;
          xs = xs_try;
          us = us_try;
          is_feasible = was_feasible || (steplength == 1.0);
          cost = cost_try;
          recalc = true;
          break;
        }
      }
      if (steplength > c.th_stepdec) decrease_reg();
      if (steplength <= c.th_stepinc) {
        increase_reg();
        if (preg == c.reg_max) return false;
      }
      stop = rabs(d1 + 0.5 * d2);
      if (!fixed_iters && was_feasible && stop < c.th_stop) return true;
    }
    return false;
  }
};

void make_problem(const ffmpc_config &cfg, const ffmpc_model &mdl, const ffmpc_batch_in &in, int b, Problem &P) {
  P.mdl = mdl;
  P.cfg = cfg;
  P.N = cfg.horizon;
  P.nx = cfg.variant == FFMPC_VARIANT_FF ? 21 : 14;
  P.mode = in.mode[b];
  const int N = P.N;
  P.p_ref.assign(in.p_ref + (size_t)b * (N + 1) * 3, in.p_ref + (size_t)(b + 1) * (N + 1) * 3);
  P.v_ref.assign(in.v_ref + (size_t)b * (N + 1) * 3, in.v_ref + (size_t)(b + 1) * (N + 1) * 3);
  for (int i = 0; i < P.nx; ++i) P.x0[i] = in.x0[(size_t)b * P.nx + i];
  for (int i = 0; i < 14; ++i) P.x_reg_ref[i] = in.x_reg_ref[(size_t)b * 14 + i];
  if (in.tau_ref) {
    for (int i = 0; i < NJ; ++i) P.tau_ref[i] = in.tau_ref[(size_t)b * NJ + i];
  } else {
    // _compute_tau_reference "gravity_x0": pin.rnea(q0, 0, 0)  (crocoddyl_classical.py:447-460)
    Kin<real> k;
    real zero[NJ] = {0, 0, 0, 0, 0, 0, 0};
    kinematics<real>(mdl, P.x0, zero, k);
    rnea<real>(mdl, k, zero, nullptr, P.tau_ref);
  }
  for (int i = 0; i < 21; ++i) P.y_ref[i] = (i < P.nx) ? P.x0[i] : 0.0;  // y_ref = y0 (FF :782)
}

std::string g_err;

}  // namespace

// ==============================================================================================
// C API (same shapes as include/ffmpc.h, HOST pointers, prefix ffmpc_oracle_)
// ==============================================================================================
extern "C" {

void ffmpc_oracle_default_config(ffmpc_config *c, int variant) {
  std::memset(c, 0, sizeof(*c));
  c->variant = variant;
  c->horizon = 20;
  c->contact_nc = 1;
  c->use_box_fddp = 1;
  c->use_inner_state_reg = 1;
  c->use_inner_tau_reg = 1;
  c->dt_ocp = 0.01;
  c->w_ee_pos = 2.0e2; c->w_ee_ori = 1.0e1;
  c->ori_weights[0] = 2.0; c->ori_weights[1] = 2.0; c->ori_weights[2] = 0.15;
  for (int i = 0; i < 9; ++i) c->R_des[i] = (i % 4 == 0) ? 1.0 : 0.0;
  c->w_posture = 5.0e-1; c->w_v = 2.5e-1;
  const double vd[7] = {1.0, 1.0, 1.0, 1.0, 0.4, 0.4, 0.4};
  const double tl[7] = {87, 87, 87, 87, 12, 12, 12};
  for (int i = 0; i < 7; ++i) {
    c->v_damp_weights[i] = vd[i];
    c->tau_limits[i] = tl[i];
    c->q_soft_lb[i] = -std::numeric_limits<double>::infinity();
    c->q_soft_ub[i] = std::numeric_limits<double>::infinity();
  }
  c->w_q_soft_limits = (variant == FFMPC_VARIANT_FF) ? 8.0 : 0.0;
  c->w_wdamp = 2.0e1;
  c->w_wdamp_weights[0] = 1.5; c->w_wdamp_weights[1] = 1.5; c->w_wdamp_weights[2] = 0.2;
  c->w_tau = 1.0e-3;
  c->w_tau_soft_limits = (variant == FFMPC_VARIANT_FF) ? 1.5 : 0.0;
  c->tau_soft_limit_margin = 0.2;
  c->z_press = 0.0020;
  c->w_tangent_pos = 2.0e2; c->w_tangent_vel = 1.0e2; c->w_plane_z = 0.0; c->w_vz = 0.0;
  c->w_unilateral = 5.0e1; c->friction_margin = 1e-3; c->w_fn = 2.0e1; c->fn_des = 8.0;
  c->contact_gains[0] = 0.0; c->contact_gains[1] = 60.0; c->contact_inv_damping = 1.0e-8;
  if (variant == FFMPC_VARIANT_FF) {
  c->ff_alpha = std::exp(-2.0 * M_PI * 18.0 * 0.01);
  c->w_w = 8.0e-4; c->w_w_soft_limits = 2.0; c->w_y = 3.0e-3;
  const double yq[7] = {0.2, 0.2, 0.2, 0.2, 0.1, 0.1, 0.1}, yv[7] = {0.08, 0.08, 0.08, 0.08, 0.05, 0.05, 0.05},
               yt[7] = {0.35, 0.35, 0.35, 0.35, 0.2, 0.2, 0.2};
  for (int i = 0; i < 7; ++i) { c->y_reg_weights[i] = yq[i]; c->y_reg_weights[7 + i] = yv[i]; c->y_reg_weights[14 + i] = yt[i]; }
  }
  c->reg_min = 1e-9; c->reg_max = 1e9; c->reg_incfactor = 10.0; c->reg_decfactor = 10.0;
  c->th_stepdec = 0.5; c->th_stepinc = 0.01; c->th_acceptstep = 0.1; c->th_acceptnegstep = 2.0; c->th_grad = 1e-12;
  c->th_stop = 5e-5;  // Box solvers; plain SolverFDDP uses 1e-9
  c->n_alphas = 10;
  c->boxqp_maxiter = 100; c->boxqp_th_acceptstep = 0.1; c->boxqp_th_grad = 1e-5; c->boxqp_reg = 0.0;
  c->w_friction_cone = 0.0; c->mu = 0.6;
}

const char *ffmpc_oracle_last_error(void) { return g_err.c_str(); }

static int check_cfg(const ffmpc_config *cfg) {
  if (!cfg || cfg->horizon < 1 || (cfg->contact_nc != 1 && cfg->contact_nc != 3) || !(cfg->dt_ocp > 0.0) ||
      (cfg->variant != FFMPC_VARIANT_CLASSICAL && cfg->variant != FFMPC_VARIANT_FF) || cfg->n_alphas < 1) {
    g_err = "invalid ffmpc_config";
    return FFMPC_ERR_ARG;
  }
  return FFMPC_OK;
}

// Batched solve; std::thread pool over instances when nthreads > 1.
int ffmpc_oracle_solve_batch(const ffmpc_config *cfg, const ffmpc_model *mdl, const ffmpc_batch_in *in, int max_iters,
                             int fixed_iters, const ffmpc_batch_out *out, int nthreads) {
  if (int e = check_cfg(cfg)) return e;
  if (!in || !in->mode || !in->x0 || !in->p_ref || !in->v_ref || !in->x_reg_ref) { g_err = "null input"; return FFMPC_ERR_ARG; }
  const int N = cfg->horizon, nx = cfg->variant == FFMPC_VARIANT_FF ? 21 : 14, B = in->B;
  if (nthreads < 1) nthreads = 1;
  auto work = [&](int b) {
    Problem P;
    make_problem(*cfg, *mdl, *in, b, P);
    Solver S(P);
    std::vector<real> xs0((N + 1) * nx), us0(N * NJ);
    if (in->xs_init) std::copy(in->xs_init + (size_t)b * (N + 1) * nx, in->xs_init + (size_t)(b + 1) * (N + 1) * nx, xs0.begin());
    else for (int t = 0; t <= N; ++t) for (int i = 0; i < nx; ++i) xs0[t * nx + i] = P.x0[i];
    if (in->us_init) std::copy(in->us_init + (size_t)b * N * NJ, in->us_init + (size_t)(b + 1) * N * NJ, us0.begin());
    else for (int t = 0; t < N; ++t) for (int i = 0; i < NJ; ++i) us0[t * NJ + i] = (nx == 21) ? P.x0[14 + i] : P.tau_ref[i];
    const bool ok = S.solve(xs0.data(), us0.data(), max_iters, fixed_iters != 0);
    if (out->xs) copy_out(S.xs.begin(), S.xs.end(), out->xs + (size_t)b * (N + 1) * nx);
    if (out->us) copy_out(S.us.begin(), S.us.end(), out->us + (size_t)b * N * NJ);
    if (out->K) copy_out(S.K.begin(), S.K.end(), out->K + (size_t)b * N * NJ * nx);
    if (out->k) copy_out(S.k.begin(), S.k.end(), out->k + (size_t)b * N * NJ);
    if (out->gaps) copy_out(S.fs.begin(), S.fs.end(), out->gaps + (size_t)b * (N + 1) * nx);
    if (out->cost) out->cost[b] = to_d(S.cost);
    if (out->stop) out->stop[b] = to_d(S.stop);
    if (out->iters) out->iters[b] = S.iter;
    if (out->ok) out->ok[b] = ok ? 1 : 0;
    if (out->f_pred) for (int t = 0; t < 2; ++t) for (int i = 0; i < 3; ++i) out->f_pred[(size_t)b * 6 + t * 3 + i] = to_d(S.f_pred[t][i]);
#ifdef FFMPC_ORACLE_COUNT
    g_flops.fetch_add(g_flops_tl, std::memory_order_relaxed);
    g_flops_tl = 0;
#endif
  };
  if (nthreads == 1) {
    for (int b = 0; b < B; ++b) work(b);
  } else {
    std::atomic<int> next(0);
    std::vector<std::thread> pool;
    for (int w = 0; w < nthreads; ++w)
      pool.emplace_back([&]() { for (int b = next++; b < B; b = next++) work(b); });
    for (auto &th : pool) th.join();
  }
  return FFMPC_OK;
}

// ActionModel.calc + calcDiff at M points (instance inst[i], knot knot[i]).
int ffmpc_oracle_node_eval(const ffmpc_config *cfg, const ffmpc_model *mdl, const ffmpc_batch_in *in, int M, const int *inst,
                           const int *knot, const double *x, const double *u, const ffmpc_node_out *out) {
  if (int e = check_cfg(cfg)) return e;
  const int nx = cfg->variant == FFMPC_VARIANT_FF ? 21 : 14;
  for (int m = 0; m < M; ++m) {
    Problem P;
    make_problem(*cfg, *mdl, *in, inst[m], P);
    if (knot[m] < 0 || knot[m] > P.N) { g_err = "knot out of range"; return FFMPC_ERR_ARG; }
    NodeData d;
    real xr[21], ur[NJ];
    for (int i = 0; i < nx; ++i) xr[i] = x[(size_t)m * nx + i];
    for (int i = 0; i < NJ; ++i) ur[i] = u[(size_t)m * NJ + i];
    node_eval(P, knot[m], xr, ur, true, d);
    if (out->xnext) for (int i = 0; i < nx; ++i) out->xnext[(size_t)m * nx + i] = to_d(d.xnext[i]);
    if (out->cost) out->cost[m] = to_d(d.cost);
    if (out->force) for (int i = 0; i < 3; ++i) out->force[(size_t)m * 3 + i] = to_d(d.lam[i]);
    for (int i = 0; i < nx; ++i) {
      for (int j = 0; j < nx; ++j) {
        if (out->Fx) out->Fx[((size_t)m * nx + i) * nx + j] = to_d(d.Fx[i][j]);
        if (out->Lxx) out->Lxx[((size_t)m * nx + i) * nx + j] = to_d(d.Lxx[i][j]);
      }
      for (int j = 0; j < NJ; ++j) {
        if (out->Fu) out->Fu[((size_t)m * nx + i) * NJ + j] = to_d(d.Fu[i][j]);
        if (out->Lxu) out->Lxu[((size_t)m * nx + i) * NJ + j] = to_d(d.Lxu[i][j]);
      }
      if (out->Lx) out->Lx[(size_t)m * nx + i] = to_d(d.Lx[i]);
    }
    for (int i = 0; i < NJ; ++i) {
      if (out->Lu) out->Lu[(size_t)m * NJ + i] = to_d(d.Lu[i]);
      for (int j = 0; j < NJ; ++j) if (out->Luu) out->Luu[((size_t)m * NJ + i) * NJ + j] = to_d(d.Luu[i][j]);
    }
  }
  return FFMPC_OK;
}

// value-only action model calc (used for finite-difference self-validation)
int ffmpc_oracle_node_calc(const ffmpc_config *cfg, const ffmpc_model *mdl, const ffmpc_batch_in *in, int M, const int *inst,
                           const int *knot, const double *x, const double *u, double *xnext, double *cost, double *force) {
  if (int e = check_cfg(cfg)) return e;
  const int nx = cfg->variant == FFMPC_VARIANT_FF ? 21 : 14;
  for (int m = 0; m < M; ++m) {
    Problem P;
    make_problem(*cfg, *mdl, *in, inst[m], P);
    NodeData d;
    real xr[21], ur[NJ];
    for (int i = 0; i < nx; ++i) xr[i] = x[(size_t)m * nx + i];
    for (int i = 0; i < NJ; ++i) ur[i] = u[(size_t)m * NJ + i];
    node_eval(P, knot[m], xr, ur, false, d);
    if (xnext) for (int i = 0; i < nx; ++i) xnext[(size_t)m * nx + i] = to_d(d.xnext[i]);
    if (cost) cost[m] = to_d(d.cost);
    if (force) for (int i = 0; i < 3; ++i) force[(size_t)m * 3 + i] = to_d(d.lam[i]);
  }
  return FFMPC_OK;
}

// rigid-body primitives for the identity tests: FK of the EE frame, RNEA, CRBA, forward dynamics
static void to_real7(const double *a, real *r) { for (int i = 0; i < NJ; ++i) r[i] = a ? real(a[i]) : real(0.0); }

void ffmpc_oracle_fk(const ffmpc_model *mdl, const double *q, double *p_ee, double *R_ee) {
  Kin<real> k;
  real qr[NJ], zero[NJ];
  to_real7(q, qr); to_real7(nullptr, zero);
  kinematics<real>(*mdl, qr, zero, k);
  p_ee[0] = to_d(k.pf.x); p_ee[1] = to_d(k.pf.y); p_ee[2] = to_d(k.pf.z);
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) R_ee[3 * i + j] = to_d(k.Rf.m[i][j]);
}
void ffmpc_oracle_rnea(const ffmpc_model *mdl, const double *q, const double *v, const double *a, double *tau) {
  Kin<real> k;
  real qr[NJ], vr[NJ], ar[NJ], tr[NJ];
  to_real7(q, qr); to_real7(v, vr); to_real7(a, ar);
  kinematics<real>(*mdl, qr, vr, k);
  rnea<real>(*mdl, k, ar, nullptr, tr);
  for (int i = 0; i < NJ; ++i) tau[i] = to_d(tr[i]);
}
void ffmpc_oracle_crba(const ffmpc_model *mdl, const double *q, double *M) {
  Kin<real> k;
  real qr[NJ], zero[NJ];
  to_real7(q, qr); to_real7(nullptr, zero);
  kinematics<real>(*mdl, qr, zero, k);
  real Mm[NJ][NJ];
  crba<real>(k, Mm);
  for (int i = 0; i < NJ; ++i) for (int j = 0; j < NJ; ++j) M[i * NJ + j] = to_d(Mm[i][j]);
}
// EE frame Jacobian (LOCAL_WORLD_ALIGNED, 6x7, rows [linear; angular]) and frame velocity
void ffmpc_oracle_frame_jacobian(const ffmpc_model *mdl, const double *q, double *J6) {
  Kin<real> k;
  real qr[NJ], zero[NJ];
  to_real7(q, qr); to_real7(nullptr, zero);
  kinematics<real>(*mdl, qr, zero, k);
  for (int j = 0; j < NJ; ++j) {
    const V3<real> jl = k.J[j].l + cross(k.J[j].a, k.pf);
    J6[0 * NJ + j] = to_d(jl.x); J6[1 * NJ + j] = to_d(jl.y); J6[2 * NJ + j] = to_d(jl.z);
    J6[3 * NJ + j] = to_d(k.J[j].a.x); J6[4 * NJ + j] = to_d(k.J[j].a.y); J6[5 * NJ + j] = to_d(k.J[j].a.z);
  }
}

// counting build only: scalar operations executed since the last reset, summed over the worker threads (-1: not a counting build)
long long ffmpc_oracle_flops(int reset) {
#ifdef FFMPC_ORACLE_COUNT
  const long long v = g_flops.load();
  if (reset) g_flops = 0;
  return v;
#else
  (void)reset;
  return -1;
#endif
}

}  // extern "C"

extern "C" void ffmpc_oracle_rollout_stats(long long *rollouts, long long *iterations, long long *trial_hist32, int reset) {
  if (rollouts) *rollouts = g_rollouts.load();
  if (iterations) *iterations = g_iterations.load();
  if (trial_hist32)
    for (int i = 0; i < 32; ++i) trial_hist32[i] = g_trial_hist[i].load();
  if (reset) {
    g_rollouts = 0;
    g_iterations = 0;
    for (int i = 0; i < 32; ++i) g_trial_hist[i] = 0;
  }
}
