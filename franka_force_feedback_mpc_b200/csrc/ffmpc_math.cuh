// ffmpc_math.cuh -- small fixed-size algebra shared by the CUDA kernels: 3-vectors, 3x3 matrices, spatial
// (6-D) vectors [linear; angular], spatial inertias about the world origin, the composite "B" operator of the
// analytic RNEA derivatives, the quadratic-barrier activation, and the layout of one knot record.
//
// Everything is expressed in the WORLD frame at the world origin (Carpentier & Mansard-style recursions):
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
// Math spec: SURVEY.md App. A.1-A.6; the knot evaluation itself is ffmpc_coop.cuh.
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

// ------------------------------------------------------------------------------------------------
// Knot record: written by k_node_diff_coop (calcDiff), streamed by the Riccati kernel (ffmpc_riccati.cuh).
// The Riccati sweep runs its products on the fp64 tensor-core MMA (mma.sync m8n8k4), so the record is
// stored directly in the register layouts of that instruction, over z = (q, v, u_inner) as 3 blocks of
// 7 padded to 8 (u_inner = tau: the third STATE block of the force-feedback model, the control of the
// classical one).  With lane = 4 g + t (g = 0..7, t = 0..3):
//   "C layout" of an 8x8 tile M : lane holds M[g][2t], M[g][2t+1]          (mma accumulator)
//   "fragment"  of an 8x8 tile M : lane holds M[g][t],  M[g][4+t]           (mma A operand of M, B operand of M^T)
//   [REC_GH,  +192)  fragments of Gh_a^T, a = 0..2, Gh = dt * d a/d z (7 x 21): element (k, 7a + c) at rec_gh(k, a, c)
//   [REC_LZZ, +384)  the six upper tiles (0,0) (0,1) (0,2) (1,1) (1,2) (2,2) of the Gauss-Newton Hessian Lzz (already
//                    scaled by dt on running knots) in C layout: element (7a + r, 7b + c) at rec_lzz(a, r, b, c); diagonal
//                    tiles hold both triangles.  ROW 7 of the tiles (0, b) carries the gradient Lz[7b + c]: the sweep
//                    keeps (Vxx, Vx) as one matrix in homogeneous coordinates, the gradient riding in the padding row.
//   [REC_LW,  +8)    FF only: gradient wrt the unfiltered control w
//   [REC_LWW, +8)    FF only: diagonal of the Hessian wrt w
// Padding entries are never written: the buffer is zeroed once when the handle is created.
// ------------------------------------------------------------------------------------------------
constexpr int REC_GH = 0;
constexpr int REC_LZZ = 192;
constexpr int REC_LW = 576;
constexpr int REC_LWW = 584;
constexpr int REC_SIZE = 592;  // 4736 bytes, a multiple of 16 (one bulk copy per knot)
FFMPC_DEV int rec_tile(int a, int b) { return a == 0 ? b : (a == 1 ? 2 + b : 5); }
FFMPC_DEV int rec_gh(int k, int a, int c) { return REC_GH + a * 64 + (c * 4 + (k & 3)) * 2 + (k >> 2); }
FFMPC_DEV int rec_lzz(int a, int r, int b, int c) { return REC_LZZ + rec_tile(a, b) * 64 + (r * 4 + (c >> 1)) * 2 + (c & 1); }  // a <= b
FFMPC_DEV int rec_lz(int b, int c) { return rec_lzz(0, 7, b, c); }
// element (i, j) of Lzz / component i of Lz for plain z indices 0..20 (test hooks)
FFMPC_DEV int rec_lzz_any(int i, int j) {
  const int a = i / NJ, b = j / NJ;
  return a <= b ? rec_lzz(a, i - a * NJ, b, j - b * NJ) : rec_lzz(b, j - b * NJ, a, i - a * NJ);
}
FFMPC_DEV int rec_lz_any(int i) { return rec_lz(i / NJ, i % NJ); }

}  // namespace ffmpc
