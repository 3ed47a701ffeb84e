// SE(3) / pinhole arithmetic of the smoother factors in float64, written once for the device
// kernels (factors.cu).  Also compiles as plain host C++ so tests/ can exercise the exact same
// arithmetic on a box without a GPU (tests/host_shim.cpp) — the product path is the CUDA one.
//
// Follows perseus/smoother/factors.py:79-142 (dynamics), :160-171 (constant velocity),
// :245-275 (projection) and the GTSAM 4.2 conventions they rely on:
// tangent order [rot; trans], Jacobians w.r.t. right perturbations.
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define PB_HD __host__ __device__ __forceinline__
#else
#define PB_HD inline
#endif

namespace pb200 {
namespace fm {

struct V3 {
  double x, y, z;
};
struct M3 {  // row-major
  double m[9];
};
struct Pose {
  M3 R;
  V3 t;
};
struct M6 {  // row-major 6x6
  double m[36];
};

PB_HD V3 v3(double x, double y, double z) { return V3{x, y, z}; }
PB_HD V3 add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
PB_HD V3 sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
PB_HD V3 scale(double s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
PB_HD double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
PB_HD V3 cross(V3 a, V3 b) {
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
PB_HD M3 eye3() {
  M3 r;
  for (int i = 0; i < 9; ++i) r.m[i] = 0.0;
  r.m[0] = r.m[4] = r.m[8] = 1.0;
  return r;
}
PB_HD M3 skew(V3 w) {
  M3 r;
  r.m[0] = 0.0; r.m[1] = -w.z; r.m[2] = w.y;
  r.m[3] = w.z; r.m[4] = 0.0; r.m[5] = -w.x;
  r.m[6] = -w.y; r.m[7] = w.x; r.m[8] = 0.0;
  return r;
}
PB_HD M3 mul(const M3& a, const M3& b) {
  M3 r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      r.m[i * 3 + j] = a.m[i * 3] * b.m[j] + a.m[i * 3 + 1] * b.m[3 + j] + a.m[i * 3 + 2] * b.m[6 + j];
  return r;
}
PB_HD M3 transpose(const M3& a) {
  M3 r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r.m[i * 3 + j] = a.m[j * 3 + i];
  return r;
}
PB_HD V3 mul(const M3& a, V3 v) {
  return v3(a.m[0] * v.x + a.m[1] * v.y + a.m[2] * v.z, a.m[3] * v.x + a.m[4] * v.y + a.m[5] * v.z,
            a.m[6] * v.x + a.m[7] * v.y + a.m[8] * v.z);
}
PB_HD V3 mulT(const M3& a, V3 v) {  // a^T v
  return v3(a.m[0] * v.x + a.m[3] * v.y + a.m[6] * v.z, a.m[1] * v.x + a.m[4] * v.y + a.m[7] * v.z,
            a.m[2] * v.x + a.m[5] * v.y + a.m[8] * v.z);
}
// r = a + sa*A + sb*B (3x3)
PB_HD M3 lin3(const M3& a, double sa, const M3& A, double sb, const M3& B) {
  M3 r;
  for (int i = 0; i < 9; ++i) r.m[i] = a.m[i] + sa * A.m[i] + sb * B.m[i];
  return r;
}

#define PB_EPS_THETA2 1e-10

// Rot3::Expmap
PB_HD M3 so3_exp(V3 w) {
  const double th2 = dot(w, w);
  double a, b;
  if (th2 < PB_EPS_THETA2) {
    a = 1.0 - th2 / 6.0;
    b = 0.5 - th2 / 24.0;
  } else {
    const double th = sqrt(th2);
    a = sin(th) / th;
    b = (1.0 - cos(th)) / th2;
  }
  const M3 W = skew(w);
  return lin3(eye3(), a, W, b, mul(W, W));
}

// Rot3::Logmap, GTSAM 4.2 SO3::Logmap branch structure
PB_HD V3 so3_log(const M3& R) {
  const double R11 = R.m[0], R12 = R.m[1], R13 = R.m[2];
  const double R21 = R.m[3], R22 = R.m[4], R23 = R.m[5];
  const double R31 = R.m[6], R32 = R.m[7], R33 = R.m[8];
  const double tr = R11 + R22 + R33;
  if (tr + 1.0 < 1e-3) {
    // theta near pi: largest diagonal entry a, (a,b,c) cyclic
    double W, Q1, Q2, Q3;
    int a;
    if (R33 > R22 && R33 > R11) {
      a = 2; W = R21 - R12; Q1 = 2.0 + 2.0 * R33; Q2 = R31 + R13; Q3 = R23 + R32;
    } else if (R22 > R11) {
      a = 1; W = R13 - R31; Q1 = 2.0 + 2.0 * R22; Q2 = R23 + R32; Q3 = R12 + R21;
    } else {
      a = 0; W = R32 - R23; Q1 = 2.0 + 2.0 * R11; Q2 = R12 + R21; Q3 = R31 + R13;
    }
    const double r = sqrt(Q1);
    const double nrm = sqrt(Q1 * Q1 + Q2 * Q2 + Q3 * Q3 + W * W);
    const double sgn = W < 0 ? -1.0 : 1.0;
    const double mag = M_PI - (2.0 * sgn * W) / nrm;
    const double sc = sgn * 0.5 * mag / r;
    if (a == 2) return v3(sc * Q2, sc * Q3, sc * Q1);
    if (a == 1) return v3(sc * Q3, sc * Q1, sc * Q2);
    return v3(sc * Q1, sc * Q2, sc * Q3);
  }
  double mag;
  const double tr3 = tr - 3.0;
  if (tr3 < -1e-6) {
    double c = (tr - 1.0) / 2.0;
    c = c < -1.0 ? -1.0 : (c > 1.0 ? 1.0 : c);
    const double th = acos(c);
    mag = th / (2.0 * sin(th));
  } else {
    mag = 0.5 - tr3 / 12.0 + tr3 * tr3 / 60.0;
  }
  return v3(mag * (R32 - R23), mag * (R13 - R31), mag * (R21 - R12));
}

// SO3 right Jacobian and its inverse
PB_HD M3 so3_jr(V3 w) {
  const double th2 = dot(w, w);
  double b, c;
  if (th2 < PB_EPS_THETA2) {
    b = 0.5 - th2 / 24.0;
    c = 1.0 / 6.0 - th2 / 120.0;
  } else {
    const double th = sqrt(th2);
    b = (1.0 - cos(th)) / th2;
    c = (th - sin(th)) / (th2 * th);
  }
  const M3 W = skew(w);
  return lin3(eye3(), -b, W, c, mul(W, W));
}
PB_HD M3 so3_jr_inv(V3 w) {
  const double th2 = dot(w, w);
  double c;
  if (th2 < PB_EPS_THETA2) {
    c = 1.0 / 12.0 + th2 / 720.0;
  } else {
    const double th = sqrt(th2);
    c = 1.0 / th2 - (1.0 + cos(th)) / (2.0 * th * sin(th));
  }
  const M3 W = skew(w);
  return lin3(eye3(), 0.5, W, c, mul(W, W));
}

// Pose3::Expmap([w; v])
PB_HD Pose pose_exp(V3 w, V3 v) {
  Pose p;
  p.R = so3_exp(w);
  const double th2 = dot(w, w);
  if (th2 < PB_EPS_THETA2) {
    p.t = v;
  } else {
    const V3 wxv = cross(w, v);
    const V3 t = add(sub(wxv, mul(p.R, wxv)), scale(dot(w, v), w));
    p.t = scale(1.0 / th2, t);
  }
  return p;
}
// Pose3::Logmap -> (w, u)
PB_HD void pose_log(const Pose& p, V3* w_out, V3* u_out) {
  const V3 w = so3_log(p.R);
  const double th = sqrt(dot(w, w));
  if (th < 1e-10) {
    *w_out = w;
    *u_out = p.t;
    return;
  }
  const M3 W = skew(scale(1.0 / th, w));
  const V3 Wt = mul(W, p.t);
  const V3 WWt = mul(W, Wt);
  const double coef = 1.0 - th / (2.0 * tan(0.5 * th));
  *w_out = w;
  *u_out = add(sub(p.t, scale(0.5 * th, Wt)), scale(coef, WWt));
}
PB_HD Pose compose(const Pose& a, const Pose& b) {
  Pose r;
  r.R = mul(a.R, b.R);
  r.t = add(mul(a.R, b.t), a.t);
  return r;
}
PB_HD Pose inverse(const Pose& a) {
  Pose r;
  r.R = transpose(a.R);
  r.t = scale(-1.0, mul(r.R, a.t));
  return r;
}

// 6x6 helpers (row-major); blocks are [rot | trans]
PB_HD void m6_set_block(M6& A, int bi, int bj, const M3& B, double s) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) A.m[(bi * 3 + i) * 6 + bj * 3 + j] = s * B.m[i * 3 + j];
}
PB_HD M6 m6_zero() {
  M6 r;
  for (int i = 0; i < 36; ++i) r.m[i] = 0.0;
  return r;
}
PB_HD M6 m6_mul(const M6& a, const M6& b) {
  M6 r;
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) {
      double s = 0.0;
      for (int k = 0; k < 6; ++k) s += a.m[i * 6 + k] * b.m[k * 6 + j];
      r.m[i * 6 + j] = s;
    }
  return r;
}
// Pose3::AdjointMap = [[R, 0], [[t]x R, R]], scaled by s
PB_HD M6 adjoint(const Pose& p, double s) {
  M6 A = m6_zero();
  m6_set_block(A, 0, 0, p.R, s);
  m6_set_block(A, 1, 1, p.R, s);
  m6_set_block(A, 1, 0, mul(skew(p.t), p.R), s);
  return A;
}

// Q block of Pose3::ExpmapDerivative (Barfoot 7.86b with right-Jacobian signs)
PB_HD M3 q_matrix(V3 w, V3 v) {
  const M3 V = skew(v), W = skew(w);
  const M3 WV = mul(W, V), VW = mul(V, W);
  const M3 WVW = mul(WV, W);
  const double th2 = dot(w, w);
  double a, b, c;
  if (th2 < PB_EPS_THETA2) {
    a = 1.0 / 6.0;
    b = -1.0 / 24.0;
    c = 1.0 / 120.0;
  } else {
    const double th = sqrt(th2);
    const double s = sin(th), co = cos(th);
    const double th3 = th2 * th, th4 = th2 * th2, th5 = th4 * th;
    a = (th - s) / th3;
    b = (1.0 - 0.5 * th2 - co) / th4;
    c = -0.5 * (b - 3.0 * (th - s - th3 / 6.0) / th5);
  }
  const M3 WWV = mul(W, WV), VWW = mul(VW, W);
  const M3 WVWW = mul(WVW, W), WWVW = mul(W, WVW);
  M3 Q;
  for (int i = 0; i < 9; ++i)
    Q.m[i] = -0.5 * V.m[i] + a * (WV.m[i] + VW.m[i] - WVW.m[i]) +
             b * (WWV.m[i] + VWW.m[i] - 3.0 * WVW.m[i]) + c * (WVWW.m[i] + WWVW.m[i]);
  return Q;
}
// Pose3::ExpmapDerivative([w; v]) = [[Jr, 0], [Q, Jr]]
PB_HD M6 pose_expmap_derivative(V3 w, V3 v) {
  const M3 Jr = so3_jr(w);
  M6 J = m6_zero();
  m6_set_block(J, 0, 0, Jr, 1.0);
  m6_set_block(J, 1, 1, Jr, 1.0);
  m6_set_block(J, 1, 0, q_matrix(w, v), 1.0);
  return J;
}
// Pose3::LogmapDerivative at xi = (w, u) = Log(T): [[Jr^-1, 0], [-Jr^-1 Q Jr^-1, Jr^-1]]
PB_HD M6 pose_logmap_derivative(V3 w, V3 u) {
  const M3 Ji = so3_jr_inv(w);
  const M3 Q = q_matrix(w, u);
  M6 J = m6_zero();
  m6_set_block(J, 0, 0, Ji, 1.0);
  m6_set_block(J, 1, 1, Ji, 1.0);
  m6_set_block(J, 1, 0, mul(mul(Ji, Q), Ji), -1.0);
  return J;
}

PB_HD Pose load_pose12(const double* p) {
  Pose r;
  for (int i = 0; i < 9; ++i) r.R.m[i] = p[i];
  r.t = v3(p[9], p[10], p[11]);
  return r;
}

// ---- PoseDynamicsFactor.error_func (factors.py:79-142) -------------------------------------
// resid[6] = [rot; trans].  H0,H3: 6x6 ; H1,H2: 6x3 ; all COLUMN-major (GTSAM / Fortran order),
// written only when jac != 0.
// DENSE form: every Jacobian product as a full 6x6 (kept as the cross-check of dynamics_factor below; tests/host_shim.cpp)
PB_HD void dynamics_factor_dense(const Pose& x1, V3 ang_vel, V3 vel, const Pose& x2, double dt,
                                 int world_frame, int jac, double* resid, double* H0, double* H1,
                                 double* H2, double* H3) {
  const V3 vb = world_frame ? mulT(x1.R, vel) : vel;            // factors.py:100-101 / :134
  const V3 xw = scale(dt, ang_vel), xv = scale(dt, vb);         // factors.py:104
  const Pose inc = pose_exp(xw, xv);
  const Pose pred = compose(x1, inc);                           // factors.py:105
  const Pose rel = compose(inverse(pred), x2);                  // factors.py:108
  V3 ew, eu;
  pose_log(rel, &ew, &eu);                                      // factors.py:109
  resid[0] = ew.x; resid[1] = ew.y; resid[2] = ew.z;
  resid[3] = eu.x; resid[4] = eu.y; resid[5] = eu.z;
  if (!jac) return;
  const M6 dlog = pose_logmap_derivative(ew, eu);               // factors.py:112
  const M6 A = m6_mul(dlog, adjoint(inverse(rel), -1.0));       // dlog @ drel_dpred
  M6 h0 = m6_mul(A, adjoint(inverse(inc), 1.0));                // factors.py:114
  M6 D = m6_mul(A, pose_expmap_derivative(xw, xv));             // factors.py:117
  for (int i = 0; i < 36; ++i) D.m[i] *= dt;
  if (world_frame) {
    const M3 Vb = skew(vb);
    const M3 R1T = transpose(x1.R);
    for (int i = 0; i < 6; ++i) {
      const double d0 = D.m[i * 6 + 3], d1 = D.m[i * 6 + 4], d2 = D.m[i * 6 + 5];
      for (int j = 0; j < 3; ++j) {
        h0.m[i * 6 + j] += d0 * Vb.m[j] + d1 * Vb.m[3 + j] + d2 * Vb.m[6 + j];  // factors.py:122
        H2[j * 6 + i] = d0 * R1T.m[j] + d1 * R1T.m[3 + j] + d2 * R1T.m[6 + j];  // factors.py:125
      }
    }
  } else {
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 3; ++j) H2[j * 6 + i] = D.m[i * 6 + 3 + j];           // factors.py:128
  }
  for (int i = 0; i < 6; ++i) {
    for (int j = 0; j < 6; ++j) {
      H0[j * 6 + i] = h0.m[i * 6 + j];
      H3[j * 6 + i] = dlog.m[i * 6 + j];                                         // factors.py:130
    }
    for (int j = 0; j < 3; ++j) H1[j * 6 + i] = D.m[i * 6 + j];                  // factors.py:118
  }
}

// ---- the same factor with block-triangular Jacobian algebra -------------------------------------------------------
// In the tangent order [rot; trans] every 6x6 matrix of this factor -- Pose3::AdjointMap, ExpmapDerivative,
// LogmapDerivative and all their products -- has the form [[A, 0], [B, A]] (3x3 blocks).  The form is closed under
// multiplication: [[A1,0],[B1,A1]] [[A2,0],[B2,A2]] = [[A1 A2, 0], [B1 A2 + A1 B2, A1 A2]], so a product costs three 3x3
// products (81 multiply-adds) instead of 216 and a matrix is 18 doubles instead of 36.  The dense form above spends 60 %
// of its arithmetic and registers on structural zeros (the CUDA kernel ran at 168-255 registers per thread with spills);
// the sums below add the same non-zero terms in the same order, so the results agree with the dense form to the last
// bit (the zeros the dense form adds are exact), except that a structural zero may come out as +0 where the dense form
// computes -0.
// ---- lean forms of the SO(3) / SE(3) coefficient matrices (what dynamics_factor_io evaluates) -----------------------
// The product forms above (so3_jr, so3_jr_inv, q_matrix: GTSAM's own expressions, kept for the dense cross-check) build
// every term as a 3x3 product of skew matrices: q_matrix alone holds eight products = 72 doubles live at once, twice per
// factor, and each function takes sin / cos of the same angle again.  With W = [w]x, V = [v]x, d = w.v, n = w x v:
//   W W = w w^T - |w|^2 I        W V = v w^T - d I        V W = w v^T - d I        W V W = -d W
//   W W V = n w^T - d W          V W W = -w n^T - d W     W V W W = W W V W = -d W W
// so every matrix is a short sum of outer products (same values up to rounding; tests/test_host_factor_math.py).
struct Trig {  // angle of a rotation vector, evaluated once per twist
  double th2, th, s, c;
  int small;  // |w|^2 < PB_EPS_THETA2: the series branches
};
PB_HD Trig trig_of(V3 w) {
  Trig t;
  t.th2 = dot(w, w);
  t.small = t.th2 < PB_EPS_THETA2;
  t.th = 0.0; t.s = 0.0; t.c = 1.0;
  if (!t.small) {
    t.th = sqrt(t.th2);
    t.s = sin(t.th);
    t.c = cos(t.th);
  }
  return t;
}
// I + sa [w]x + sb [w]x^2
PB_HD M3 rodrigues_form(V3 w, double th2, double sa, double sb) {
  M3 r;
  const double k = 1.0 - sb * th2;
  r.m[0] = k + sb * w.x * w.x;         r.m[1] = sb * w.x * w.y - sa * w.z;  r.m[2] = sb * w.x * w.z + sa * w.y;
  r.m[3] = sb * w.y * w.x + sa * w.z;  r.m[4] = k + sb * w.y * w.y;         r.m[5] = sb * w.y * w.z - sa * w.x;
  r.m[6] = sb * w.z * w.x - sa * w.y;  r.m[7] = sb * w.z * w.y + sa * w.x;  r.m[8] = k + sb * w.z * w.z;
  return r;
}
PB_HD M3 so3_exp_t(V3 w, const Trig& t) {  // Rot3::Expmap
  const double a = t.small ? 1.0 - t.th2 / 6.0 : t.s / t.th;
  const double b = t.small ? 0.5 - t.th2 / 24.0 : (1.0 - t.c) / t.th2;
  return rodrigues_form(w, t.th2, a, b);
}
PB_HD M3 so3_jr_t(V3 w, const Trig& t) {
  const double b = t.small ? 0.5 - t.th2 / 24.0 : (1.0 - t.c) / t.th2;
  const double c = t.small ? 1.0 / 6.0 - t.th2 / 120.0 : (t.th - t.s) / (t.th2 * t.th);
  return rodrigues_form(w, t.th2, -b, c);
}
PB_HD M3 so3_jr_inv_t(V3 w, const Trig& t) {
  const double c = t.small ? 1.0 / 12.0 + t.th2 / 720.0 : 1.0 / t.th2 - (1.0 + t.c) / (2.0 * t.th * t.s);
  return rodrigues_form(w, t.th2, 0.5, c);
}
PB_HD Pose pose_exp_t(V3 w, V3 v, const Trig& t) {  // Pose3::Expmap([w; v])
  Pose p;
  p.R = so3_exp_t(w, t);
  if (t.small) {
    p.t = v;
  } else {
    const V3 wxv = cross(w, v);
    p.t = scale(1.0 / t.th2, add(sub(wxv, mul(p.R, wxv)), scale(dot(w, v), w)));
  }
  return p;
}
// Q block of Pose3::ExpmapDerivative (q_matrix) from outer products
PB_HD M3 q_matrix_t(V3 w, V3 v, const Trig& t) {
  double a, b, c;
  if (t.small) {
    a = 1.0 / 6.0;
    b = -1.0 / 24.0;
    c = 1.0 / 120.0;
  } else {
    const double th3 = t.th2 * t.th, th4 = t.th2 * t.th2, th5 = th4 * t.th;
    a = (t.th - t.s) / th3;
    b = (1.0 - 0.5 * t.th2 - t.c) / th4;
    c = -0.5 * (b - 3.0 * (t.th - t.s - th3 / 6.0) / th5);
  }
  const double d = dot(w, v);
  const V3 n = cross(w, v);
  // Q = -V/2 + a (v w^T + w v^T - 2 d I + d W) + b (n w^T - w n^T + d W) - 2 c d (w w^T - |w|^2 I)
  const double e = -2.0 * c * d;
  const double diag = -2.0 * a * d - e * t.th2;
  const double sw = (a + b) * d;  // coefficient of W
  const double wv[3] = {w.x, w.y, w.z}, vv[3] = {v.x, v.y, v.z}, nv[3] = {n.x, n.y, n.z};
  M3 Q;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      Q.m[i * 3 + j] = a * (vv[i] * wv[j] + wv[i] * vv[j]) + b * (nv[i] * wv[j] - wv[i] * nv[j]) + e * wv[i] * wv[j] +
                       (i == j ? diag : 0.0);
  // skew parts: sw * W - 0.5 * V
  const double kx = sw * w.x - 0.5 * v.x, ky = sw * w.y - 0.5 * v.y, kz = sw * w.z - 0.5 * v.z;
  Q.m[1] -= kz; Q.m[2] += ky;
  Q.m[3] += kz; Q.m[5] -= kx;
  Q.m[6] -= ky; Q.m[7] += kx;
  return Q;
}

struct BT {  // [[A, 0], [B, A]]
  M3 A, B;
};
PB_HD BT bt_mul(const BT& a, const BT& b) {
  BT r;
  r.A = mul(a.A, b.A);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0.0;
      for (int k = 0; k < 3; ++k) s += a.B.m[i * 3 + k] * b.A.m[k * 3 + j];
      for (int k = 0; k < 3; ++k) s += a.A.m[i * 3 + k] * b.B.m[k * 3 + j];
      r.B.m[i * 3 + j] = s;
    }
  return r;
}
PB_HD M3 scale3(double s, const M3& a) {
  M3 r;
  for (int i = 0; i < 9; ++i) r.m[i] = s * a.m[i];
  return r;
}
PB_HD BT bt_adjoint(const Pose& p, double s) {  // Pose3::AdjointMap = [[R, 0], [[t]x R, R]], scaled by s
  BT r;
  r.A = scale3(s, p.R);
  r.B = scale3(s, mul(skew(p.t), p.R));
  return r;
}
// a 6x6 [[A,0],[B,A]] written COLUMN-major (GTSAM / Fortran order)
PB_HD void bt_store_colmajor(const BT& a, double* H) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      H[j * 6 + i] = a.A.m[i * 3 + j];
      H[j * 6 + 3 + i] = a.B.m[i * 3 + j];
      H[(3 + j) * 6 + i] = 0.0;
      H[(3 + j) * 6 + 3 + i] = a.A.m[i * 3 + j];
    }
}

// The factor talks to its caller through `io` so that a CUDA thread can keep next to nothing in registers across the
// expensive middle of the computation (ncu, r02: with plain arguments ~48 doubles -- x1, the increment, the relative pose,
// the twists -- stayed live across the logarithm's Jacobian, the kernel spilled 500 bytes per thread at 128 registers and
// waited on local memory 58 % of the time):
//   inputs   io.x1(), io.x2() (Pose), io.w(), io.v() (V3): may be re-read at any time (the kernels re-read shared memory);
//   stash    io.put_pose(slot, Pose) / io.get_pose(slot), slots 0..1: parked values (shared memory on the GPU);
//   outputs  io.resid(double[6]), io.H3, io.H1, io.H2, io.H0 (column-major blocks), each handed over AS SOON AS it is
//            complete, in that order, so no block waits in registers while another is computed.
template <class IO>
PB_HD void dynamics_factor_io(IO& io, double dt, int world_frame, int jac) {
  V3 ew, eu;
  {
    const Pose x1 = io.x1();
    const V3 vb = world_frame ? mulT(x1.R, io.v()) : io.v();     // factors.py:100-101 / :134
    const V3 xw = scale(dt, io.w()), xv = scale(dt, vb);         // factors.py:104
    const Pose inc = pose_exp_t(xw, xv, trig_of(xw));
    io.put_pose(0, inc);
    const Pose pred = compose(x1, inc);                          // factors.py:105
    const Pose rel = compose(inverse(pred), io.x2());            // factors.py:108
    io.put_pose(1, rel);
    pose_log(rel, &ew, &eu);                                     // factors.py:109
    const double r[6] = {ew.x, ew.y, ew.z, eu.x, eu.y, eu.z};
    io.resid(r);
  }
  if (!jac) return;
  double blk[36];
  BT A;
  {
    BT dlog;                                                     // Pose3::LogmapDerivative, factors.py:112
    const Trig te = trig_of(ew);
    dlog.A = so3_jr_inv_t(ew, te);
    dlog.B = scale3(-1.0, mul(mul(dlog.A, q_matrix_t(ew, eu, te)), dlog.A));
    bt_store_colmajor(dlog, blk);
    io.H3(blk);                                                  // factors.py:130
    A = bt_mul(dlog, bt_adjoint(inverse(io.get_pose(1)), -1.0)); // dlog @ drel_dpred
  }
  // the twists again (15 multiply-adds from re-read inputs instead of 9 doubles held across the block above)
  const M3 R1 = io.x1().R;
  const V3 vb = world_frame ? mulT(R1, io.v()) : io.v();
  M3 DA;  // D[:, 3:6] = [[0], [D.A]]: the velocity columns touch only the translation rows
  {
    const V3 xw = scale(dt, io.w()), xv = scale(dt, vb);
    BT dexp;                                                     // Pose3::ExpmapDerivative
    const Trig tx = trig_of(xw);
    dexp.A = so3_jr_t(xw, tx);
    dexp.B = q_matrix_t(xw, xv, tx);
    BT D = bt_mul(A, dexp);                                      // factors.py:117
    D.A = scale3(dt, D.A);
    D.B = scale3(dt, D.B);
    double h1[18];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        h1[j * 6 + i] = D.A.m[i * 3 + j];                                              // factors.py:118
        h1[j * 6 + 3 + i] = D.B.m[i * 3 + j];
      }
    io.H1(h1);
    DA = D.A;
  }
  BT h0 = bt_mul(A, bt_adjoint(inverse(io.get_pose(0)), 1.0));   // factors.py:114
  {
    double h2[18];
    if (world_frame) {
      const M3 Vb = skew(vb);
      const M3 R1T = transpose(R1);
      for (int i = 0; i < 3; ++i) {
        const double d0 = DA.m[i * 3], d1 = DA.m[i * 3 + 1], d2 = DA.m[i * 3 + 2];
        for (int j = 0; j < 3; ++j) {
          h0.B.m[i * 3 + j] += d0 * Vb.m[j] + d1 * Vb.m[3 + j] + d2 * Vb.m[6 + j];      // factors.py:122
          h2[j * 6 + i] = 0.0;
          h2[j * 6 + 3 + i] = d0 * R1T.m[j] + d1 * R1T.m[3 + j] + d2 * R1T.m[6 + j];    // factors.py:125
        }
      }
    } else {
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
          h2[j * 6 + i] = 0.0;
          h2[j * 6 + 3 + i] = DA.m[i * 3 + j];                                          // factors.py:128
        }
    }
    io.H2(h2);
  }
  bt_store_colmajor(h0, blk);
  io.H0(blk);
}

struct DynArrayIO {  // plain values and arrays (host callers, tests)
  Pose p1, p2;
  V3 av, lv;
  double *r, *h0, *h1, *h2, *h3;
  Pose stash[2];
  PB_HD Pose x1() const { return p1; }
  PB_HD Pose x2() const { return p2; }
  PB_HD V3 w() const { return av; }
  PB_HD V3 v() const { return lv; }
  PB_HD void put_pose(int k, const Pose& p) { stash[k] = p; }
  PB_HD Pose get_pose(int k) const { return stash[k]; }
  PB_HD void resid(const double (&x)[6]) { for (int i = 0; i < 6; ++i) r[i] = x[i]; }
  PB_HD void H0(const double (&x)[36]) { for (int i = 0; i < 36; ++i) h0[i] = x[i]; }
  PB_HD void H1(const double (&x)[18]) { for (int i = 0; i < 18; ++i) h1[i] = x[i]; }
  PB_HD void H2(const double (&x)[18]) { for (int i = 0; i < 18; ++i) h2[i] = x[i]; }
  PB_HD void H3(const double (&x)[36]) { for (int i = 0; i < 36; ++i) h3[i] = x[i]; }
};
PB_HD void dynamics_factor(const Pose& x1, V3 ang_vel, V3 vel, const Pose& x2, double dt,
                           int world_frame, int jac, double* resid, double* H0, double* H1,
                           double* H2, double* H3) {
  DynArrayIO io{x1, x2, ang_vel, vel, resid, H0, H1, H2, H3, {}};
  dynamics_factor_io(io, dt, world_frame, jac);
}

// ---- KeypointProjectionFactor.error_func (factors.py:245-275) ------------------------------
// resid[2]; J 2x6 column-major (12 doubles); returns 1 if the point is behind the camera.
PB_HD int projection_factor(const Pose& body, V3 p_body, const double* K5, const Pose& cam,
                            double mu, double mv, int jac, double* resid, double* J) {
  const V3 pw = add(mul(body.R, p_body), body.t);               // transformFrom, factors.py:257
  const V3 q = mulT(cam.R, sub(pw, cam.t));                     // camera_pose.transformTo
  const double iz = 1.0 / q.z;
  const double x = q.x * iz, y = q.y * iz;
  const double fx = K5[0], fy = K5[1], s = K5[2], u0 = K5[3], v0 = K5[4];
  resid[0] = fx * x + s * y + u0 - mu;                          // factors.py:265
  resid[1] = fy * y + v0 - mv;
  if (jac) {
    // d pix / d q = [[fx, s], [0, fy]] * (1/z) [[1, 0, -x], [0, 1, -y]]
    const double a00 = fx * iz, a01 = s * iz, a02 = -(fx * x + s * y) * iz;
    const double a11 = fy * iz, a12 = -fy * y * iz;
    // d q / d p_w = Rc^T  ->  G = dpix/dq * Rc^T  (2x3)
    double G[6];
    for (int j = 0; j < 3; ++j) {
      const double c0 = cam.R.m[j * 3 + 0], c1 = cam.R.m[j * 3 + 1], c2 = cam.R.m[j * 3 + 2];
      G[j] = a00 * c0 + a01 * c1 + a02 * c2;
      G[3 + j] = a11 * c1 + a12 * c2;
    }
    // d p_w / d pose = [ -R [p]x , R ]  (3x6)
    const M3 RP = mul(body.R, skew(p_body));
    for (int r = 0; r < 2; ++r)
      for (int j = 0; j < 3; ++j) {
        double a = 0.0, b = 0.0;
        for (int k = 0; k < 3; ++k) {
          a -= G[r * 3 + k] * RP.m[k * 3 + j];
          b += G[r * 3 + k] * body.R.m[k * 3 + j];
        }
        J[j * 2 + r] = a;        // factors.py:264
        J[(3 + j) * 2 + r] = b;
      }
  }
  return q.z <= 0.0;
}

}  // namespace fm
}  // namespace pb200
