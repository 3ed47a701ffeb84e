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
PB_HD void dynamics_factor(const Pose& x1, V3 ang_vel, V3 vel, const Pose& x2, double dt,
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
