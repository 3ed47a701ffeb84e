// TEST-ONLY: compiles the device arithmetic headers of perseus_b200/csrc as plain host C++ so the
// CPU test-suite can check the exact expressions the CUDA kernels evaluate (against the numpy
// oracle) on a box without a GPU.  Never loaded by the product path.
#include <stdint.h>

#include "../perseus_b200/csrc/factors_math.h"

using namespace pb200::fm;

extern "C" {

void shim_dynamics(const double* pose1, const double* angvel, const double* vel, const double* pose2,
                   double dt, int world_frame, int jac, double* resid, double* H0, double* H1,
                   double* H2, double* H3, int64_t N) {
  for (int64_t i = 0; i < N; ++i) {
    const Pose x1 = load_pose12(pose1 + 12 * i), x2 = load_pose12(pose2 + 12 * i);
    const V3 w = v3(angvel[3 * i], angvel[3 * i + 1], angvel[3 * i + 2]);
    const V3 v = v3(vel[3 * i], vel[3 * i + 1], vel[3 * i + 2]);
    dynamics_factor(x1, w, v, x2, dt, world_frame, jac, resid + 6 * i, jac ? H0 + 36 * i : nullptr,
                    jac ? H1 + 18 * i : nullptr, jac ? H2 + 18 * i : nullptr,
                    jac ? H3 + 36 * i : nullptr);
  }
}

// the dense 6x6 form of the same factor (every structural zero multiplied out): cross-check of the block-triangular algebra
void shim_dynamics_dense(const double* pose1, const double* angvel, const double* vel, const double* pose2,
                         double dt, int world_frame, double* resid, double* H0, double* H1, double* H2, double* H3,
                         int64_t N) {
  for (int64_t i = 0; i < N; ++i) {
    const Pose x1 = load_pose12(pose1 + 12 * i), x2 = load_pose12(pose2 + 12 * i);
    const V3 w = v3(angvel[3 * i], angvel[3 * i + 1], angvel[3 * i + 2]);
    const V3 v = v3(vel[3 * i], vel[3 * i + 1], vel[3 * i + 2]);
    dynamics_factor_dense(x1, w, v, x2, dt, world_frame, 1, resid + 6 * i, H0 + 36 * i, H1 + 18 * i, H2 + 18 * i, H3 + 36 * i);
  }
}

void shim_projection(const double* poses, const double* meas, const double* pts, const double* K5,
                     const double* cam12, int jac, double* resid, double* J, uint8_t* cheir,
                     int64_t N, int P) {
  Pose cam;
  if (cam12) {
    cam = load_pose12(cam12);
  } else {
    cam.R = eye3();
    cam.t = v3(0, 0, 0);
  }
  for (int64_t n = 0; n < N; ++n) {
    const Pose body = load_pose12(poses + 12 * n);
    for (int p = 0; p < P; ++p) {
      const int64_t f = n * P + p;
      const V3 pb = v3(pts[3 * p], pts[3 * p + 1], pts[3 * p + 2]);
      cheir[f] = (uint8_t)projection_factor(body, pb, K5, cam, meas[2 * f], meas[2 * f + 1], jac,
                                            resid + 2 * f, jac ? J + 12 * f : nullptr);
    }
  }
}

void shim_so3_log(const double* R9, double* w3) {
  M3 R;
  for (int i = 0; i < 9; ++i) R.m[i] = R9[i];
  const V3 w = so3_log(R);
  w3[0] = w.x; w3[1] = w.y; w3[2] = w.z;
}
}
