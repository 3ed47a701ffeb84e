// Fixed-lag smoother linearisation on the device: the consumer of the factor kernels.
//
// The reference stops at the factors (perseus/smoother/factors.py) and hands them to GTSAM, which whitens each factor
// by its diagonal noise model (factors.py:42,49: noiseModel.Diagonal sigmas), builds the Gauss-Newton system and
// solves it.  For the cube-tracking chain the graph is a chain: state k = [pose_k (rot, trans); ang_vel_k; vel_k]
// (12 tangent dims), P projection factors on pose_k, one dynamics factor on (pose_k, ang_vel_k, vel_k, pose_k+1) and
// constant-velocity factors on (ang_vel_k, ang_vel_k+1), (vel_k, vel_k+1)  ->  H = sum J^T W J is block-tridiagonal
// with 12x12 blocks.
//
//   normal_equations_kernel : one CTA per (window, step), one thread per entry of the diagonal block D_k / the
//                             sub-diagonal block O_k = H[k+1][k]; reads the factor kernels' outputs in place
//                             (column-major Jacobians exactly as pb200_factors_* write them), no atomics
//   chain_solve_kernel      : one warp per window: block-tridiagonal Cholesky (L_k, M_k = O_k L_k^-T), forward and
//                             backward substitution, fp64, everything staged in shared memory
#include "common.cuh"
#include "factors_math.h"

namespace pb200 {

constexpr int kS = 12;        // tangent dims per step
constexpr int kSS = kS * kS;  // 144

struct NormalEqArgs {
  const double *proj_r, *proj_J;                 // [W][T][P][2], [W][T][P][12] (2x6 column-major)
  const double *dyn_r, *H0, *H1, *H2, *H3;       // [W][T-1][6], [..][36], [..][18], [..][18], [..][36] (column-major)
  const double *cvw_r, *cvv_r;                   // [W][T-1][3] ang_vel / vel constant-velocity residuals (x_k+1 - x_k)
  double w_px, w_dyn[6], w_cvw[3], w_cvv[3];     // 1 / sigma^2
  double lambda, prior;                          // Levenberg damping on every diagonal; extra precision on step 0
  double *D, *O, *g;                             // [W][T][144] row-major, [W][T-1][144] row-major, [W][T][12]
  int W, T, P;
};

// Jacobian of the dynamics factor k w.r.t. state k: 6 x 12 = [H0 | H1 | H2]; w.r.t. state k+1: [H3 | 0 | 0]
__device__ __forceinline__ double dynJa(const NormalEqArgs& a, long long f, int r, int c) {
  if (c < 6) return a.H0[f * 36 + c * 6 + r];
  if (c < 9) return a.H1[f * 18 + (c - 6) * 6 + r];
  return a.H2[f * 18 + (c - 9) * 6 + r];
}

__global__ void __launch_bounds__(kSS) normal_equations_kernel(NormalEqArgs a) {
  const int k = blockIdx.x, w = blockIdx.y;
  const int i = threadIdx.x / kS, j = threadIdx.x % kS;
  const long long s = static_cast<long long>(w) * a.T + k;        // state index
  const long long f = static_cast<long long>(w) * (a.T - 1) + k;  // factor (k, k+1); f - 1 = factor (k-1, k)
  double d = 0.0, o = 0.0, gi = 0.0;
  // ---- projection factors on pose_k
  if (i < 6 && j < 6) {
    for (int p = 0; p < a.P; ++p) {
      const double* J = a.proj_J + (s * a.P + p) * 12;
      d += a.w_px * (J[i * 2] * J[j * 2] + J[i * 2 + 1] * J[j * 2 + 1]);
      if (j == 0) {
        const double* r = a.proj_r + (s * a.P + p) * 2;
        gi += a.w_px * (J[i * 2] * r[0] + J[i * 2 + 1] * r[1]);
      }
    }
  }
  // ---- factors to the next state: dynamics (k, k+1) and the two constant-velocity factors
  if (k + 1 < a.T) {
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      const double ji = dynJa(a, f, r, i), wr = a.w_dyn[r];
      d += wr * ji * dynJa(a, f, r, j);
      if (j == 0) gi += wr * ji * a.dyn_r[f * 6 + r];
      // O_k[i][j] = sum_r Jb[r][i] w Ja[r][j], Jb = [H3 | 0 | 0]
      if (i < 6) o += wr * a.H3[f * 36 + i * 6 + r] * dynJa(a, f, r, j);
    }
    if (i >= 6 && i == j) {  // const-vel: J_k = -I, J_k+1 = +I
      const double wc = i < 9 ? a.w_cvw[i - 6] : a.w_cvv[i - 9];
      d += wc;
      o -= wc;
      // (j == i here, so the gradient is added by the diagonal thread's j == 0 peer below)
    }
    if (i >= 6 && j == 0) {
      const double wc = i < 9 ? a.w_cvw[i - 6] : a.w_cvv[i - 9];
      const double r = i < 9 ? a.cvw_r[f * 3 + i - 6] : a.cvv_r[f * 3 + i - 9];
      gi -= wc * r;
    }
  }
  // ---- factors from the previous state: dynamics (k-1, k) w.r.t. pose_k, const-vel (+I)
  if (k > 0) {
    if (i < 6 && j < 6) {
#pragma unroll
      for (int r = 0; r < 6; ++r) {
        const double hi = a.H3[(f - 1) * 36 + i * 6 + r], wr = a.w_dyn[r];
        d += wr * hi * a.H3[(f - 1) * 36 + j * 6 + r];
        if (j == 0) gi += wr * hi * a.dyn_r[(f - 1) * 6 + r];
      }
    }
    if (i >= 6 && i == j) d += i < 9 ? a.w_cvw[i - 6] : a.w_cvv[i - 9];
    if (i >= 6 && j == 0) {
      const double wc = i < 9 ? a.w_cvw[i - 6] : a.w_cvv[i - 9];
      const double r = i < 9 ? a.cvw_r[(f - 1) * 3 + i - 6] : a.cvv_r[(f - 1) * 3 + i - 9];
      gi += wc * r;
    }
  }
  if (i == j) d += a.lambda + (k == 0 ? a.prior : 0.0);
  a.D[s * kSS + threadIdx.x] = d;
  if (k + 1 < a.T) a.O[f * kSS + threadIdx.x] = o;
  if (j == 0) a.g[s * kS + i] = gi;
}

// ---- block-tridiagonal Cholesky, one warp per window ---------------------------------------------------------------
// work: [W][T][2][144] doubles (L_k lower-triangular row-major, M_k = O_k L_k^-T) + the intermediate y in `delta`.
__global__ void __launch_bounds__(32)
chain_solve_kernel(const double* __restrict__ D, const double* __restrict__ O, const double* __restrict__ g,
                   double* __restrict__ work, double* __restrict__ delta, int* __restrict__ status, int T) {
  __shared__ double A[kSS], L[kSS], M[kSS], y[kS], t[kS];
  const int w = blockIdx.x, lane = threadIdx.x;
  const double* Dw = D + static_cast<long long>(w) * T * kSS;
  const double* Ow = O + static_cast<long long>(w) * (T - 1) * kSS;
  const double* gw = g + static_cast<long long>(w) * T * kS;
  double* Ww = work + static_cast<long long>(w) * T * 2 * kSS;
  double* xw = delta + static_cast<long long>(w) * T * kS;
  bool ok = true;
  for (int k = 0; k < T; ++k) {
    // A = D_k - M_{k-1} M_{k-1}^T
    for (int e = lane; e < kSS; e += 32) {
      double v = Dw[k * kSS + e];
      if (k > 0) {
        const int i = e / kS, j = e % kS;
#pragma unroll
        for (int p = 0; p < kS; ++p) v -= M[i * kS + p] * M[j * kS + p];
      }
      A[e] = v;
    }
    // rhs: t = -g_k - M_{k-1} y_{k-1}
    if (lane < kS) {
      double v = -gw[k * kS + lane];
      if (k > 0) {
#pragma unroll
        for (int p = 0; p < kS; ++p) v -= M[lane * kS + p] * y[p];
      }
      t[lane] = v;
    }
    __syncwarp();
    // L = chol(A) (lower), column by column; lane i owns row i
    for (int e = lane; e < kSS; e += 32) L[e] = 0.0;
    __syncwarp();
    for (int j = 0; j < kS; ++j) {
      double diag = A[j * kS + j];
      for (int p = 0; p < j; ++p) diag -= L[j * kS + p] * L[j * kS + p];
      if (!(diag > 0.0)) {
        ok = false;
        diag = 1.0;
      }
      const double ljj = sqrt(diag);
      if (lane == j) L[j * kS + j] = ljj;
      if (lane > j && lane < kS) {
        double v = A[lane * kS + j];
        for (int p = 0; p < j; ++p) v -= L[lane * kS + p] * L[j * kS + p];
        L[lane * kS + j] = v / ljj;
      }
      __syncwarp();
    }
    // y_k = L^-1 t (lane 0, 78 MACs) while lanes 1..12 solve M_k L^T = O_k row by row
    if (lane == 0) {
      for (int i = 0; i < kS; ++i) {
        double v = t[i];
        for (int p = 0; p < i; ++p) v -= L[i * kS + p] * y[p];
        y[i] = v / L[i * kS + i];
      }
    }
    __syncwarp();
    if (k + 1 < T && lane < kS) {
      const double* Or = Ow + k * kSS + lane * kS;
      double m[kS];
#pragma unroll
      for (int c = 0; c < kS; ++c) {
        double v = Or[c];
#pragma unroll
        for (int p = 0; p < kS; ++p)
          if (p < c) v -= m[p] * L[c * kS + p];
        m[c] = v / L[c * kS + c];
      }
#pragma unroll
      for (int c = 0; c < kS; ++c) M[lane * kS + c] = m[c];
    }
    __syncwarp();
    for (int e = lane; e < kSS; e += 32) {
      Ww[(k * 2) * kSS + e] = L[e];
      if (k + 1 < T) Ww[(k * 2 + 1) * kSS + e] = M[e];
    }
    if (lane < kS) xw[k * kS + lane] = y[lane];
    __syncwarp();
  }
  // backward: x_k = L_k^-T (y_k - M_k^T x_{k+1})
  for (int k = T - 1; k >= 0; --k) {
    for (int e = lane; e < kSS; e += 32) {
      L[e] = Ww[(k * 2) * kSS + e];
      if (k + 1 < T) M[e] = Ww[(k * 2 + 1) * kSS + e];
    }
    if (lane < kS) t[lane] = xw[k * kS + lane];
    __syncwarp();
    if (lane < kS && k + 1 < T) {
      double v = t[lane];
#pragma unroll
      for (int p = 0; p < kS; ++p) v -= M[p * kS + lane] * y[p];  // y holds x_{k+1}
      t[lane] = v;
    }
    __syncwarp();
    if (lane == 0) {
      for (int i = kS - 1; i >= 0; --i) {
        double v = t[i];
        for (int p = i + 1; p < kS; ++p) v -= L[p * kS + i] * y[p];
        y[i] = v / L[i * kS + i];
      }
    }
    __syncwarp();
    if (lane < kS) xw[k * kS + lane] = y[lane];
    __syncwarp();
  }
  if (lane == 0 && status != nullptr) status[w] = ok ? 0 : 1;
}

// ---- retraction: x <- x (+) delta for every state (gtsam Values::retract: Pose3::retract = pose * Expmap(xi), vectors add)
__global__ void retract_kernel(const double* __restrict__ poses, const double* __restrict__ ang_vel,
                               const double* __restrict__ vel, const double* __restrict__ delta, double* __restrict__ poses_out,
                               double* __restrict__ ang_vel_out, double* __restrict__ vel_out, long long n) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* d = delta + i * kS;
  const fm::Pose x = fm::load_pose12(poses + i * 12);
  const fm::Pose y = fm::compose(x, fm::pose_exp(fm::v3(d[0], d[1], d[2]), fm::v3(d[3], d[4], d[5])));
#pragma unroll
  for (int k = 0; k < 9; ++k) poses_out[i * 12 + k] = y.R.m[k];
  poses_out[i * 12 + 9] = y.t.x; poses_out[i * 12 + 10] = y.t.y; poses_out[i * 12 + 11] = y.t.z;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    ang_vel_out[i * 3 + k] = ang_vel[i * 3 + k] + d[6 + k];
    vel_out[i * 3 + k] = vel[i * 3 + k] + d[9 + k];
  }
}

// ---- fixed-lag window advance: drop the oldest state, append the prediction of the newest, insert its measurement
// One CTA per window.  New state = the dynamics factor's own prediction (factors.py:100-105): pose_T-1 * Exp(dt [w; v_b]),
// v_b = R^T v for a world-frame velocity; velocities carried over (the constant-velocity prior).  The measurement row is
// the detector's pixel keypoints (float32 [W][P][2], what head_kernel writes) widened to float64 in place.
__global__ void __launch_bounds__(128) window_push_kernel(double* __restrict__ poses, double* __restrict__ ang_vel,
                                                          double* __restrict__ vel, double* __restrict__ meas,
                                                          const float* __restrict__ keypoints_px, double dt, int world_frame,
                                                          int T, int P) {
  extern __shared__ double sm[];
  const int w = blockIdx.x;
  const int per = 12 + 3 + 3 + 2 * P;  // doubles per state
  double* pw = poses + static_cast<long long>(w) * T * 12;
  double* aw = ang_vel + static_cast<long long>(w) * T * 3;
  double* vw = vel + static_cast<long long>(w) * T * 3;
  double* mw = meas + static_cast<long long>(w) * T * 2 * P;
  // stage states 1 .. T-1 (the survivors), then write them back one slot earlier
  for (int i = threadIdx.x; i < (T - 1) * per; i += blockDim.x) {
    const int k = i / per + 1, j = i % per;
    sm[i] = j < 12 ? pw[k * 12 + j] : (j < 15 ? aw[k * 3 + j - 12] : (j < 18 ? vw[k * 3 + j - 15] : mw[k * 2 * P + j - 18]));
  }
  __syncthreads();
  for (int i = threadIdx.x; i < (T - 1) * per; i += blockDim.x) {
    const int k = i / per, j = i % per;
    if (j < 12) pw[k * 12 + j] = sm[i];
    else if (j < 15) aw[k * 3 + j - 12] = sm[i];
    else if (j < 18) vw[k * 3 + j - 15] = sm[i];
    else mw[k * 2 * P + j - 18] = sm[i];
  }
  // the new newest state, predicted from the old newest (the last staged row; for T == 1 the state itself)
  if (threadIdx.x == 0) {
    const double* last = T > 1 ? sm + (T - 2) * per : nullptr;
    double p12[12], wv[3], vv[3];
    for (int j = 0; j < 12; ++j) p12[j] = last ? last[j] : pw[j];
    for (int j = 0; j < 3; ++j) {
      wv[j] = last ? last[12 + j] : aw[j];
      vv[j] = last ? last[15 + j] : vw[j];
    }
    const fm::Pose x = fm::load_pose12(p12);
    const fm::V3 om = fm::v3(wv[0], wv[1], wv[2]);
    fm::V3 vb = fm::v3(vv[0], vv[1], vv[2]);
    if (world_frame) vb = fm::mulT(x.R, vb);
    const fm::Pose y = fm::compose(x, fm::pose_exp(fm::scale(dt, om), fm::scale(dt, vb)));
    double* po = pw + (T - 1) * 12;
    for (int k = 0; k < 9; ++k) po[k] = y.R.m[k];
    po[9] = y.t.x; po[10] = y.t.y; po[11] = y.t.z;
    for (int j = 0; j < 3; ++j) {
      aw[(T - 1) * 3 + j] = wv[j];
      vw[(T - 1) * 3 + j] = vv[j];
    }
  }
  for (int i = threadIdx.x; i < 2 * P; i += blockDim.x)
    mw[(T - 1) * 2 * P + i] = static_cast<double>(keypoints_px[static_cast<long long>(w) * 2 * P + i]);
}

}  // namespace pb200

using namespace pb200;

extern "C" int pb200_smoother_retract(const double* poses, const double* ang_vel, const double* vel, const double* delta,
                                      double* poses_out, double* ang_vel_out, double* vel_out, int64_t n,
                                      pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(n >= 0, "smoother_retract: n < 0");
  if (n == 0) return PB200_OK;
  PB_CHECK(poses && ang_vel && vel && delta && poses_out && ang_vel_out && vel_out, "smoother_retract: null argument");
  retract_kernel<<<static_cast<unsigned>((n + 127) / 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      poses, ang_vel, vel, delta, poses_out, ang_vel_out, vel_out, n);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_smoother_push(double* poses, double* ang_vel, double* vel, double* meas, const float* keypoints_px,
                                   double dt, int world_frame, int W, int T, int P, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(W >= 0 && T >= 1 && P >= 1, "smoother_push: bad W/T/P");
  if (W == 0) return PB200_OK;
  PB_CHECK(poses && ang_vel && vel && meas && keypoints_px, "smoother_push: null argument");
  const size_t smem = sizeof(double) * static_cast<size_t>(T > 1 ? T - 1 : 1) * (18 + 2 * P);
  PB_CHECK(smem <= 48 * 1024, "smoother_push: window of %d states x %d points does not fit the staging buffer", T, P);
  window_push_kernel<<<W, 128, smem, static_cast<cudaStream_t>(stream)>>>(poses, ang_vel, vel, meas, keypoints_px, dt,
                                                                         world_frame, T, P);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_smoother_normal_equations(const double* proj_resid, const double* proj_jac, const double* dyn_resid,
                                               const double* H0, const double* H1, const double* H2, const double* H3,
                                               const double* cv_angvel_resid, const double* cv_vel_resid,
                                               const double* sigmas13 /* host */, double lambda, double prior,
                                               double* D, double* O, double* g, int W, int T, int P,
                                               pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(proj_resid && proj_jac && sigmas13 && D && g && W >= 0 && T >= 1 && P >= 0, "smoother: bad argument");
  PB_CHECK(T == 1 || (dyn_resid && H0 && H1 && H2 && H3 && cv_angvel_resid && cv_vel_resid && O),
           "smoother: T > 1 needs the dynamics / constant-velocity factor outputs");
  PB_CHECK(W <= 65535, "smoother: at most 65535 windows per call");
  if (W == 0) return PB200_OK;
  NormalEqArgs a;
  a.proj_r = proj_resid; a.proj_J = proj_jac; a.dyn_r = dyn_resid; a.H0 = H0; a.H1 = H1; a.H2 = H2; a.H3 = H3;
  a.cvw_r = cv_angvel_resid; a.cvv_r = cv_vel_resid;
  for (int i = 0; i < 13; ++i) PB_CHECK(sigmas13[i] > 0.0, "smoother: sigma[%d] must be positive", i);
  a.w_px = 1.0 / (sigmas13[0] * sigmas13[0]);
  for (int i = 0; i < 6; ++i) a.w_dyn[i] = 1.0 / (sigmas13[1 + i] * sigmas13[1 + i]);
  for (int i = 0; i < 3; ++i) a.w_cvw[i] = 1.0 / (sigmas13[7 + i] * sigmas13[7 + i]);
  for (int i = 0; i < 3; ++i) a.w_cvv[i] = 1.0 / (sigmas13[10 + i] * sigmas13[10 + i]);
  a.lambda = lambda; a.prior = prior; a.D = D; a.O = O; a.g = g; a.W = W; a.T = T; a.P = P;
  normal_equations_kernel<<<dim3(T, W), kSS, 0, static_cast<cudaStream_t>(stream)>>>(a);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_smoother_solve(const double* D, const double* O, const double* g, double* work, double* delta,
                                    int32_t* status_or_null, int W, int T, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(D && g && work && delta && W >= 0 && T >= 1 && (T == 1 || O), "smoother_solve: bad argument");
  if (W == 0) return PB200_OK;
  chain_solve_kernel<<<W, 32, 0, static_cast<cudaStream_t>(stream)>>>(D, O, g, work, delta, status_or_null, T);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}
