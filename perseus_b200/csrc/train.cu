// Training-step operators of the keypoint detector behind the C ABI (pb200_train_*): the pieces of
// perseus/detector/train.py:259-304 (autocast forward with batch-statistics BatchNorm, SmoothL1 loss, backward,
// gradient-norm clip, AdamW) as hand-written sm_100a kernels.  Convolutions (forward, data gradient, weight
// gradient) run on tcgen05 (conv_umma.cu / wgrad_umma.cu); everything else here is memory-bound SIMT work on
// NHWC bf16 activations with fp32 statistics, gradients and optimizer state.
//
// All entries are asynchronous on the caller's stream and allocate nothing; tensors are device pointers.
#include <math.h>
#include <string.h>

#include "common.cuh"
#include "conv_umma.cuh"
#include "layout.h"
#include "wgrad_umma.cuh"

namespace pb200 {

// ---- weight packing -------------------------------------------------------------------------------------------
// mode 0: forward operand  [Cout][(r, s, ci)]          (K-major rows for the TMA-im2col kernels)
// mode 1: data-gradient operand [Cin][(R-1-r, R-1-s, co)]: conv of dz with the flipped, transposed filter
__global__ void pack_weights_kernel(const float* __restrict__ w, int Cout, int Cin, int R, int mode,
                                    __nv_bfloat16* __restrict__ out) {
  const long long total = static_cast<long long>(Cout) * Cin * R * R;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  // i indexes the OUTPUT so the stores are coalesced
  if (mode == 0) {
    const int ci = static_cast<int>(i % Cin);
    long long t = i / Cin;
    const int s = static_cast<int>(t % R);
    t /= R;
    const int r = static_cast<int>(t % R);
    const int co = static_cast<int>(t / R);
    out[i] = __float2bfloat16_rn(w[((static_cast<long long>(co) * Cin + ci) * R + r) * R + s]);
  } else {
    const int co = static_cast<int>(i % Cout);
    long long t = i / Cout;
    const int s2 = static_cast<int>(t % R);
    t /= R;
    const int r2 = static_cast<int>(t % R);
    const int ci = static_cast<int>(t / R);
    const int r = R - 1 - r2, s = R - 1 - s2;
    out[i] = __float2bfloat16_rn(w[((static_cast<long long>(co) * Cin + ci) * R + r) * R + s]);
  }
}

// fp32 scratch gradient [R*R][Cin][Cout] (wgrad_umma.cu) -> OIHW fp32 (the parameter's layout)
__global__ void unpack_wgrad_kernel(const float* __restrict__ scratch, int Cout, int Cin, int R, int cin_real,
                                    float* __restrict__ dw) {
  const long long total = static_cast<long long>(Cout) * cin_real * R * R;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int s = static_cast<int>(i % R);
  long long t = i / R;
  const int r = static_cast<int>(t % R);
  t /= R;
  const int ci = static_cast<int>(t % cin_real);
  const int co = static_cast<int>(t / cin_real);
  dw[i] = scratch[(static_cast<long long>(r * R + s) * Cin + ci) * Cout + co];
}

}  // namespace pb200

using namespace pb200;

extern "C" int pb200_train_pack_weights(const float* w_oihw, int Cout, int Cin, int R, int mode, void* out_bf16,
                                        pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(w_oihw && out_bf16 && (mode == 0 || mode == 1), "train_pack_weights: bad argument");
  const long long total = static_cast<long long>(Cout) * Cin * R * R;
  pack_weights_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w_oihw, Cout, Cin, R, mode, static_cast<__nv_bfloat16*>(out_bf16));
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_train_conv(const void* in_nhwc, int B, int H, int W, int Cin, const void* wpacked,
                                const float* bias, int Cout, int R, int stride, int pad,
                                const void* residual_or_null, int relu, void* out_nhwc, double* stats_or_null,
                                pb200_stream stream) {
  if (int rc = require_device()) return rc;
  if (int rc = umma_conv_init()) return rc;
  PB_CHECK(in_nhwc && wpacked && bias && out_nhwc, "train_conv: null argument");
  UmmaConvLaunch L;
  // same tile choice as the inference plan (conv_umma.cu: shared-memory-traffic cost model); PB200_FORCE_CLUSTER still
  // overrides the cluster shape in the tests
  int bn = 64, cluster = 0;
  {
    const int Ho = (H + 2 * pad - R) / stride + 1, Wo = (W + 2 * pad - R) / stride + 1;
    const long long m_tiles = (static_cast<long long>(B) * Ho * Wo + 127) / 128;
    umma_choose_tile(m_tiles, Cout, 1, &bn, &cluster);
    if (getenv("PB200_FORCE_CLUSTER") != nullptr) cluster = 0;
  }
  if (int rc = build_umma_conv(&L, in_nhwc, B, H, W, Cin, wpacked, bias, residual_or_null, out_nhwc, Cout, R, stride,
                               pad, relu, bn, cluster, 0, nullptr, 0, 0, 0, stats_or_null))
    return rc;
  return launch_umma_conv(L, static_cast<cudaStream_t>(stream));
}

extern "C" int pb200_train_dgrad_s2(const void* dz_nhwc, int B, int Ho, int Wo, int Cout, const void* w_dgrad_packed,
                                    const float* zero_bias, int Cin, int R, const void* residual00_or_null,
                                    void* dx_nhwc, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  if (int rc = umma_conv_init()) return rc;
  PB_CHECK(dz_nhwc && w_dgrad_packed && zero_bias && dx_nhwc, "train_dgrad_s2: null argument");
  int bn = 64;
  {
    const long long m_tiles = (static_cast<long long>(B) * Ho * Wo + 127) / 128;
    long long best_cost = -1;
    for (int c = 64; c <= 256 && c <= Cin; c *= 2) {
      if (Cin % c) continue;
      const long long tiles = m_tiles * (Cin / c);
      const long long rounds = (tiles + num_sms() - 1) / num_sms();
      const long long cost = rounds * (c / 2 < 48 ? 48 : c / 2);
      if (best_cost < 0 || cost <= best_cost) {
        bn = c;
        best_cost = cost;
      }
    }
  }
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      UmmaConvLaunch L;
      const void* res = (ph == 0 && pw == 0) ? residual00_or_null : nullptr;
      if (int rc = build_umma_dgrad_s2(&L, dz_nhwc, B, Ho, Wo, Cout, w_dgrad_packed, zero_bias, res, dx_nhwc, Cin, R, ph, pw, bn))
        return rc;
      if (int rc = launch_umma_conv(L, static_cast<cudaStream_t>(stream))) return rc;
    }
  return PB200_OK;
}

extern "C" int pb200_train_wgrad(const void* x_nhwc, const void* dz_nhwc, int B, int H, int W, int Cin, int Cout,
                                 int R, int stride, int pad, float* scratch, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  if (int rc = wgrad_umma_init()) return rc;
  PB_CHECK(x_nhwc && dz_nhwc && scratch, "train_wgrad: null argument");
  WgradLaunch L;
  if (int rc = build_wgrad(&L, x_nhwc, dz_nhwc, scratch, B, H, W, Cin, Cout, R, stride, pad)) return rc;
  return launch_wgrad(L, static_cast<cudaStream_t>(stream));
}

extern "C" int pb200_train_unpack_wgrad(const float* scratch, int Cout, int Cin, int R, int cin_real, float* dw_oihw,
                                        pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(scratch && dw_oihw && cin_real >= 1 && cin_real <= Cin, "train_unpack_wgrad: bad argument");
  const long long total = static_cast<long long>(Cout) * cin_real * R * R;
  unpack_wgrad_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      scratch, Cout, Cin, R, cin_real, dw_oihw);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

// =================================================================================================================
// Batch-statistics BatchNorm on NHWC bf16 [M][C] (train-mode nn.BatchNorm2d, resnet.py:93-99 / 269; eps 1e-5,
// momentum 0.1).  Memory-bound: a thread owns 8 adjacent channels (one 16-byte load) of a strided set of rows.
//   forward : bn_stats (sum, sum of squares -> fp64 atomics)  +  bn_apply (scale/shift derived from the sums in
//             every CTA's prologue; y = [relu](z * scale + shift [+ residual]); CTA 0 stores mean / invstd and updates
//             the running statistics)
//   backward: bn_bwd_reduce (sum g, sum g * xhat with g = dy * [y > 0])  +  bn_bwd_apply
//             (dz = gamma * invstd * (g - mean(g) - xhat * mean(g * xhat)); also g itself for the identity branch, the
//             zero-stuffed copy a stride-2 data gradient convolves, and dgamma / dbeta)
// =================================================================================================================
namespace pb200 {

constexpr int kBnThreads = 256;

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}

// rows [row0, row1) of this CTA; thread -> (channel group cg, row lane rl)
struct RowSplit {
  long long row0, row1;
  int cg, rl, lanes;
};
__device__ __forceinline__ RowSplit row_split(long long M, int C8) {
  RowSplit s;
  const long long per = (M + gridDim.x - 1) / gridDim.x;
  s.row0 = per * blockIdx.x;
  s.row1 = s.row0 + per < M ? s.row0 + per : M;
  s.cg = threadIdx.x % C8;
  s.rl = threadIdx.x / C8;
  s.lanes = kBnThreads / C8;
  return s;
}

__global__ void __launch_bounds__(kBnThreads, 4)
bn_stats_kernel(const uint4* __restrict__ z, long long M, int C, double* __restrict__ sums /* [2][C] */) {
  __shared__ float s_acc[2 * 512];
  const int C8 = C / 8;
  for (int i = threadIdx.x; i < 2 * C; i += kBnThreads) s_acc[i] = 0.f;
  __syncthreads();
  const RowSplit rs = row_split(M, C8);
  float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 4
  for (long long r = rs.row0 + rs.rl; r < rs.row1; r += rs.lanes) {
    float f[8];
    unpack8(__ldg(z + r * C8 + rs.cg), f);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      a[e] += f[e];
      q[e] = fmaf(f[e], f[e], q[e]);
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    atomicAdd(&s_acc[rs.cg * 8 + e], a[e]);
    atomicAdd(&s_acc[C + rs.cg * 8 + e], q[e]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * C; i += kBnThreads) atomicAdd(&sums[i], static_cast<double>(s_acc[i]));
}

__global__ void __launch_bounds__(kBnThreads, 4)
bn_apply_kernel(const uint4* __restrict__ z, const double* __restrict__ sums, const float* __restrict__ gamma,
                const float* __restrict__ beta, float* __restrict__ running_mean, float* __restrict__ running_var,
                float momentum, float eps, float* __restrict__ save_mean, float* __restrict__ save_invstd,
                const uint4* __restrict__ residual, int relu, uint4* __restrict__ y, long long M, int C) {
  __shared__ float s_scale[512], s_shift[512];
  for (int c = threadIdx.x; c < C; c += kBnThreads) {
    const double mean = sums[c] / static_cast<double>(M);
    double var = sums[C + c] / static_cast<double>(M) - mean * mean;
    if (var < 0.0) var = 0.0;
    const float invstd = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
    const float sc = gamma[c] * invstd;
    s_scale[c] = sc;
    s_shift[c] = beta[c] - static_cast<float>(mean) * sc;
    if (blockIdx.x == 0) {
      save_mean[c] = static_cast<float>(mean);
      save_invstd[c] = invstd;
      if (running_mean != nullptr) {  // nn.BatchNorm2d: unbiased variance in the running estimate
        const double unb = M > 1 ? var * static_cast<double>(M) / static_cast<double>(M - 1) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * static_cast<float>(mean);
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * static_cast<float>(unb);
      }
    }
  }
  __syncthreads();
  const int C8 = C / 8;
  // C8 is a power of two <= 64 (host check) and both the CTA size and the grid stride are multiples of it, so a thread
  // keeps ONE channel group for the whole loop: its 8 scales / shifts live in registers
  // The tensor is walked from its END: z was just written front to back by the producing conv, so its tail is what L2
  // still holds -- a front-to-back pass would miss on the (evicted) head and push the tail out before reaching it.
  // Element index total-1-j: the channel group of a thread is C8-1 - (tid mod C8), still fixed.
  const int c0 = (C8 - 1 - (static_cast<int>(threadIdx.x) & (C8 - 1))) * 8;
  float sc[8], sh[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    sc[e] = s_scale[c0 + e];
    sh[e] = s_shift[c0 + e];
  }
  const long long total = M * C8;
  const long long step = static_cast<long long>(gridDim.x) * kBnThreads;
#pragma unroll 4
  for (long long j = static_cast<long long>(blockIdx.x) * kBnThreads + threadIdx.x; j < total; j += step) {
    const long long i = total - 1 - j;
    float f[8];
    unpack8(__ldg(z + i), f);
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = fmaf(f[e], sc[e], sh[e]);
    if (residual != nullptr) {
      float r[8];
      unpack8(__ldg(residual + i), r);
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] += r[e];
    }
    if (relu) {
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
    }
    y[i] = pack8(f);
  }
}

// sums[0..C) = sum g, sums[C..2C) = sum g * (z - mean)   (g = dy under the layer's ReLU mask); the consumer multiplies
// the second by invstd to get sum g * xhat
__global__ void __launch_bounds__(kBnThreads, 3)
bn_bwd_reduce_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y_mask, const uint4* __restrict__ z,
                     const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                     const float* __restrict__ beta, int mask_from_z, long long M, int C,
                     double* __restrict__ sums /* [2][C] */) {
  __shared__ float s_acc[2 * 512];
  const int C8 = C / 8;
  for (int i = threadIdx.x; i < 2 * C; i += kBnThreads) s_acc[i] = 0.f;
  __syncthreads();
  const RowSplit rs = row_split(M, C8);
  // mask_from_z: the ReLU mask of a layer without a residual input is [z*sc + sh > 0] = [(z - mu) * sc > -beta]
  // (the forward's affine map, bn_apply_kernel); thr = -beta / sc folded per sign of sc
  float mu[8], sc[8], nb[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int c = rs.cg * 8 + e;
    mu[e] = mean[c];
    sc[e] = mask_from_z ? gamma[c] * invstd[c] : 0.f;
    nb[e] = mask_from_z ? beta[c] - mu[e] * sc[e] : 0.f;
  }
  float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  // rows interleaved over the CTAs (not one contiguous range per CTA): the pass ends on the tensor's tail, which is
  // where bn_bwd_apply_kernel (walking backwards) starts and finds dy / z still in L2
  const long long rstep = static_cast<long long>(gridDim.x) * rs.lanes;
#pragma unroll 2
  for (long long r = static_cast<long long>(blockIdx.x) * rs.lanes + rs.rl; r < M; r += rstep) {
    const long long i = r * C8 + rs.cg;
    float g[8], zz[8];
    unpack8(__ldg(dy + i), g);
    unpack8(__ldg(z + i), zz);
    if (mask_from_z) {
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] = fmaf(zz[e], sc[e], nb[e]) > 0.f ? g[e] : 0.f;
    } else if (y_mask != nullptr) {
      float m[8];
      unpack8(__ldg(y_mask + i), m);
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] = m[e] > 0.f ? g[e] : 0.f;
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      a[e] += g[e];
      q[e] = fmaf(g[e], zz[e] - mu[e], q[e]);
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    atomicAdd(&s_acc[rs.cg * 8 + e], a[e]);
    atomicAdd(&s_acc[C + rs.cg * 8 + e], q[e]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * C; i += kBnThreads) atomicAdd(&sums[i], static_cast<double>(s_acc[i]));
}

// dz = k * (g - mean(g) - xhat * mean(g * xhat)), k = gamma * invstd, evaluated as k*g + A*z + B with per-channel
// constants (A = -k * invstd * mean(g*xhat), B = -k * mean(g) - A * mean) held in registers.
// stuffed_out: [B][2*Ho][2*Wo][C] with dz at the even positions (the caller zeroes it); Wo / HoWo describe dz's rows
__global__ void __launch_bounds__(kBnThreads, 3)
bn_bwd_apply_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y_mask, const uint4* __restrict__ z,
                    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                    const float* __restrict__ beta, int mask_from_z, const double* __restrict__ sums, long long M, int C,
                    uint4* __restrict__ dz,
                    uint4* __restrict__ g_out, uint4* __restrict__ stuffed_out, int Ho, int Wo,
                    float* __restrict__ dgamma, float* __restrict__ dbeta) {
  __shared__ float s_k[512], s_a[512], s_b[512], s_sh[512];
  const double invM = 1.0 / static_cast<double>(M);
  for (int c = threadIdx.x; c < C; c += kBnThreads) {
    const double is = static_cast<double>(invstd[c]), mu = static_cast<double>(mean[c]);
    const double sg = sums[c], sgx = sums[C + c] * is;  // sum g, sum g * xhat
    const double k = static_cast<double>(gamma[c]) * is;
    const double A = -k * is * (sgx * invM);
    s_k[c] = static_cast<float>(k);
    s_a[c] = static_cast<float>(A);
    s_b[c] = static_cast<float>(-k * (sg * invM) - A * mu);
    s_sh[c] = mask_from_z ? beta[c] - mean[c] * (gamma[c] * invstd[c]) : 0.f;
    if (blockIdx.x == 0) {
      dgamma[c] = static_cast<float>(sgx);
      dbeta[c] = static_cast<float>(sg);
    }
  }
  __syncthreads();
  const int C8 = C / 8;
  const int cg = C8 - 1 - (static_cast<int>(threadIdx.x) & (C8 - 1));  // fixed per thread; back-to-front walk (bn_apply_kernel)
  float k[8], A[8], Bc[8], sh[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    k[e] = s_k[cg * 8 + e];
    A[e] = s_a[cg * 8 + e];
    Bc[e] = s_b[cg * 8 + e];
    sh[e] = s_sh[cg * 8 + e];
  }
  const int c8_shift = __ffs(C8) - 1;
  const long long total = M * C8;
  const long long step = static_cast<long long>(gridDim.x) * kBnThreads;
#pragma unroll 2
  for (long long j = static_cast<long long>(blockIdx.x) * kBnThreads + threadIdx.x; j < total; j += step) {
    const long long i = total - 1 - j;
    float g[8], zz[8], o[8];
    unpack8(__ldg(dy + i), g);
    unpack8(__ldg(z + i), zz);
    if (mask_from_z) {
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] = fmaf(zz[e], k[e], sh[e]) > 0.f ? g[e] : 0.f;
    } else if (y_mask != nullptr) {
      float m[8];
      unpack8(__ldg(y_mask + i), m);
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] = m[e] > 0.f ? g[e] : 0.f;
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = fmaf(k[e], g[e], fmaf(A[e], zz[e], Bc[e]));
    const uint4 ov = pack8(o);
    dz[i] = ov;
    if (g_out != nullptr) g_out[i] = pack8(g);
    if (stuffed_out != nullptr) {
      const unsigned row = static_cast<unsigned>(i >> c8_shift);  // < 2^31 rows
      const unsigned wo = row % static_cast<unsigned>(Wo);
      const unsigned t = row / static_cast<unsigned>(Wo);
      const unsigned ho = t % static_cast<unsigned>(Ho);
      const long long b = t / static_cast<unsigned>(Ho);
      stuffed_out[((b * (2 * Ho) + 2 * ho) * (2LL * Wo) + 2 * wo) * C8 + cg] = ov;
    }
  }
}

// one resident wave (per_sm = the kernel's __launch_bounds__ occupancy), and at least 16 vectors per thread: a small
// tensor gets few CTAs, so its 2*C fp64 atomics per CTA and the per-CTA constant prologue stay negligible
static inline int bn_grid(long long M, int C, int per_sm) {
  const long long vec = M * (C / 8);
  long long g = (vec + kBnThreads * 16 - 1) / (kBnThreads * 16);
  const long long cap = static_cast<long long>(num_sms()) * per_sm;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

}  // namespace pb200

#define PB_BN_CHECK(C) PB_CHECK((C) >= 8 && (C) <= 512 && ((C) & ((C) - 1)) == 0, "bn: C = %d unsupported (power of two in [8, 512])", (C))

extern "C" int pb200_train_bn_stats(const void* z, int64_t M, int C, double* sums, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(z && sums && M >= 1, "train_bn_stats: bad argument");
  PB_BN_CHECK(C);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PB_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * 2 * C, st));
  bn_stats_kernel<<<bn_grid(M, C, 4), kBnThreads, 0, st>>>(static_cast<const uint4*>(z), M, C, sums);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_train_bn_apply(const void* z, const double* sums, const float* gamma, const float* beta,
                                    float* running_mean, float* running_var, float momentum, float eps,
                                    float* save_mean, float* save_invstd, const void* residual_or_null, int relu,
                                    void* y, int64_t M, int C, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(z && sums && gamma && beta && save_mean && save_invstd && y, "train_bn_apply: null argument");
  PB_BN_CHECK(C);
  bn_apply_kernel<<<bn_grid(M, C, 4), kBnThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(z), sums, gamma, beta, running_mean, running_var, momentum, eps, save_mean, save_invstd,
      static_cast<const uint4*>(residual_or_null), relu, static_cast<uint4*>(y), M, C);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_train_bn_backward(const void* dy, const void* y_mask_or_null, const void* z, const float* save_mean,
                                       const float* save_invstd, const float* gamma, const float* beta_or_null,
                                       double* sums, int64_t M, int C,
                                       void* dz, void* g_out_or_null, void* stuffed_out_or_null, int Ho, int Wo,
                                       float* dgamma, float* dbeta, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(dy && z && save_mean && save_invstd && gamma && sums && dz && dgamma && dbeta, "train_bn_backward: null argument");
  PB_BN_CHECK(C);
  PB_CHECK(stuffed_out_or_null == nullptr || (Ho > 0 && Wo > 0 && M % (static_cast<int64_t>(Ho) * Wo) == 0),
           "train_bn_backward: bad zero-stuffing geometry");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PB_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * 2 * C, st));
  if (stuffed_out_or_null != nullptr)
    PB_CUDA(cudaMemsetAsync(stuffed_out_or_null, 0, static_cast<size_t>(M) * 4 * C * 2, st));
  const int grid = bn_grid(M, C, 3);
  // beta given (and no y tensor): the layer had no residual input, so its ReLU mask is recomputed from z with the
  // forward's own affine map instead of reading y (one tensor less per pass)
  const int mask_from_z = (beta_or_null != nullptr && y_mask_or_null == nullptr) ? 1 : 0;
  bn_bwd_reduce_kernel<<<grid, kBnThreads, 0, st>>>(static_cast<const uint4*>(dy), static_cast<const uint4*>(y_mask_or_null),
                                                    static_cast<const uint4*>(z), save_mean, save_invstd, gamma, beta_or_null,
                                                    mask_from_z, M, C, sums);
  bn_bwd_apply_kernel<<<grid, kBnThreads, 0, st>>>(static_cast<const uint4*>(dy), static_cast<const uint4*>(y_mask_or_null),
                                                   static_cast<const uint4*>(z), save_mean, save_invstd, gamma, beta_or_null,
                                                   mask_from_z, sums, M, C,
                                                   static_cast<uint4*>(dz), static_cast<uint4*>(g_out_or_null),
                                                   static_cast<uint4*>(stuffed_out_or_null), Ho, Wo, dgamma, dbeta);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

// =================================================================================================================
// Stem im2col, max-pool forward (with argmax) / backward
// =================================================================================================================
namespace pb200 {

// conv1 (7x7 / 2, pad 3) as a GEMM: A[m][k], m = (b, ho, wo), k = (r*7 + s)*C + c for k < 49*C, zero up to K = 256.
// x is the reference-style NCHW fp32 batch (train.py:266-268).  One CTA = one output row (b, ho): the 7 input rows it
// touches are staged in shared memory with coalesced loads (zero padded), then every thread emits 16-byte chunks
// (8 consecutive k) of the 128 A rows, so the 2 GB write is fully coalesced.
template <int C>
__global__ void __launch_bounds__(256)
stem_im2col_kernel(const float* __restrict__ x, int H, int W, int Ho, int Wo, uint4* __restrict__ A) {
  extern __shared__ float s_in[];  // [C][7][W + 6]
  const int Wp = W + 6;
  const int b = blockIdx.x / Ho, ho = blockIdx.x - b * Ho;
  for (int i = threadIdx.x; i < C * 7 * Wp; i += 256) {
    const int wi = i % Wp - 3;
    const int t = i / Wp;
    const int r = t % 7, c = t / 7;
    const int hi = 2 * ho + r - 3;
    float v = 0.f;
    if (hi >= 0 && hi < H && wi >= 0 && wi < W) v = __ldg(x + ((static_cast<long long>(b) * C + c) * H + hi) * W + wi);
    s_in[i] = v;
  }
  __syncthreads();
  uint4* rowbase = A + static_cast<long long>(blockIdx.x) * Wo * 32;  // 32 chunks of 16 B per A row
  for (int i = threadIdx.x; i < Wo * 32; i += 256) {
    const int j = i & 31, wo = i >> 5;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int k = 8 * j + e;
      float v = 0.f;
      if (k < 49 * C) {
        const int tap = k / C, c = k - tap * C;
        const int r = tap / 7, sx = tap - r * 7;
        v = s_in[(c * 7 + r) * Wp + 2 * wo + sx];
      }
      f[e] = v;
    }
    rowbase[i] = make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
  }
}

// conv1.weight fp32 [64][C][7][7] -> bf16 [64][256] with k = (r*7+s)*C + c (zero tail)
__global__ void stem_pack_kernel(const float* __restrict__ w, int C, __nv_bfloat16* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * 256) return;
  const int co = i >> 8, k = i & 255;
  float v = 0.f;
  if (k < 49 * C) {
    const int tap = k / C, c = k - tap * C;
    v = w[(co * C + c) * 49 + tap];
  }
  out[i] = __float2bfloat16_rn(v);
}

// scratch [1][256][64] -> conv1.weight.grad [64][C][7][7]
__global__ void stem_unpack_kernel(const float* __restrict__ scratch, int C, float* __restrict__ dw) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * C * 49) return;
  const int tap = i % 49;
  const int t = i / 49;
  const int c = t % C, co = t / C;
  dw[i] = scratch[(tap * C + c) * 64 + co];
}

// 3x3 / 2 pad 1 max pool, NHWC bf16, 8 channels per thread; idx = position (r*3+s) of the FIRST maximum in scan order
// (torch's tie rule), one byte per channel
__global__ void __launch_bounds__(256)
maxpool_train_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, uint2* __restrict__ idx, int B, int H, int W,
                     int C8, int Ho, int Wo) {
  const unsigned total = static_cast<unsigned>(B) * Ho * Wo * C8;  // host: < 2^31
  const unsigned gstep = gridDim.x * blockDim.x;
  for (unsigned gid = blockIdx.x * blockDim.x + threadIdx.x; gid < total; gid += gstep) {  // one resident wave
  const int c = static_cast<int>(gid % C8);
  unsigned p = gid / C8;
  const int wo = static_cast<int>(p % Wo);
  p /= Wo;
  const int ho = static_cast<int>(p % Ho);
  const int b = static_cast<int>(p / Ho);
  float m[8];
  uint32_t am[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    m[e] = -INFINITY;
    am[e] = 0;
  }
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int hi = 2 * ho + r - 1;
    if (hi < 0 || hi >= H) continue;
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int wi = 2 * wo + s - 1;
      if (wi < 0 || wi >= W) continue;
      float f[8];
      unpack8(__ldg(in + ((static_cast<long long>(b) * H + hi) * W + wi) * C8 + c), f);
#pragma unroll
      for (int e = 0; e < 8; ++e)
        if (f[e] > m[e]) {
          m[e] = f[e];
          am[e] = r * 3 + s;
        }
    }
  }
  out[gid] = pack8(m);
  idx[gid] = make_uint2(am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24),
                        am[4] | (am[5] << 8) | (am[6] << 16) | (am[7] << 24));
  }
}

// gather form: every input pixel sums the gradients of the (<= 4) windows whose argmax it is.  A thread owns the 2x2
// input block (2a..2a+1, 2b..2b+1) x 8 channels: its four pixels only see windows (a | a+1, b | b+1), so four
// (argmax, gradient) loads serve four outputs (one pixel per thread visited nine windows for the same four); one
// resident wave of CTAs walks the blocks (131 k one-vector CTAs with 64-bit div/mod ran at 1.3 TB/s).
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const uint4* __restrict__ dout, const uint2* __restrict__ idx, uint4* __restrict__ din, int B, int H,
                   int W, int C8, int Ho, int Wo) {
  const unsigned total = static_cast<unsigned>(B) * Ho * Wo * C8;  // host: < 2^31
  const unsigned step = gridDim.x * blockDim.x;
  for (unsigned gid = blockIdx.x * blockDim.x + threadIdx.x; gid < total; gid += step) {
    const unsigned c = gid % C8;
    unsigned p = gid / C8;
    const unsigned wb = p % Wo;
    p /= Wo;
    const unsigned ha = p % Ho;
    const unsigned b = p / Ho;
    float acc[2][2][8];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[i][j][e] = 0.f;
#pragma unroll
    for (int dh = 0; dh < 2; ++dh) {
      const unsigned ho = ha + dh;
      if (ho >= static_cast<unsigned>(Ho)) continue;
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        const unsigned wo = wb + dw;
        if (wo >= static_cast<unsigned>(Wo)) continue;
        const long long o = ((static_cast<long long>(b) * Ho + ho) * Wo + wo) * C8 + c;
        const uint2 iv = __ldg(idx + o);
        float g[8];
        unpack8(__ldg(dout + o), g);
        // window (ho, wo) covers input rows 2ho-1..2ho+1: of this block, rows i with r = 2ha + i - (2ho - 1) in [0, 3)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int r = i + 1 - 2 * dh;  // dh = 0: r = 1, 2;  dh = 1: r = -1 (not covered), 0
          if (r < 0) continue;
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int sft = j + 1 - 2 * dw;
            if (sft < 0) continue;
            const uint32_t me = static_cast<uint32_t>(r * 3 + sft);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const uint32_t word = e < 4 ? iv.x : iv.y;
              if (((word >> (8 * (e & 3))) & 0xffu) == me) acc[i][j][e] += g[e];
            }
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j)
        din[((static_cast<long long>(b) * H + 2 * ha + i) * W + 2 * wb + j) * C8 + c] = pack8(acc[i][j]);
  }
}

}  // namespace pb200

extern "C" int pb200_train_stem_im2col(const float* x_nchw, int B, int C, int H, int W, void* A_bf16, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(x_nchw && A_bf16 && C >= 1 && C <= 4 && H % 2 == 0 && W % 2 == 0, "train_stem_im2col: bad argument");
  const int Ho = H / 2, Wo = W / 2;
  const size_t smem = sizeof(float) * C * 7 * (W + 6);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint4* A = static_cast<uint4*>(A_bf16);
  const unsigned grid = static_cast<unsigned>(B) * Ho;
  switch (C) {
    case 1: stem_im2col_kernel<1><<<grid, 256, smem, st>>>(x_nchw, H, W, Ho, Wo, A); break;
    case 2: stem_im2col_kernel<2><<<grid, 256, smem, st>>>(x_nchw, H, W, Ho, Wo, A); break;
    case 3: stem_im2col_kernel<3><<<grid, 256, smem, st>>>(x_nchw, H, W, Ho, Wo, A); break;
    default: stem_im2col_kernel<4><<<grid, 256, smem, st>>>(x_nchw, H, W, Ho, Wo, A); break;
  }
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_train_stem_pack(const float* w_oihw, int C, void* out_bf16, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(w_oihw && out_bf16 && C >= 1 && C <= 4, "train_stem_pack: bad argument");
  stem_pack_kernel<<<64, 256, 0, static_cast<cudaStream_t>(stream)>>>(w_oihw, C, static_cast<__nv_bfloat16*>(out_bf16));
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_train_stem_unpack(const float* scratch, int C, float* dw_oihw, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(scratch && dw_oihw && C >= 1 && C <= 4, "train_stem_unpack: bad argument");
  const int total = 64 * C * 49;
  stem_unpack_kernel<<<(total + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(scratch, C, dw_oihw);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_train_maxpool(const void* in, void* out, void* idx, int B, int H, int W, int C, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(in && out && idx && C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "train_maxpool: bad argument");
  const int Ho = H / 2, Wo = W / 2;
  const long long total = static_cast<long long>(B) * Ho * Wo * (C / 8);
  PB_CHECK(total < (1LL << 31), "train_maxpool: tensor too large");
  long long grid = (total + 255) / 256;
  if (grid > static_cast<long long>(num_sms()) * 8) grid = static_cast<long long>(num_sms()) * 8;
  maxpool_train_kernel<<<static_cast<unsigned>(grid), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(in), static_cast<uint4*>(out), static_cast<uint2*>(idx), B, H, W, C / 8, Ho, Wo);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_train_maxpool_backward(const void* dout, const void* idx, void* din, int B, int H, int W, int C,
                                            pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(dout && idx && din && C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "train_maxpool_backward: bad argument");
  const int Ho = H / 2, Wo = W / 2;
  const long long total = static_cast<long long>(B) * Ho * Wo * (C / 8);
  PB_CHECK(total < (1LL << 31), "train_maxpool_backward: tensor too large");
  long long grid = (total + 255) / 256;
  if (grid > static_cast<long long>(num_sms()) * 8) grid = static_cast<long long>(num_sms()) * 8;
  maxpool_bwd_kernel<<<static_cast<unsigned>(grid), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(dout), static_cast<const uint2*>(idx), static_cast<uint4*>(din), B, H, W, C / 8, Ho, Wo);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

// =================================================================================================================
// Regression head (avgpool + fc), SmoothL1 loss, their backward, and the fused optimizer
// =================================================================================================================
namespace pb200 {

// forward: feat[b][c] = mean over HW of y[b][hw][c]; pred[b][o] = feat . W[o] + bias[o]   (one CTA per frame)
__global__ void __launch_bounds__(256)
head_train_fwd_kernel(const __nv_bfloat16* __restrict__ y, const float* __restrict__ fc_w, const float* __restrict__ fc_b,
                      float* __restrict__ feat, float* __restrict__ pred, int HW, int C, int n_out) {
  extern __shared__ float s_mean[];
  const int b = blockIdx.x;
  const float inv = 1.0f / static_cast<float>(HW);
  for (int c2 = threadIdx.x; c2 < C / 2; c2 += blockDim.x) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(y) + static_cast<long long>(b) * HW * (C / 2) + c2;
    float a0 = 0.f, a1 = 0.f;
    for (int i = 0; i < HW; ++i) {
      const uint32_t v = __ldg(p + static_cast<long long>(i) * (C / 2));
      a0 += bf16_lo(v);
      a1 += bf16_hi(v);
    }
    s_mean[2 * c2] = a0 * inv;
    s_mean[2 * c2 + 1] = a1 * inv;
    feat[static_cast<long long>(b) * C + 2 * c2] = a0 * inv;
    feat[static_cast<long long>(b) * C + 2 * c2 + 1] = a1 * inv;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int o = warp; o < n_out; o += (blockDim.x >> 5)) {
    float acc = 0.f;
    for (int c = lane; c < C; c += 32) acc = fmaf(s_mean[c], __ldg(fc_w + o * C + c), acc);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if (lane == 0) pred[static_cast<long long>(b) * n_out + o] = acc + __ldg(fc_b + o);
  }
}

// SmoothL1Loss(beta, reduction = mean) (train.py:119,290) and d loss / d pred, one CTA
__global__ void __launch_bounds__(256)
smooth_l1_kernel(const float* __restrict__ pred, const float* __restrict__ target, int n, float beta,
                 float* __restrict__ loss, float* __restrict__ dpred) {
  __shared__ float s_part[256];
  float acc = 0.f;
  const float invn = 1.0f / static_cast<float>(n);
  for (int i = threadIdx.x; i < n; i += 256) {
    const float d = pred[i] - target[i];
    const float ad = fabsf(d);
    if (ad < beta) {
      acc += 0.5f * d * d / beta;
      dpred[i] = d / beta * invn;
    } else {
      acc += ad - 0.5f * beta;
      dpred[i] = (d > 0.f ? 1.f : -1.f) * invn;
    }
  }
  s_part[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) s_part[threadIdx.x] += s_part[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) *loss = s_part[0] * invn;
}

// d fc.weight[o][c] = sum_b dpred[b][o] feat[b][c]; d fc.bias[o] = sum_b dpred[b][o].  CTA = 32 channels (lanes) x 8 batch
// slices (warps); a thread keeps all n_out (<= 32) partial sums of its channel, the warps are combined through shared
// memory.  (One CTA per output walking the whole batch sequentially took 110 us for 0.5 MB of input.)
__global__ void __launch_bounds__(256)
head_bwd_params_kernel(const float* __restrict__ dpred, const float* __restrict__ feat, int B, int C, int n_out,
                       float* __restrict__ dW, float* __restrict__ db) {
  __shared__ float s_part[8][32][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float acc[32];
#pragma unroll
  for (int o = 0; o < 32; ++o) acc[o] = 0.f;
  if (c < C) {
    for (int b = warp; b < B; b += 8) {
      const float f = __ldg(feat + static_cast<long long>(b) * C + c);
#pragma unroll
      for (int o = 0; o < 32; ++o)
        if (o < n_out) acc[o] = fmaf(__ldg(dpred + b * n_out + o), f, acc[o]);
    }
  }
#pragma unroll
  for (int o = 0; o < 32; ++o) s_part[warp][o][lane] = acc[o];
  __syncthreads();
  for (int i = threadIdx.x; i < n_out * 32; i += 256) {
    const int o = i >> 5, l = i & 31;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += s_part[w][o][l];
    if (blockIdx.x * 32 + l < C) dW[o * C + blockIdx.x * 32 + l] = t;
  }
  if (blockIdx.x == 0 && threadIdx.x < n_out) {
    float t = 0.f;
    for (int b = 0; b < B; ++b) t += dpred[b * n_out + threadIdx.x];
    db[threadIdx.x] = t;
  }
}

// dy[b][hw][c] = (sum_o dpred[b][o] W[o][c]) / HW, broadcast over the pixels (one CTA per frame)
__global__ void __launch_bounds__(256)
head_bwd_input_kernel(const float* __restrict__ dpred, const float* __restrict__ fc_w, int HW, int C, int n_out,
                      __nv_bfloat16* __restrict__ dy) {
  extern __shared__ float s_df[];
  const int b = blockIdx.x;
  const float inv = 1.0f / static_cast<float>(HW);
  for (int c = threadIdx.x; c < C; c += 256) {
    float acc = 0.f;
    for (int o = 0; o < n_out; ++o) acc = fmaf(dpred[b * n_out + o], __ldg(fc_w + o * C + c), acc);
    s_df[c] = acc * inv;
  }
  __syncthreads();
  __nv_bfloat16* out = dy + static_cast<long long>(b) * HW * C;
  for (int i = threadIdx.x; i < HW * C; i += 256) out[i] = __float2bfloat16_rn(s_df[i % C]);
}

// ---- optimizer: global gradient norm, clip, AdamW (train.py:301-303; torch.optim.AdamW defaults) ----------------
__global__ void __launch_bounds__(256)
grad_sqnorm_kernel(const float* __restrict__ g, long long n, float grad_scale, double* __restrict__ out) {
  __shared__ float s_part[256];
  float acc = 0.f;
  for (long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * 256) {
    const float v = g[i] * grad_scale;
    acc = fmaf(v, v, acc);
  }
  s_part[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) s_part[threadIdx.x] += s_part[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) atomicAdd(out, static_cast<double>(s_part[0]));
}

// decay[i] (0/1 per element would cost a byte per parameter): torch's AdamW applies weight decay to EVERY parameter of
// the group (train.py:199 passes model.parameters()), so one scalar wd is exact.
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
             long long n, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt, float grad_scale,
             float max_norm, const double* __restrict__ sqnorm) {
  float clip = 1.f;
  if (max_norm > 0.f) {  // clip_grad_norm_: coef = max_norm / (total_norm + 1e-6), clamped to 1
    const float total = static_cast<float>(sqrt(*sqnorm));
    clip = fminf(max_norm / (total + 1e-6f), 1.f);
  }
  const float gs = grad_scale * clip;
  for (long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * 256) {
    const float gi = g[i] * gs;
    float pi = p[i];
    pi -= lr * wd * pi;
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = pi - (lr / bc1) * (mi / denom);
  }
}

}  // namespace pb200

extern "C" int pb200_train_head_forward(const void* y_nhwc, const float* fc_w, const float* fc_b, float* feat, float* pred,
                                        int B, int HW, int C, int n_out, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(y_nhwc && fc_w && fc_b && feat && pred && C % 2 == 0, "train_head_forward: bad argument");
  head_train_fwd_kernel<<<B, 256, sizeof(float) * C, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(y_nhwc), fc_w, fc_b, feat, pred, HW, C, n_out);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_train_smooth_l1(const float* pred, const float* target, int n, float beta, float* loss, float* dpred,
                                     pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(pred && target && loss && dpred && n >= 1 && beta > 0.f, "train_smooth_l1: bad argument");
  smooth_l1_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(pred, target, n, beta, loss, dpred);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_train_head_backward(const float* dpred, const float* feat, const float* fc_w, int B, int HW, int C,
                                         int n_out, float* d_fc_w, float* d_fc_b, void* dy_nhwc, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(dpred && feat && fc_w && d_fc_w && d_fc_b && dy_nhwc, "train_head_backward: null argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PB_CHECK(n_out >= 1 && n_out <= 32, "train_head_backward: n_out = %d unsupported (1..32)", n_out);
  head_bwd_params_kernel<<<(C + 31) / 32, 256, 0, st>>>(dpred, feat, B, C, n_out, d_fc_w, d_fc_b);
  head_bwd_input_kernel<<<B, 256, sizeof(float) * C, st>>>(dpred, fc_w, HW, C, n_out, static_cast<__nv_bfloat16*>(dy_nhwc));
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

extern "C" int pb200_train_adamw_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                                      float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                                      float grad_scale, float max_norm, double* sqnorm_scratch, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(params && grads && exp_avg && exp_avg_sq && sqnorm_scratch && n >= 1 && step >= 1, "train_adamw_step: bad argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = num_sms() * 4;
  PB_CUDA(cudaMemsetAsync(sqnorm_scratch, 0, sizeof(double), st));
  if (max_norm > 0.f) grad_sqnorm_kernel<<<grid, 256, 0, st>>>(grads, n, grad_scale, sqnorm_scratch);
  const float bc1 = 1.f - powf(beta1, static_cast<float>(step));
  const float bc2_sqrt = sqrtf(1.f - powf(beta2, static_cast<float>(step)));
  adamw_kernel<<<grid, 256, 0, st>>>(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, bc1,
                                     bc2_sqrt, grad_scale, max_norm, sqnorm_scratch);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

// ---- all weight packs of a step in ONE launch -----------------------------------------------------------------
// table[n][8] (int64, device): {w fp32 OIHW, out forward bf16, out data-gradient bf16 (0: none), Cout, Cin, R,
// first TILE index, kind (0: conv, 1: conv1 as [64][256] im2col GEMM operand with Cin real channels)}.
// A tile of a conv = 32 output x 32 input channels x R*R taps, staged through shared memory: the OIHW rows are read as
// 32 runs of 32*R*R contiguous floats, the forward operand [co][(r,s)][ci] and the data-gradient operand
// [ci][(2-r,2-s)][co] are both written in 64-byte runs.  (One thread per element gathered 4-byte words 36 bytes apart
// and scattered 2-byte words: 178 us for 45 MB.)  A conv1 "tile" = 256 elements of its [64][256] operand.
namespace pb200 {
constexpr int kPackT = 32;
__global__ void __launch_bounds__(256)
pack_all_kernel(const long long* __restrict__ table, int n, long long total_tiles) {
  __shared__ long long s_t[32 * 8];
  __shared__ float s_w[kPackT][kPackT * 9 + 1];
  for (int i = threadIdx.x; i < n * 8; i += 256) s_t[i] = table[i];
  __syncthreads();
  const long long tile = blockIdx.x;
  if (tile >= total_tiles) return;
  int e = 0;
  while (e + 1 < n && tile >= s_t[(e + 1) * 8 + 6]) ++e;
  const long long* d = s_t + e * 8;
  const float* w = reinterpret_cast<const float*>(d[0]);
  __nv_bfloat16* of = reinterpret_cast<__nv_bfloat16*>(d[1]);
  __nv_bfloat16* od = reinterpret_cast<__nv_bfloat16*>(d[2]);
  const int Cout = static_cast<int>(d[3]), Cin = static_cast<int>(d[4]), R = static_cast<int>(d[5]);
  const int t = static_cast<int>(tile - d[6]);
  if (d[7] == 1) {  // conv1: elements [256 t, 256 t + 256) of [64][256]
    const int i = t * 256 + threadIdx.x;
    const int co = i >> 8, k = i & 255;
    float v = 0.f;
    if (k < 49 * Cin) {
      const int tap = k / Cin, c = k - tap * Cin;
      v = w[(co * Cin + c) * 49 + tap];
    }
    of[i] = __float2bfloat16_rn(v);
    return;
  }
  const int RR = R * R;
  const int cbs = Cin / kPackT;
  const int ob = t / cbs, cb = t - ob * cbs;
  const int run = kPackT * RR;  // contiguous floats of one output channel's 32 input channels
  for (int i = threadIdx.x; i < kPackT * run; i += 256) {
    const int col = i / run, j = i - col * run;
    s_w[col][j] = w[(static_cast<long long>(ob * kPackT + col) * Cin + cb * kPackT) * RR + j];
  }
  __syncthreads();
  // forward operand: [co][q = (r, s)][ci]
  for (int i = threadIdx.x; i < kPackT * run; i += 256) {
    const int cil = i & (kPackT - 1);
    const int rest = i >> 5;
    const int q = rest % RR, col = rest / RR;
    of[(static_cast<long long>(ob * kPackT + col) * RR + q) * Cin + cb * kPackT + cil] = __float2bfloat16_rn(s_w[col][cil * RR + q]);
  }
  if (od != nullptr) {  // data-gradient operand: [ci][q2 = flipped (r, s)][co]
    for (int i = threadIdx.x; i < kPackT * run; i += 256) {
      const int col = i & (kPackT - 1);
      const int rest = i >> 5;
      const int q2 = rest % RR, cil = rest / RR;
      od[(static_cast<long long>(cb * kPackT + cil) * RR + q2) * Cout + ob * kPackT + col] =
          __float2bfloat16_rn(s_w[col][cil * RR + (RR - 1 - q2)]);
    }
  }
}
}  // namespace pb200

extern "C" int pb200_train_pack_all(const int64_t* table_dev, int n, int64_t total, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(table_dev && n >= 1 && n <= 32 && total >= 1 && total < (1LL << 31), "train_pack_all: bad argument");
  pack_all_kernel<<<static_cast<unsigned>(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const long long*>(table_dev), n, total);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}
