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
                                const void* residual_or_null, int relu, void* out_nhwc, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  if (int rc = umma_conv_init()) return rc;
  PB_CHECK(in_nhwc && wpacked && bias && out_nhwc, "train_conv: null argument");
  UmmaConvLaunch L;
  int bn = Cout < 256 ? Cout : 256;
  // same tile choice as the inference plan: rounds x MMA cost
  {
    const int Ho = (H + 2 * pad - R) / stride + 1, Wo = (W + 2 * pad - R) / stride + 1;
    const long long m_tiles = (static_cast<long long>(B) * Ho * Wo + 127) / 128;
    long long best_cost = -1;
    for (int c = 64; c <= 256 && c <= Cout; c *= 2) {
      if (Cout % c) continue;
      const long long tiles = m_tiles * (Cout / c);
      const long long rounds = (tiles + num_sms() - 1) / num_sms();
      const long long cost = rounds * (c / 2 < 48 ? 48 : c / 2);
      if (best_cost < 0 || cost <= best_cost) {
        bn = c;
        best_cost = cost;
      }
    }
  }
  if (int rc = build_umma_conv(&L, in_nhwc, B, H, W, Cin, wpacked, bias, residual_or_null, out_nhwc, Cout, R, stride,
                               pad, relu, bn))
    return rc;
  return launch_umma_conv(L, static_cast<cudaStream_t>(stream));
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
