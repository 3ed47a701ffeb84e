// fp32 SIMT implicit-GEMM convolution (NHWC) — the arithmetic of PB200_PREC_FP32, the mode that
// reproduces the reference's fp32 forward to <= 1e-3 px.  FFMA only, fp32 accumulation in K
// order; eval-mode BatchNorm is already folded into w / bias, ReLU and the residual add are
// fused in the epilogue (torchvision BasicBlock.forward, resnet.py:89-105).
#include "common.cuh"
#include "layout.h"

namespace pb200 {

constexpr int SBM = 64, SBN = 64, SBK = 16;

__global__ void __launch_bounds__(256) conv_simt_kernel(SimtConv c) {
  __shared__ float As[SBK][SBM + 4];
  __shared__ float Bs[SBK][SBN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const long long M = static_cast<long long>(c.B) * c.Ho * c.Wo;
  const int K = c.R * c.S * c.Cin;
  const long long m0 = static_cast<long long>(blockIdx.x) * SBM;
  const int n0 = blockIdx.y * SBN;

  // the four pixels this thread gathers for the A tile
  const int kk_a = tid & 15;
  int hi0[4], wi0[4];
  long long pbase[4];
  bool pvalid[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + (tid >> 4) + 16 * i;
    pvalid[i] = m < M;
    const long long mm = pvalid[i] ? m : 0;
    const int wo = static_cast<int>(mm % c.Wo);
    const long long t = mm / c.Wo;
    const int ho = static_cast<int>(t % c.Ho);
    const int b = static_cast<int>(t / c.Ho);
    hi0[i] = ho * c.stride - c.pad;
    wi0[i] = wo * c.stride - c.pad;
    pbase[i] = static_cast<long long>(b) * c.Hin * c.Win * c.Cin;
  }
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += SBK) {
    {
      const int k = k0 + kk_a;
      const bool kv = k < K;
      const int kc = kv ? k : 0;
      const int tap = kc / c.Cin, ch = kc - tap * c.Cin;
      const int r = tap / c.S, s = tap - r * c.S;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int hi = hi0[i] + r, wi = wi0[i] + s;
        float v = 0.f;
        if (kv && pvalid[i] && hi >= 0 && hi < c.Hin && wi >= 0 && wi < c.Win)
          v = __ldg(c.in + pbase[i] + (static_cast<long long>(hi) * c.Win + wi) * c.Cin + ch);
        As[kk_a][(tid >> 4) + 16 * i] = v;
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + 256 * i;
      const int kk = idx >> 6, nn = idx & 63;
      float v = 0.f;
      if (k0 + kk < K && n0 + nn < c.Cout) v = __ldg(c.w + static_cast<long long>(k0 + kk) * c.Cout + n0 + nn);
      Bs[kk][nn] = v;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < SBK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= c.Cout) continue;
      float v = acc[i][j] + __ldg(c.bias + n);
      if (c.residual != nullptr) v += __ldg(c.residual + m * c.Cout + n);
      if (c.relu) v = fmaxf(v, 0.f);
      c.out[m * c.Cout + n] = v;
    }
  }
}

int launch_conv_simt(const SimtConv& c, cudaStream_t st) {
  const long long M = static_cast<long long>(c.B) * c.Ho * c.Wo;
  dim3 grid(static_cast<unsigned>((M + SBM - 1) / SBM), static_cast<unsigned>((c.Cout + SBN - 1) / SBN));
  conv_simt_kernel<<<grid, 256, 0, st>>>(c);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

}  // namespace pb200
