// fp32 SIMT implicit-GEMM convolution (NHWC) — the arithmetic of PB200_PREC_FP32, the mode that
// reproduces the reference's fp32 forward to <= 1e-3 px.  FFMA only, fp32 accumulation in K
// order; eval-mode BatchNorm is already folded into w / bias, ReLU and the residual add are
// fused in the epilogue (torchvision BasicBlock.forward, resnet.py:89-105).
#include <cstdlib>
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

// ---- dedicated fp32 stem (conv1 7x7/2 on the zero-margined frames buffer) ------------------------------
// The generic kernel above gathers one scalar per (pixel, k) with a division per K step; on the stem (the
// largest GEMM of the fp32 and split-bf16 modes: M = B*128*128, K = 224) it is shared-memory bound.  Here a
// persistent CTA keeps the whole [224][64] filter in shared memory and walks 8 x 32 output tiles: the input
// patch (21 rows x 70 pixels x 4 channels) is staged once, split by column parity so that the 32 lanes of a
// warp (= 32 consecutive output columns of a stride-2 conv) read consecutive 16-byte pixels for every tap;
// a warp owns 8 output channels (filter reads are broadcasts), a thread 8 output rows x 8 channels.
// Same K order (r, tap, channel) as the generic kernel, so the sums are bit-identical; the zero-weight
// tap 0 (layout.h) and, for RGB models, the empty fourth channel are skipped.
constexpr int ST_ROWS = 8, ST_COLS = 32;
constexpr int ST_PR = 2 * ST_ROWS + 5;   // patch rows
constexpr int ST_PCOLS = 2 * ST_COLS + 6;  // patch pixels per row
constexpr int ST_PC = 36;                // float4 per parity plane and row (35 used)
constexpr int ST_W_BYTES = kStemK * 64 * 4;
constexpr int ST_SMEM = ST_W_BYTES + 2 * ST_PR * ST_PC * 16;

template <int NC, int MODE>  // MODE 0: scalar FFMA, 1: FFMA2, 2: FFMA2 for six of the eight rows + scalar FFMA for two
__global__ void __launch_bounds__(256, 2)
stem_f32_kernel(const float4* __restrict__ frames, const float* __restrict__ w, const float* __restrict__ bias,
                float* __restrict__ out, int B, int H, int W, int relu) {
  extern __shared__ __align__(16) unsigned char st_smem[];
  float4* ws = reinterpret_cast<float4*>(st_smem);                  // [k][16] float4 = [k][co]
  float4* patch = reinterpret_cast<float4*>(st_smem + ST_W_BYTES);  // [parity][row][ST_PC]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < kStemK * 16; i += 256) ws[i] = __ldg(reinterpret_cast<const float4*>(w) + i);
  const int Ho = H / 2, Wo = W / 2, FH = H + kFrameMarginTop + kFrameMarginBottom,
            FW = W + kFrameMarginLeft + kFrameMarginRight;
  const int tiles_x = Wo / ST_COLS, tiles_y = Ho / ST_ROWS;
  const long long ntiles = static_cast<long long>(B) * tiles_y * tiles_x;
  float bv[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) bv[j] = __ldg(bias + warp * 8 + j);

  for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int tx = static_cast<int>(t % tiles_x);
    const int ty = static_cast<int>((t / tiles_x) % tiles_y);
    const long long b = t / (static_cast<long long>(tiles_x) * tiles_y);
    const float4* src = frames + (b * FH + 2 * ty * ST_ROWS) * FW + 2 * tx * ST_COLS;
    __syncthreads();  // the previous tile's reads of the patch are done
    for (int i = tid; i < ST_PR * ST_PCOLS; i += 256) {
      const int r = i / ST_PCOLS, c = i - r * ST_PCOLS;
      patch[((c & 1) * ST_PR + r) * ST_PC + (c >> 1)] = __ldg(src + static_cast<long long>(r) * FW + c);
    }
    __syncthreads();
    // accumulators as channel pairs: one FFMA2 (fma.rn.f32x2) per pair, same per-element arithmetic and order
    float2 acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = make_float2(0.f, 0.f);
#pragma unroll 1
    for (int r = 0; r < 7; ++r) {
#pragma unroll
      for (int s8 = 1; s8 < kStemTaps; ++s8) {
        const float4* pa = patch + ((s8 & 1) * ST_PR + r) * ST_PC + (s8 >> 1) + lane;
        float4 a[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = pa[2 * i * ST_PC];  // output row i reads patch row 2i + r
        const float4* wk = ws + (r * (kStemTaps * kFrameC) + s8 * kFrameC) * 16 + warp * 2;
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          const float4 w0 = wk[c * 16], w1 = wk[c * 16 + 1];
          const float2 wv[4] = {make_float2(w0.x, w0.y), make_float2(w0.z, w0.w), make_float2(w1.x, w1.y),
                                make_float2(w1.z, w1.w)};
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float av = c == 0 ? a[i].x : (c == 1 ? a[i].y : (c == 2 ? a[i].z : a[i].w));
            const float2 av2 = make_float2(av, av);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (MODE == 1 || (MODE == 2 && (i % 4) != 3)) {
                acc[i][j] = __ffma2_rn(av2, wv[j], acc[i][j]);
              } else {
                acc[i][j].x = fmaf(av, wv[j].x, acc[i][j].x);
                acc[i][j].y = fmaf(av, wv[j].y, acc[i][j].y);
              }
            }
          }
        }
      }
    }
    float* o = out + ((b * Ho + ty * ST_ROWS) * Wo + tx * ST_COLS + lane) * 64 + warp * 8;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float v[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[2 * j] = acc[i][j].x + bv[2 * j];
        v[2 * j + 1] = acc[i][j].y + bv[2 * j + 1];
      }
      if (relu) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
      }
      float4* dst = reinterpret_cast<float4*>(o + static_cast<long long>(i) * Wo * 64);
      dst[0] = make_float4(v[0], v[1], v[2], v[3]);
      dst[1] = make_float4(v[4], v[5], v[6], v[7]);
    }
  }
}

bool stem_f32_supported(const SimtConv& c) {
  return c.Cin == kFrameC && c.R == 7 && c.S == kStemTaps && c.stride == 2 && c.pad == 0 && c.Cout == 64 &&
         c.residual == nullptr && c.Ho % ST_ROWS == 0 && c.Wo % ST_COLS == 0 &&
         c.Hin == 2 * c.Ho + kFrameMarginTop + kFrameMarginBottom &&
         c.Win == 2 * c.Wo + kFrameMarginLeft + kFrameMarginRight;
}

int launch_stem_f32(const SimtConv& c, int real_channels, cudaStream_t st) {
  // PB200_STEM_F32_MODE (read per call, for A/B and the parity test): 0 scalar FFMA, 1 packed FFMA2 (default),
  // 2 mixed, -1 the generic conv_simt_kernel on this layer
  const char* env = getenv("PB200_STEM_F32_MODE");
  const int mode = env ? atoi(env) : 1;
  if (mode < 0) return launch_conv_simt(c, st);
  typedef void (*kern_t)(const float4*, const float*, const float*, float*, int, int, int, int);
  static const kern_t table[2][3] = {{stem_f32_kernel<3, 0>, stem_f32_kernel<3, 1>, stem_f32_kernel<3, 2>},
                                     {stem_f32_kernel<4, 0>, stem_f32_kernel<4, 1>, stem_f32_kernel<4, 2>}};
  static bool attr_set[kMaxDevices] = {};  // function attributes belong to the device's context
  if (!attr_set[current_device()]) {
    for (int a = 0; a < 2; ++a)
      for (int m = 0; m < 3; ++m)
        PB_CUDA(cudaFuncSetAttribute(table[a][m], cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM));
    attr_set[current_device()] = true;
  }
  const long long ntiles = static_cast<long long>(c.B) * (c.Ho / ST_ROWS) * (c.Wo / ST_COLS);
  const long long cap = 2LL * num_sms();
  const unsigned grid = static_cast<unsigned>(ntiles < cap ? ntiles : cap);
  const float4* in = reinterpret_cast<const float4*>(c.in);
  const kern_t k = table[real_channels <= 3 ? 0 : 1][mode >= 0 && mode <= 2 ? mode : 1];
  k<<<grid, 256, ST_SMEM, st>>>(in, c.w, c.bias, c.out, c.B, 2 * c.Ho, 2 * c.Wo, c.relu);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

int launch_conv_simt(const SimtConv& c, cudaStream_t st) {
  const long long M = static_cast<long long>(c.B) * c.Ho * c.Wo;
  dim3 grid(static_cast<unsigned>((M + SBM - 1) / SBM), static_cast<unsigned>((c.Cout + SBN - 1) / SBN));
  conv_simt_kernel<<<grid, 256, 0, st>>>(c);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

}  // namespace pb200
