// Frame ingest: camera / dataset pixels -> the plan's "frames" layout, in one pass.
//
// frames layout: [B][H+6][W+8][4] with a zero margin of 3 rows top/bottom and
// 4 pixels left/right, channels (R, G, B, D).  bf16 in the tcgen05 modes — RGB stored as the
// exact integers 0..255 (1/255 is folded into the stem weights); fp16 in the "fp16x2" mode (k * 2^-12, depth * 2^-4) — fp32 in PB200_PREC_FP32, where
// RGB is k/255 exactly as the reference computes it.
//
// Restates scripts/streaming.py:68-80 (BGRA slice + reverse + /255, NaN/Inf -> 0, /0.035,
// centre crop), perseus/detector/data.py:84-85 (u8 -> f32 /255, HWC -> CHW),
// validate.py:122-124 (RGBD concat) and augmentations.py:137-168 (val-mode near/far clip).
#include "common.cuh"
#include "layout.h"

namespace pb200 {

template <typename T>
__device__ __forceinline__ void store_px4(T* dst, const float (&v)[16]);
template <>
__device__ __forceinline__ void store_px4<float>(float* dst, const float (&v)[16]) {
  float4* d = reinterpret_cast<float4*>(dst);
#pragma unroll
  for (int i = 0; i < 4; ++i) d[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
}
template <>
__device__ __forceinline__ void store_px4<__nv_bfloat16>(__nv_bfloat16* dst, const float (&v)[16]) {
  uint4* d = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int i = 0; i < 2; ++i)
    d[i] = make_uint4(pack_bf16x2(v[8 * i], v[8 * i + 1]), pack_bf16x2(v[8 * i + 2], v[8 * i + 3]),
                      pack_bf16x2(v[8 * i + 4], v[8 * i + 5]), pack_bf16x2(v[8 * i + 6], v[8 * i + 7]));
}

// fp16 frames: RGB scaled by 2^-12, depth by 2^-4 (layout.h), both exact
template <>
__device__ __forceinline__ void store_px4<__half>(__half* dst, const float (&v)[16]) {
  uint4* d = reinterpret_cast<uint4*>(dst);
  float s[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) s[i] = v[i] * ((i & 3) == 3 ? kF16DepthScale : kF16RgbScale);
#pragma unroll
  for (int i = 0; i < 2; ++i)
    d[i] = make_uint4(pack_f16x2(s[8 * i], s[8 * i + 1]), pack_f16x2(s[8 * i + 2], s[8 * i + 3]),
                      pack_f16x2(s[8 * i + 4], s[8 * i + 5]), pack_f16x2(s[8 * i + 6], s[8 * i + 7]));
}

// One thread = 4 horizontally adjacent pixels of the zero-margined output (one 16/12-byte source
// read, one 16-byte depth read, 32/64 bytes written with 128-bit stores).
template <typename T, int CH>
__global__ void __launch_bounds__(256) ingest_kernel(IngestArgs a, T* __restrict__ out) {
  const int Hp = a.H + kFrameMarginTop + kFrameMarginBottom;
  const int Wp = a.W + kFrameMarginLeft + kFrameMarginRight;
  const int gpr = Wp / 4;  // groups per row
  const long long total = static_cast<long long>(a.B) * Hp * gpr;
  const long long gid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int g = static_cast<int>(gid % gpr);
  const long long row = gid / gpr;
  const int hp = static_cast<int>(row % Hp);
  const int b = static_cast<int>(row / Hp);
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = 0.f;
  const int y = hp - kFrameMarginTop;
  const int x = g * 4 - kFrameMarginLeft;  // margins are whole groups (left margin = 4)
  if (y >= 0 && y < a.H && x >= 0 && x < a.W) {
    const int sy = a.y0 + y, sx = a.x0 + x;
    // ---- colour: 4 pixels = 12 or 16 source bytes
    const uint8_t* src = a.rgb + b * a.frame_pitch + sy * a.pitch + static_cast<long long>(sx) * CH;
    uint8_t px[16];
    if (CH == 4) {
      if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        *reinterpret_cast<uint4*>(px) = __ldg(reinterpret_cast<const uint4*>(src));
      } else if ((reinterpret_cast<uintptr_t>(src) & 3) == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          reinterpret_cast<uint32_t*>(px)[i] = __ldg(reinterpret_cast<const uint32_t*>(src) + i);
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) px[i] = __ldg(src + i);
      }
    } else {
      if ((reinterpret_cast<uintptr_t>(src) & 3) == 0) {
#pragma unroll
        for (int i = 0; i < 3; ++i)
          reinterpret_cast<uint32_t*>(px)[i] = __ldg(reinterpret_cast<const uint32_t*>(src) + i);
      } else {
#pragma unroll
        for (int i = 0; i < 12; ++i) px[i] = __ldg(src + i);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float c0 = static_cast<float>(px[i * CH]);
      const float gr = static_cast<float>(px[i * CH + 1]);
      const float c2 = static_cast<float>(px[i * CH + 2]);
      const float r = a.bgr ? c2 : c0;   // streaming.py:69 reverses BGR -> RGB
      const float bl = a.bgr ? c0 : c2;
      if (a.rgb_integer) {
        v[4 * i] = r; v[4 * i + 1] = gr; v[4 * i + 2] = bl;
      } else {
        // IEEE fp32 division: identical to numpy's float32(k)/255 (data.py:84) and to
        // float32(float64(k)/255.0) (streaming.py:69,127) for every k in 0..255
        v[4 * i] = __fdiv_rn(r, 255.0f);
        v[4 * i + 1] = __fdiv_rn(gr, 255.0f);
        v[4 * i + 2] = __fdiv_rn(bl, 255.0f);
      }
    }
    // ---- depth
    if (a.depth != nullptr) {
      const float* dsrc = a.depth + (static_cast<long long>(b) * a.src_h + sy) * a.src_w + sx;
      float d[4];
      if ((reinterpret_cast<uintptr_t>(dsrc) & 15) == 0) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(dsrc));
        d[0] = t.x; d[1] = t.y; d[2] = t.z; d[3] = t.w;
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) d[i] = __ldg(dsrc + i);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float z = d[i];
        if (isnan(z) || isinf(z)) z = 0.f;                 // streaming.py:74-75
        z = __fdiv_rn(z, a.depth_divisor);                 // streaming.py:76
        if (a.clip) {                                      // augmentations.py:137-168
          float s = __fmul_rn(a.cube_scale, z);
          float jn = 0.f, jf = 0.f;
          if (a.jitter != nullptr) {
            const long long jb = ((static_cast<long long>(b) * 2) * a.H + y) * a.W + x + i;
            jn = __ldg(a.jitter + jb);
            jf = __ldg(a.jitter + jb + static_cast<long long>(a.H) * a.W);
          }
          if (s < __fadd_rn(a.near_plane, jn)) s = 0.f;
          if (s > __fadd_rn(a.far_plane, jf)) s = 0.f;
          z = __fdiv_rn(s, a.cube_scale);
        }
        v[4 * i + 3] = z;
      }
    }
  }
  store_px4<T>(out + ((static_cast<long long>(b) * Hp + hp) * Wp + g * 4) * 4, v);
}

// NCHW fp32 (the reference forward() argument) -> frames.  C = 3 or 4.
template <typename T>
__global__ void __launch_bounds__(256)
repack_nchw_kernel(const float* __restrict__ x, int C, int H, int W, int B, int rgb_integer,
                   T* __restrict__ out) {
  const int Hp = H + kFrameMarginTop + kFrameMarginBottom;
  const int Wp = W + kFrameMarginLeft + kFrameMarginRight;
  const int gpr = Wp / 4;
  const long long total = static_cast<long long>(B) * Hp * gpr;
  const long long gid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int g = static_cast<int>(gid % gpr);
  const long long row = gid / gpr;
  const int hp = static_cast<int>(row % Hp);
  const int b = static_cast<int>(row / Hp);
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = 0.f;
  const int y = hp - kFrameMarginTop;
  const int xx = g * 4 - kFrameMarginLeft;
  if (y >= 0 && y < H && xx >= 0 && xx < W) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if (c >= C) break;
      const float* src = x + ((static_cast<long long>(b) * C + c) * H + y) * W + xx;
      float d[4];
      if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(src));
        d[0] = t.x; d[1] = t.y; d[2] = t.z; d[3] = t.w;
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) d[i] = __ldg(src + i);
      }
      const float sc = (rgb_integer && c < 3) ? 255.0f : 1.0f;
#pragma unroll
      for (int i = 0; i < 4; ++i) v[4 * i + c] = d[i] * sc;
    }
  }
  store_px4<T>(out + ((static_cast<long long>(b) * Hp + hp) * Wp + g * 4) * 4, v);
}

int launch_ingest(const IngestArgs& a, void* out, int elt, cudaStream_t st) {
  const int Hp = a.H + kFrameMarginTop + kFrameMarginBottom;
  const int Wp = a.W + kFrameMarginLeft + kFrameMarginRight;
  const long long total = static_cast<long long>(a.B) * Hp * (Wp / 4);
  const unsigned grid = static_cast<unsigned>((total + 255) / 256);
  if (elt == FRAME_F16) {
    if (a.ch == 4)
      ingest_kernel<__half, 4><<<grid, 256, 0, st>>>(a, static_cast<__half*>(out));
    else
      ingest_kernel<__half, 3><<<grid, 256, 0, st>>>(a, static_cast<__half*>(out));
  } else if (elt == FRAME_BF16) {
    if (a.ch == 4)
      ingest_kernel<__nv_bfloat16, 4><<<grid, 256, 0, st>>>(a, static_cast<__nv_bfloat16*>(out));
    else
      ingest_kernel<__nv_bfloat16, 3><<<grid, 256, 0, st>>>(a, static_cast<__nv_bfloat16*>(out));
  } else {
    if (a.ch == 4)
      ingest_kernel<float, 4><<<grid, 256, 0, st>>>(a, static_cast<float*>(out));
    else
      ingest_kernel<float, 3><<<grid, 256, 0, st>>>(a, static_cast<float*>(out));
  }
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

int launch_repack_nchw(const float* x, int C, int H, int W, int B, void* out, int elt,
                       cudaStream_t st) {
  const int Hp = H + kFrameMarginTop + kFrameMarginBottom;
  const int Wp = W + kFrameMarginLeft + kFrameMarginRight;
  const long long total = static_cast<long long>(B) * Hp * (Wp / 4);
  const unsigned grid = static_cast<unsigned>((total + 255) / 256);
  if (elt == FRAME_F16)
    repack_nchw_kernel<__half><<<grid, 256, 0, st>>>(x, C, H, W, B, 1, static_cast<__half*>(out));
  else if (elt == FRAME_BF16)
    repack_nchw_kernel<__nv_bfloat16>
        <<<grid, 256, 0, st>>>(x, C, H, W, B, 1, static_cast<__nv_bfloat16*>(out));
  else
    repack_nchw_kernel<float><<<grid, 256, 0, st>>>(x, C, H, W, B, 0, static_cast<float*>(out));
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

}  // namespace pb200
