// Memory-bound tail kernels of the detector: 3x3/2 max pooling (NHWC), the regression head
// (global average pool + Linear + pixel decode, warp-shuffle reductions, fp32), and the
// stand-alone coordinate decode.
//
// Replaces torchvision ResNet.maxpool / avgpool / flatten / fc (resnet.py:271,278-280 via
// perseus/detector/models.py:31-32) and kornia.geometry.denormalize_pixel_coordinates
// (validate.py:144-153, streaming.py:129-131): u = (x+1)(W-1)/2, v = (y+1)(H-1)/2.
#include "common.cuh"
#include "layout.h"

namespace pb200 {

// kornia computes factor = 2/(size-1) (clamped at 1e-8) and returns (1/factor)*(x+1), all in the
// tensor's dtype; reproduce those fp32 roundings on the host.
static float denorm_scale(int size) {
  volatile float f = 2.0f / fmaxf(static_cast<float>(size - 1), 1e-8f);
  volatile float inv = 1.0f / f;
  return inv;
}

// ---- maxpool 3x3 s2 p1 ---------------------------------------------------------------------
// one thread = one output pixel x 8 channels (bf16, one 128-bit load per tap)
__global__ void __launch_bounds__(256)
maxpool_bf16_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, int B, int H, int W,
                    int C8, int Ho, int Wo) {
  const long long total = static_cast<long long>(B) * Ho * Wo * C8;
  const long long gid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int c = static_cast<int>(gid % C8);
  long long p = gid / C8;
  const int wo = static_cast<int>(p % Wo);
  p /= Wo;
  const int ho = static_cast<int>(p % Ho);
  const int b = static_cast<int>(p / Ho);
  const __nv_bfloat162 ninf = __float2bfloat162_rn(-INFINITY);
  __nv_bfloat162 m[4] = {ninf, ninf, ninf, ninf};
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int hi = 2 * ho + r - 1;
    if (hi < 0 || hi >= H) continue;
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int wi = 2 * wo + s - 1;
      if (wi < 0 || wi >= W) continue;
      const uint4 v = __ldg(in + ((static_cast<long long>(b) * H + hi) * W + wi) * C8 + c);
      const __nv_bfloat162* pv = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int k = 0; k < 4; ++k) m[k] = __hmax2(m[k], pv[k]);
    }
  }
  uint4 o;
  __nv_bfloat162* po = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int k = 0; k < 4; ++k) po[k] = m[k];
  out[gid] = o;
}

__global__ void __launch_bounds__(256)
maxpool_f32_kernel(const float4* __restrict__ in, float4* __restrict__ out, int B, int H, int W,
                   int C4, int Ho, int Wo) {
  const long long total = static_cast<long long>(B) * Ho * Wo * C4;
  const long long gid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int c = static_cast<int>(gid % C4);
  long long p = gid / C4;
  const int wo = static_cast<int>(p % Wo);
  p /= Wo;
  const int ho = static_cast<int>(p % Ho);
  const int b = static_cast<int>(p / Ho);
  float4 m = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int hi = 2 * ho + r - 1;
    if (hi < 0 || hi >= H) continue;
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int wi = 2 * wo + s - 1;
      if (wi < 0 || wi >= W) continue;
      const float4 v = __ldg(in + ((static_cast<long long>(b) * H + hi) * W + wi) * C4 + c);
      m.x = fmaxf(m.x, v.x); m.y = fmaxf(m.y, v.y); m.z = fmaxf(m.z, v.z); m.w = fmaxf(m.w, v.w);
    }
  }
  out[gid] = m;
}

// split-bf16 ("bf16x3") mode: fp32 NHWC in, [hi(C) | lo(C)] bf16 NHWC out (hi = bf16(m), lo = bf16(m - hi))
__global__ void __launch_bounds__(256)
maxpool_split_kernel(const float4* __restrict__ in, uint2* __restrict__ out, int B, int H, int W,
                     int C4, int Ho, int Wo) {
  const long long total = static_cast<long long>(B) * Ho * Wo * C4;
  const long long gid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int c = static_cast<int>(gid % C4);
  long long p = gid / C4;
  const long long pix = p;
  const int wo = static_cast<int>(p % Wo);
  p /= Wo;
  const int ho = static_cast<int>(p % Ho);
  const int b = static_cast<int>(p / Ho);
  float4 m = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int hi = 2 * ho + r - 1;
    if (hi < 0 || hi >= H) continue;
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int wi = 2 * wo + s - 1;
      if (wi < 0 || wi >= W) continue;
      const float4 v = __ldg(in + ((static_cast<long long>(b) * H + hi) * W + wi) * C4 + c);
      m.x = fmaxf(m.x, v.x); m.y = fmaxf(m.y, v.y); m.z = fmaxf(m.z, v.z); m.w = fmaxf(m.w, v.w);
    }
  }
  const uint32_t h0 = pack_bf16x2(m.x, m.y), h1 = pack_bf16x2(m.z, m.w);
  const uint32_t l0 = pack_bf16x2(m.x - bf16_lo(h0), m.y - bf16_hi(h0));
  const uint32_t l1 = pack_bf16x2(m.z - bf16_lo(h1), m.w - bf16_hi(h1));
  out[pix * (2 * C4) + c] = make_uint2(h0, h1);
  out[pix * (2 * C4) + C4 + c] = make_uint2(l0, l1);
}

int launch_maxpool_split(const void* in, void* out, int B, int H, int W, int C, cudaStream_t st) {
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  PB_CHECK(C % 4 == 0, "maxpool: C %% 4 != 0");
  const long long total = static_cast<long long>(B) * Ho * Wo * (C / 4);
  maxpool_split_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(
      static_cast<const float4*>(in), static_cast<uint2*>(out), B, H, W, C / 4, Ho, Wo);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

int launch_maxpool(const void* in, void* out, int B, int H, int W, int C, bool bf16, cudaStream_t st) {
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  if (bf16) {
    PB_CHECK(C % 8 == 0, "maxpool: C %% 8 != 0");
    const long long total = static_cast<long long>(B) * Ho * Wo * (C / 8);
    maxpool_bf16_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(
        static_cast<const uint4*>(in), static_cast<uint4*>(out), B, H, W, C / 8, Ho, Wo);
  } else {
    PB_CHECK(C % 4 == 0, "maxpool: C %% 4 != 0");
    const long long total = static_cast<long long>(B) * Ho * Wo * (C / 4);
    maxpool_f32_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(
        static_cast<const float4*>(in), static_cast<float4*>(out), B, H, W, C / 4, Ho, Wo);
  }
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

// ---- head: avgpool + fc + decode -----------------------------------------------------------
// One CTA per frame, 256 threads.  Phase 1: thread t owns channels {2t, 2t+1 (+512k)} and sums
// them over the HW pixels with coalesced 4-byte (bf16x2) / 8-byte (float2) loads; the means go to
// shared memory.  Phase 2: each warp produces outputs o = warp, warp+8, ...: lanes stride over
// the channels, then a 5-step shuffle tree.  Everything after the loads is fp32.
template <int MODE>  // 0: fp32 input, 1: bf16, 2: split bf16 rows [hi(C) | lo(C)], 3: fp16
__global__ void __launch_bounds__(256)
head_kernel(const void* __restrict__ in, const float* __restrict__ fc_w,
            const float* __restrict__ fc_b, float* __restrict__ out_norm, float* __restrict__ out_px,
            int HW, int C, int n_out, float half_wm1, float half_hm1) {
  extern __shared__ float s_mean[];  // C floats
  pdl_wait();  // launched with programmatic serialization: the last conv's output must be complete
  const int b = blockIdx.x;
  const float inv = 1.0f / static_cast<float>(HW);
  // 16-bit activations, C = 512, 64 pixels (the detector's shape): thread = (pixel group g of 4, 8-channel chunk): 16
  // independent 16-byte loads per thread, all in flight at once, then the four groups' partial sums meet in shared
  // memory.  (One thread per channel pair walked its 64 pixels with 4-byte loads, 8 in flight: ~8 dependent round trips.)
  __shared__ float s_part[(MODE == 1 || MODE == 3) ? 4 * 512 : 1];
  const bool fast = (MODE == 1 || MODE == 3) && C == 512 && HW % 4 == 0 && blockDim.x == 256;
  if (fast) {
    if constexpr (MODE == 1 || MODE == 3) {
      const int g = threadIdx.x >> 6, ch = threadIdx.x & 63;  // 64 chunks of 8 channels
      const int per = HW / 4;
      const uint4* p = static_cast<const uint4*>(in) + (static_cast<long long>(b) * HW + g * per) * (C / 8) + ch;
      float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 16
      for (int i = 0; i < per; ++i) {
        const uint4 v = __ldg(p + static_cast<long long>(i) * (C / 8));
        a[0] += lo16<MODE == 3>(v.x); a[1] += hi16<MODE == 3>(v.x);
        a[2] += lo16<MODE == 3>(v.y); a[3] += hi16<MODE == 3>(v.y);
        a[4] += lo16<MODE == 3>(v.z); a[5] += hi16<MODE == 3>(v.z);
        a[6] += lo16<MODE == 3>(v.w); a[7] += hi16<MODE == 3>(v.w);
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) s_part[g * 512 + ch * 8 + e] = a[e];
      __syncthreads();
      for (int c = threadIdx.x; c < C; c += blockDim.x)
        s_mean[c] = ((s_part[c] + s_part[512 + c]) + (s_part[1024 + c] + s_part[1536 + c])) * inv;
    }
  }
  for (int c2 = threadIdx.x; !fast && c2 < C / 2; c2 += blockDim.x) {
    float a0 = 0.f, a1 = 0.f;
    if (MODE == 2) {
      const uint32_t* p = static_cast<const uint32_t*>(in) + static_cast<long long>(b) * HW * C + c2;
#pragma unroll 8
      for (int i = 0; i < HW; ++i) {
        const uint32_t h = __ldg(p + static_cast<long long>(i) * C);
        const uint32_t l = __ldg(p + static_cast<long long>(i) * C + C / 2);
        a0 += bf16_lo(h) + bf16_lo(l);
        a1 += bf16_hi(h) + bf16_hi(l);
      }
    } else if (MODE == 1 || MODE == 3) {
      const uint32_t* p = static_cast<const uint32_t*>(in) + static_cast<long long>(b) * HW * (C / 2) + c2;
#pragma unroll 8
      for (int i = 0; i < HW; ++i) {
        const uint32_t v = __ldg(p + static_cast<long long>(i) * (C / 2));
        a0 += lo16<MODE == 3>(v);
        a1 += hi16<MODE == 3>(v);
      }
    } else {
      const float2* p = static_cast<const float2*>(in) + static_cast<long long>(b) * HW * (C / 2) + c2;
#pragma unroll 8
      for (int i = 0; i < HW; ++i) {
        const float2 v = __ldg(p + static_cast<long long>(i) * (C / 2));
        a0 += v.x;
        a1 += v.y;
      }
    }
    s_mean[2 * c2] = a0 * inv;
    s_mean[2 * c2 + 1] = a1 * inv;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int o = warp; o < n_out; o += (blockDim.x >> 5)) {
    float acc = 0.f;
    for (int c = lane; c < C; c += 32) acc = fmaf(s_mean[c], __ldg(fc_w + o * C + c), acc);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if (lane == 0) {
      const float y = acc + __ldg(fc_b + o);
      out_norm[static_cast<long long>(b) * n_out + o] = y;
      if (out_px != nullptr)
        out_px[static_cast<long long>(b) * n_out + o] = (y + 1.0f) * ((o & 1) ? half_hm1 : half_wm1);
    }
  }
}

int launch_head(const void* in, const float* fc_w, const float* fc_b, float* out_norm, float* out_px,
                int B, int HW, int C, int n_out, int imgH, int imgW, int mode, cudaStream_t st) {
  PB_CHECK(C % 2 == 0, "head: odd channel count");
  const float hw = denorm_scale(imgW), hh = denorm_scale(imgH);
  const size_t smem = sizeof(float) * C;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(B);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  cfg.attrs = attr;
  cfg.numAttrs = pdl_launch_attrs(attr);
  if (mode == 3)
    PB_CUDA(cudaLaunchKernelEx(&cfg, head_kernel<3>, in, fc_w, fc_b, out_norm, out_px, HW, C, n_out, hw, hh));
  else if (mode == 2)
    PB_CUDA(cudaLaunchKernelEx(&cfg, head_kernel<2>, in, fc_w, fc_b, out_norm, out_px, HW, C, n_out, hw, hh));
  else if (mode == 1)
    PB_CUDA(cudaLaunchKernelEx(&cfg, head_kernel<1>, in, fc_w, fc_b, out_norm, out_px, HW, C, n_out, hw, hh));
  else
    PB_CUDA(cudaLaunchKernelEx(&cfg, head_kernel<0>, in, fc_w, fc_b, out_norm, out_px, HW, C, n_out, hw, hh));
  return PB200_OK;
}

__global__ void decode_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long n,
                              float half_wm1, float half_hm1) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float2 v = __ldg(in + i);
  out[i] = make_float2((v.x + 1.0f) * half_wm1, (v.y + 1.0f) * half_hm1);
}

int launch_decode(const float* xy_norm, float* xy_px, long long n_points, int H, int W,
                  cudaStream_t st) {
  if (n_points == 0) return PB200_OK;
  decode_kernel<<<static_cast<unsigned>((n_points + 255) / 256), 256, 0, st>>>(
      reinterpret_cast<const float2*>(xy_norm), reinterpret_cast<float2*>(xy_px), n_points,
      denorm_scale(W), denorm_scale(H));
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

}  // namespace pb200

extern "C" int pb200_decode_pixels(const float* xy_norm, float* xy_px, int64_t n_points, int H, int W,
                                   pb200_stream stream) {
  if (int rc = pb200::require_device()) return rc;
  PB_CHECK(n_points >= 0 && xy_norm && xy_px, "decode_pixels: bad arguments");
  return pb200::launch_decode(xy_norm, xy_px, n_points, H, W, static_cast<cudaStream_t>(stream));
}
