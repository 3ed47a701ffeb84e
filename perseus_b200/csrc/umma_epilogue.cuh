// Shared epilogue of the TMA-im2col tcgen05 convolutions (conv_umma.cu, conv_umma2.cu): TMEM accumulator ->
// + bias (+ residual) -> ReLU -> bf16 (fp16 in the "fp16x2" mode, F16 = true) NHWC, with every global access of a warp
// made of COMPLETE 128-byte lines.
//
// An epilogue thread owns one output pixel (TMEM lane) and kCols consecutive channels, i.e. kCols/8 16-byte chunks of
// that pixel's NHWC row.  Loading / storing those chunks directly makes each instruction of a warp touch 32 different
// lines with 16 bytes each (4x the L2 transactions, and the residual read fetches twice the sectors it uses).  Instead,
// inside every group of 8 lanes the 8 chunks of a 64-channel group are transposed with 16-byte shuffles (3 butterfly
// stages), so lane l of the group handles chunk l of the group's 8 pixels: one instruction = 4 complete lines.
#pragma once
#include "common.cuh"

namespace pb200 {

// Output row -> pixel index of the output tensor.  Identity for a normal convolution; for one output parity of a
// stride-2 data gradient (UmmaConvParams::up2) row (n, a, b) of the Ho x Wo tile grid is pixel (2a + ph, 2b + pw) of the
// [B][2Ho][2Wo] tensor (Ho, Wo powers of two).
struct RowMap {
  int up2, ph, pw, lh, lw;
  __device__ __forceinline__ long long operator()(long long gm) const {
    if (!up2) return gm;
    const long long n = gm >> (lh + lw);
    const int a = static_cast<int>(gm >> lw) & ((1 << lh) - 1), b = static_cast<int>(gm) & ((1 << lw) - 1);
    return (((n << (lh + 1)) + 2 * a + ph) << (lw + 1)) + 2 * b + pw;
  }
};

// In-place 8x8 transpose of 16-byte elements across the 8 lanes of a group: before, lane l holds o[j] = element (l, j);
// after, o[j] = element (j, l).  An involution: the same call maps back.
__device__ __forceinline__ void transpose8x8_u4(uint4 (&o)[8], int l8) {
#pragma unroll
  for (int sft = 4; sft >= 1; sft >>= 1) {
    const bool up = (l8 & sft) != 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (j & sft) continue;
      const uint4 send = up ? o[j] : o[j | sft];
      uint4 recv;
      recv.x = __shfl_xor_sync(0xffffffffu, send.x, sft);
      recv.y = __shfl_xor_sync(0xffffffffu, send.y, sft);
      recv.z = __shfl_xor_sync(0xffffffffu, send.z, sft);
      recv.w = __shfl_xor_sync(0xffffffffu, send.w, sft);
      if (up) o[j] = recv;
      else o[j | sft] = recv;
    }
  }
}

// Column sums over the 32 lanes of a warp of N per-lane values (N = 32 or 64) by recursive halving: at each step a lane
// hands half of its columns to its partner and adds the partner's contribution to the half it keeps (N - N/32 shuffles
// instead of 5 N for a butterfly per column).  Afterwards a[0 .. N/32) hold the sums of columns col_of(lane) + i.
template <int N>
__device__ __forceinline__ void warp_column_sums(float (&a)[N], int lane) {
#pragma unroll
  for (int sft = 16, n = N / 2; sft >= 1; sft >>= 1, n >>= 1) {
    const bool up = (lane & sft) != 0;
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
      if (i < n) {
        const float send = up ? a[i] : a[i + n];
        const float recv = __shfl_xor_sync(0xffffffffu, send, sft);
        a[i] = (up ? a[i + n] : a[i]) + recv;
      }
    }
  }
}
template <int N>
__device__ __forceinline__ int warp_column_of(int lane) {  // first column a lane owns after warp_column_sums<N>
  int c = 0;
#pragma unroll
  for (int sft = 16, n = N / 2; sft >= 1; sft >>= 1, n >>= 1) c += (lane & sft) ? n : 0;
  return c;
}

// Batch-statistics BatchNorm fused into the conv epilogue (training forward): per-channel sum and sum of squares of
// the bf16-ROUNDED outputs (what the separate bn_stats pass used to read back) of one 8-chunk group of this warp's 32
// pixels, added to the CTA's shared-memory accumulators s_stat[0..C) / s_stat[C..2C).
__device__ __forceinline__ void accumulate_stats64(const uint4 (&o)[8], bool row_valid, float* s_stat, int C, int col,
                                                   int lane) {
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
    float a[64];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const uint32_t w[4] = {o[c].x, o[c].y, o[c].z, o[c].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float lo = row_valid ? bf16_lo(w[e]) : 0.f, hi = row_valid ? bf16_hi(w[e]) : 0.f;
        a[c * 8 + 2 * e] = pass ? lo * lo : lo;
        a[c * 8 + 2 * e + 1] = pass ? hi * hi : hi;
      }
    }
    warp_column_sums<64>(a, lane);
    const int cc = col + warp_column_of<64>(lane);
    atomicAdd(&s_stat[pass * C + cc], a[0]);
    atomicAdd(&s_stat[pass * C + cc + 1], a[1]);
  }
}

// the same for a 32-column slice (BLOCK_N = 64 kernels: 4 chunks per thread, one column per lane afterwards)
__device__ __forceinline__ void accumulate_stats32(const uint4 (&o)[4], bool row_valid, float* s_stat, int C, int col,
                                                   int lane) {
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
    float a[32];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const uint32_t w[4] = {o[c].x, o[c].y, o[c].z, o[c].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float lo = row_valid ? bf16_lo(w[e]) : 0.f, hi = row_valid ? bf16_hi(w[e]) : 0.f;
        a[c * 8 + 2 * e] = pass ? lo * lo : lo;
        a[c * 8 + 2 * e + 1] = pass ? hi * hi : hi;
      }
    }
    warp_column_sums<32>(a, lane);
    atomicAdd(&s_stat[pass * C + col + warp_column_of<32>(lane)], a[0]);
  }
}

// Residual rows of this thread's tile slice, fetched line by line and transposed back to pixel ownership:
// rv[g * 8 + c] = chunk c of 64-channel group g of THIS lane's pixel.  gm_group = global pixel index of lane (lane & 24)
// of the warp; rows >= M_total read as zero.
template <int kCols, bool COHERENT = false>
__device__ __forceinline__ void load_residual_lines(const __nv_bfloat16* __restrict__ residual, long long gm_group,
                                                    long long M_total, int Cout, int col0, int lane,
                                                    uint4 (&rv)[kCols / 8], const RowMap rmap = RowMap{0, 0, 0, 0, 0}) {
  static_assert(kCols % 64 == 0, "line-granular epilogue needs 64-channel groups");
  const int l8 = lane & 7;
  const long long gbase = gm_group + (lane & 24);
#pragma unroll
  for (int g = 0; g < kCols / 64; ++g) {
    uint4 t[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const long long gmj = gbase + j;
      const uint4* rp = reinterpret_cast<const uint4*>(residual + rmap(gmj) * Cout + col0 + g * 64) + l8;
      t[j] = gmj < M_total ? (COHERENT ? __ldcg(rp) : __ldg(rp)) : make_uint4(0, 0, 0, 0);
    }
    transpose8x8_u4(t, l8);
#pragma unroll
    for (int c = 0; c < 8; ++c) rv[g * 8 + c] = t[c];
  }
}

// One tile slice: kCols accumulator columns of this thread's pixel.  `release()` is called once the accumulator has been
// read completely (hands the TMEM stage back to the MMA warp).
template <int kCols, bool F16 = false, typename Release>
__device__ __forceinline__ void epilogue_lines(uint32_t t_addr, const float* __restrict__ bias, bool has_res,
                                               const uint4 (&rv)[kCols / 8], __nv_bfloat16* __restrict__ out,
                                               long long gm_group, long long M_total, int Cout, int col0, int relu,
                                               int lane, Release release, float* s_stat = nullptr,
                                               const RowMap rmap = RowMap{0, 0, 0, 0, 0}) {
  static_assert(kCols % 64 == 0, "line-granular epilogue needs 64-channel groups");
  const int l8 = lane & 7;
  const long long gbase = gm_group + (lane & 24);
#pragma unroll
  for (int g = 0; g < kCols / 64; ++g) {
    uint4 o[8];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int chunk = 2 * g + h;
      uint32_t v[32];
      tmem_ld_32x32(t_addr + chunk * 32, v);
      tmem_ld_wait_regs(v);
      if (chunk == kCols / 32 - 1) release();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[8 * j + e]) + bias[chunk * 32 + 8 * j + e];
        if (has_res) {
          const uint4 r4 = rv[chunk * 4 + j];
          f[0] += lo16<F16>(r4.x); f[1] += hi16<F16>(r4.x);
          f[2] += lo16<F16>(r4.y); f[3] += hi16<F16>(r4.y);
          f[4] += lo16<F16>(r4.z); f[5] += hi16<F16>(r4.z);
          f[6] += lo16<F16>(r4.w); f[7] += hi16<F16>(r4.w);
        }
        if (relu) {
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
        }
        o[h * 4 + j] = make_uint4(pack16x2<F16>(f[0], f[1]), pack16x2<F16>(f[2], f[3]), pack16x2<F16>(f[4], f[5]),
                                  pack16x2<F16>(f[6], f[7]));
      }
    }
    if (s_stat != nullptr) accumulate_stats64(o, gm_group + lane < M_total, s_stat, Cout, col0 + g * 64, lane);
    transpose8x8_u4(o, l8);  // o[j] = chunk l8 of the group's pixel j
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const long long gmj = gbase + j;
      if (gmj < M_total) reinterpret_cast<uint4*>(out + rmap(gmj) * Cout + col0 + g * 64)[l8] = o[j];
    }
  }
}

}  // namespace pb200
