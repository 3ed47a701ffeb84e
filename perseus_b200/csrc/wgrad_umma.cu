// Weight gradient of a convolution as a TMA-fed tcgen05 GEMM with MN-major operands (sm_100a).
// Replaces cuDNN's backward-filter of the data-parallel training step (perseus/detector/train.py:300,
// loss.backward() through torchvision BasicBlock / conv1, resnet.py:89-105,268).
//
//   dW[tap][ci][co] = sum_m x[m shifted by tap][ci] * dz[m][co]        (m = output pixel, K dimension = pixels)
//
//   * Both operands are "MN-major": in the NHWC tensors the GEMM's M (ci) and N (co) indices are the contiguous
//     ones and the reduction index (pixel) is strided.  The SAME 4-D TMA boxes the forward kernel uses for its A
//     operand (128 output pixels x 64 channels, shifted by the filter tap, zero-filled outside the image, four
//     parity views for stride 2) land 128 rows of 128 bytes (SWIZZLE_128B) in shared memory; read through an
//     MN-major descriptor (SBO = 1024 B between 8-pixel groups, LBO = 16 KB between 64-channel groups) they are the
//     transposed operand, so no transposed copy of any activation or gradient is ever written.
//   * M = 128 rows of one MMA = TWO (tap, 64-channel block) units of x side by side (two boxes 16 KB apart), so the
//     dz tile is shared by two taps; N = up to 256 output channels (dz boxes of 64 channels each).
//   * split-K over the pixel dimension: a work item = (unit pair, N tile, K slice); partial sums are added into the
//     fp32 scratch gradient [tap][Cin][Cout] with 16-byte vector atomics (pb200_train_* unpacks it to OIHW).
//   * warp roles as in conv_umma.cu: TMA producer warp, MMA warp (+ TMEM owner), 8 epilogue warps; persistent CTAs.
#include <stdlib.h>

#include "common.cuh"
#include "tma_host.h"
#include "wgrad_umma.cuh"

namespace pb200 {

constexpr int kWgThreads = 320;  // TMA warp + MMA warp + 8 epilogue warps
constexpr int kWgKBlock = 128;   // pixels per k-block (= one forward M tile)

template <int BLOCK_N>
struct WgCfg {
  static constexpr int kABytes = 2 * 16384;                 // two (tap, 64-ch) units x 128 pixels x 128 B
  static constexpr int kBBytes = BLOCK_N / 64 * 16384;      // BLOCK_N/64 groups of 64 output channels
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = 192 * 1024 / kStageBytes;  // 4 / 3 / 2 for N = 64 / 128 / 256
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024;
  static constexpr int kTmemCols = 2 * BLOCK_N;
};

// kind::f16 instruction descriptor, D = f32, A = B = bf16, BOTH operands MN-major (bits 15 / 16)
__host__ __device__ constexpr uint32_t umma_idesc_bf16_mn(uint32_t M, uint32_t N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// work item -> (unit pair, N tile, K slice).  The K slice is the SLOWEST index: the ~148 items in flight at any moment
// are all (pair, N tile) combinations of a few pixel slices, so the x and dz rows of a slice are fetched from HBM once
// and served to the other pairs / taps from L2 (with the pair slowest every pair streamed both tensors from HBM again).
struct WgItem {
  int pair, n_tile, kb0, kb1;
};
__device__ __forceinline__ WgItem wg_item(const WgradParams& p, int it) {
  const int base = p.num_pairs * p.n_tiles;
  const int ks = it / base;
  const int rem = it - ks * base;
  WgItem w;
  w.pair = rem / p.n_tiles;
  w.n_tile = rem - w.pair * p.n_tiles;
  w.kb0 = static_cast<int>(static_cast<long long>(p.kblocks_total) * ks / p.ksplit);
  w.kb1 = static_cast<int>(static_cast<long long>(p.kblocks_total) * (ks + 1) / p.ksplit);
  return w;
}

template <int BLOCK_N>
__global__ void __launch_bounds__(kWgThreads, 1)
wgrad_umma_kernel(const __grid_constant__ WgradMaps maps, const WgradParams p) {
  using Cfg = WgCfg<BLOCK_N>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kStages];
  __shared__ __align__(8) uint64_t empty_bar[kStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_slot;

  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* tiles = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.x[0]);
    tma_prefetch_desc(&maps.dz);
#pragma unroll
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(&tmem_full_bar[0], 1);
    mbar_init(&tmem_full_bar[1], 1);
    mbar_init(&tmem_empty_bar[0], 8);
    mbar_init(&tmem_empty_bar[1], 8);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(&tmem_base_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

  if (warp == 0) {
    // ================================ TMA producer ================================
    int stage = 0;
    uint32_t phase = 0;
    for (int it = blockIdx.x; it < p.num_items; it += gridDim.x) {
      const WgItem w = wg_item(p, it);
      const int pair = w.pair, n_tile = w.n_tile, kb0 = w.kb0, kb1 = w.kb1;
      // the two units of this pair (the second one of an odd tail repeats the first and is discarded)
      int u_tap[2], u_cb[2];
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        int u = 2 * pair + g;
        if (u >= p.num_units) u = 2 * pair;
        u_tap[g] = u / p.cblocks;
        u_cb[g] = u - u_tap[g] * p.cblocks;
      }
      int dh[2], dw[2], mi[2];
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const int r = u_tap[g] / p.S, s = u_tap[g] - r * p.S;
        dh[g] = r - p.pad;
        dw[g] = s - p.pad;
        mi[g] = 0;
        if (p.stride == 2) {
          const int hp = dh[g] & 1, wp = dw[g] & 1;
          dh[g] = (dh[g] - hp) >> 1;
          dw[g] = (dw[g] - wp) >> 1;
          mi[g] = hp * 2 + wp;
        }
      }
      for (int kb = kb0; kb < kb1; ++kb) {
        int n0, ho0;
        if (p.tiles_per_img > 0) {
          n0 = kb / p.tiles_per_img;
          ho0 = (kb - n0 * p.tiles_per_img) * p.bh;
        } else {
          n0 = kb * p.imgs_per_tile;
          ho0 = 0;
        }
        mbar_wait(&empty_bar[stage], phase ^ 1u);
        if (elect_one()) {
          uint8_t* sa = tiles + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
#pragma unroll
          for (int g = 0; g < 2; ++g)
            tma_load_4d(sa + g * 16384, &maps.x[mi[g]], &full_bar[stage], u_cb[g] * 64, dw[g], ho0 + dh[g], n0);
#pragma unroll
          for (int j = 0; j < BLOCK_N / 64; ++j)
            tma_load_2d(sb + j * 16384, &maps.dz, &full_bar[stage], n_tile * BLOCK_N + j * 64, kb * kWgKBlock);
        }
        __syncwarp();
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    constexpr uint32_t idesc = umma_idesc_bf16_mn(128, BLOCK_N);
    const uint32_t tm = warp_uniform(tmem_base);
    // MN-major SWIZZLE_128B: SBO = 1024 B between 8-pixel (K) groups, LBO = 16 KB between 64-channel (M/N) groups
    const uint32_t d_hi = desc_hi_swizzled(128);
    int stage = 0;
    uint32_t phase = 0;
    int n = 0;
    for (int it = blockIdx.x; it < p.num_items; it += gridDim.x, ++n) {
      const WgItem w = wg_item(p, it);
      const int kb0 = w.kb0, kb1 = w.kb1;
      const int acc = n & 1;
      mbar_wait(&tmem_empty_bar[acc], ((n >> 1) & 1u) ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tm + acc * BLOCK_N;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(tiles + stage * Cfg::kStageBytes);
          const uint32_t a_lo = desc_lo(sa, 16384);
          const uint32_t b_lo = desc_lo(sa + Cfg::kABytes, 16384);
#pragma unroll
          for (int k = 0; k < kWgKBlock / 16; ++k)
            // 16 pixels further along K = 16 rows of 128 B = 2048 B: +128 in the (address >> 4) field
            umma_bf16_split(d_tmem, a_lo + 128 * k, d_hi, b_lo + 128 * k, d_hi, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          if (kb == kb1 - 1) umma_commit(&tmem_full_bar[acc]);
        }
        __syncwarp();
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1u;
        }
      }
      if (kb1 == kb0) {  // empty K slice (never produced by the host): keep the pipeline state consistent
        if (elect_one()) umma_commit(&tmem_full_bar[acc]);
        __syncwarp();
      }
    }
  } else {
    // ================================ epilogue ====================================
    // thread = one (tap, ci) row of the pair x BLOCK_N/2 output channels; partial sums -> fp32 scratch with atomics
    constexpr int kCols = BLOCK_N / 2;
    constexpr int kChunks = kCols / 32;
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = q * 32 + lane;  // 0..127: unit (row / 64), channel row % 64
    int n = 0;
    for (int it = blockIdx.x; it < p.num_items; it += gridDim.x, ++n) {
      const WgItem w = wg_item(p, it);
      const int pair = w.pair, n_tile = w.n_tile, kb0 = w.kb0, kb1 = w.kb1;
      const int acc = n & 1;
      const int u = 2 * pair + (row >> 6);
      const bool valid = u < p.num_units && kb1 > kb0;
      const int tap = u / p.cblocks;
      const int ci = (u - tap * p.cblocks) * 64 + (row & 63);
      const int col0 = n_tile * BLOCK_N + half * kCols;
      float* dst = p.dw + (static_cast<long long>(tap) * p.Cin + ci) * p.Cout + col0;
      mbar_wait(&tmem_full_bar[acc], (n >> 1) & 1u);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BLOCK_N + half * kCols;
#pragma unroll
      for (int chunk = 0; chunk < kChunks; ++chunk) {
        uint32_t v[32];
        tmem_ld_32x32(t_addr + chunk * 32, v);
        tmem_ld_wait_regs(v);
        if (chunk == kChunks - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
        }
        if (valid) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            atomicAdd(reinterpret_cast<float4*>(dst + chunk * 32 + 4 * j),
                      make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                  __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3])));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
template <int BLOCK_N>
static int wg_set_attr() {
  PB_CUDA(cudaFuncSetAttribute(wgrad_umma_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               WgCfg<BLOCK_N>::kSmemBytes));
  return PB200_OK;
}

int wgrad_umma_init() {
  static bool done[kMaxDevices] = {};  // function attributes belong to the device's context
  const int dev = current_device();
  if (done[dev]) return PB200_OK;
  if (int rc = wg_set_attr<64>()) return rc;
  if (int rc = wg_set_attr<128>()) return rc;
  if (int rc = wg_set_attr<256>()) return rc;
  done[dev] = true;
  return PB200_OK;
}

int build_wgrad(WgradLaunch* L, const void* x, const void* dz, float* dw_scratch, int B, int Hin, int Win, int Cin,
                int Cout, int R, int stride, int pad) {
  PB_CHECK(stride == 1 || stride == 2, "wgrad: stride %d unsupported", stride);
  PB_CHECK(Cin % 64 == 0 && Cout % 64 == 0, "wgrad: Cin %d / Cout %d not multiples of 64", Cin, Cout);
  const int Ho = (Hin + 2 * pad - R) / stride + 1;
  const int Wo = (Win + 2 * pad - R) / stride + 1;
  PB_CHECK(stride == 1 || (Hin % 2 == 0 && Win % 2 == 0), "wgrad: stride 2 needs even H, W");
  PB_CHECK(Wo <= kWgKBlock && kWgKBlock % Wo == 0, "wgrad: Wo=%d must divide 128", Wo);
  const int P = Ho * Wo;
  int bh, imgs, tiles_per_img;
  if (P >= kWgKBlock) {
    PB_CHECK(P % kWgKBlock == 0, "wgrad: Ho*Wo=%d must be a multiple of 128", P);
    bh = kWgKBlock / Wo;
    imgs = 1;
    tiles_per_img = P / kWgKBlock;
  } else {
    PB_CHECK(kWgKBlock % P == 0, "wgrad: Ho*Wo=%d must divide 128", P);
    bh = Ho;
    imgs = kWgKBlock / P;
    tiles_per_img = 0;
  }
  const long long M = static_cast<long long>(B) * P;
  memset(L, 0, sizeof(*L));
  WgradParams& p = L->p;
  const int block_n = Cout % 256 == 0 ? 256 : (Cout % 128 == 0 ? 128 : 64);
  L->block_n = block_n;
  p.S = R;
  p.pad = pad;
  p.stride = stride;
  p.tiles_per_img = tiles_per_img;
  p.imgs_per_tile = imgs;
  p.bh = bh;
  p.Cin = Cin;
  p.Cout = Cout;
  p.cblocks = Cin / 64;
  p.num_units = R * R * p.cblocks;
  p.num_pairs = (p.num_units + 1) / 2;
  p.n_tiles = Cout / block_n;
  p.kblocks_total = static_cast<int>((M + kWgKBlock - 1) / kWgKBlock);
  // split-K: at most two work items per SM (the second one's main loop overlaps the first one's atomics), each with
  // >= 4 k-blocks, and the count that fills the last wave best (rounding the count UP to two per SM left a third,
  // nearly empty wave: 300 items on 148 SMs cost three rounds instead of two)
  const int base = p.num_pairs * p.n_tiles;
  const int sms = num_sms();
  int ks_max = p.kblocks_total / 4;
  if (ks_max < 1) ks_max = 1;
  int ks = 1;
  double best = -1.0;
  for (int k = 1; k <= ks_max; ++k) {
    const long long items = static_cast<long long>(base) * k;
    const long long rounds = (items + sms - 1) / sms;
    if (rounds > 2) break;
    const double fill = static_cast<double>(items) / static_cast<double>(rounds * sms);
    if (fill >= best) {
      best = fill;
      ks = k;
    }
  }
  p.ksplit = ks;
  p.num_items = base * ks;
  p.dw = dw_scratch;
  L->grid = p.num_items < num_sms() ? p.num_items : num_sms();
  L->flops = 2.0 * static_cast<double>(M) * Cout * R * R * Cin;

  const uint32_t box[4] = {64u, static_cast<uint32_t>(Wo), static_cast<uint32_t>(bh), static_cast<uint32_t>(imgs)};
  const uint64_t e = 2;
  if (stride == 1) {
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)Win, (uint64_t)Hin, (uint64_t)B};
    const uint64_t str[3] = {Cin * e, (uint64_t)Win * Cin * e, (uint64_t)Hin * Win * Cin * e};
    if (int rc = encode_tma_bf16(&L->maps.x[0], x, 4, dims, str, box, 128)) return rc;
    for (int i = 1; i < 4; ++i) L->maps.x[i] = L->maps.x[0];
  } else {
    for (int hp = 0; hp < 2; ++hp)
      for (int wp = 0; wp < 2; ++wp) {
        const uint8_t* base_ptr = static_cast<const uint8_t*>(x) + (static_cast<uint64_t>(hp) * Win + wp) * Cin * e;
        const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)Win / 2, (uint64_t)Hin / 2, (uint64_t)B};
        const uint64_t str[3] = {2 * Cin * e, 2 * (uint64_t)Win * Cin * e, (uint64_t)Hin * Win * Cin * e};
        if (int rc = encode_tma_bf16(&L->maps.x[hp * 2 + wp], base_ptr, 4, dims, str, box, 128)) return rc;
      }
  }
  const uint64_t zdims[2] = {(uint64_t)Cout, (uint64_t)M};
  const uint64_t zstr[1] = {(uint64_t)Cout * e};
  const uint32_t zbox[2] = {64u, static_cast<uint32_t>(kWgKBlock)};
  return encode_tma_bf16(&L->maps.dz, dz, 2, zdims, zstr, zbox, 128);
}

template <int BLOCK_N>
static int wg_launch_one(const WgradLaunch& L, cudaStream_t st) {
  wgrad_umma_kernel<BLOCK_N><<<L.grid, kWgThreads, WgCfg<BLOCK_N>::kSmemBytes, st>>>(L.maps, L.p);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

int launch_wgrad(const WgradLaunch& L, cudaStream_t st) {
  if (L.p.num_items == 0) return PB200_OK;
  if (L.block_n == 64) return wg_launch_one<64>(L, st);
  if (L.block_n == 128) return wg_launch_one<128>(L, st);
  return wg_launch_one<256>(L, st);
}

}  // namespace pb200
