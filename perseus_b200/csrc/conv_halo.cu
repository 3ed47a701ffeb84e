// 3x3 stride-1 convolution (+ folded BN, fused 1x1/2 projection, residual add, ReLU) on NHWC bf16 where the WHOLE packed
// weight matrix fits in the shared memory of a CTA pair: layer2 of the detector's ResNet-18 (128 -> 128 channels at 32x32;
// torchvision BasicBlock.forward, resnet.py:89-105, reached from perseus/detector/models.py:34-40).
//
// The TMA-im2col kernels (conv_umma.cu / conv_umma2.cu) load one 128 x 64 activation tile per (filter tap, 64-channel block):
// every input pixel crosses L2 -> shared memory NINE times per convolution, and on layer2 that traffic (0.6 GB per conv, ~9 TB/s)
// is what the kernel waits for -- making the weight tile resident (RESB) or moving it as 512-byte rows (KT layout) changed nothing
// (DESIGN.md §4, r02).  Here:
//
//   * halo patch: per M tile and 64-channel block ONE 4-D TMA box {64 ch, W + 2 columns from column -1, 7 rows from row h0 - 1}
//     lands the tile's receptive field in shared memory, zero-filled outside the image (the conv's padding), as 128-byte rows in
//     "padded-row" order: position pi = (row - row0) * (W + 2) + (col + 1).  An M tile is 128 CONSECUTIVE padded positions
//     q = h * (W + 2) + w' of one image, so for filter tap (r, s) the A operand is the same patch read from row
//     (q0 - h0 * (W + 2)) + r * (W + 2) + s on: a SWIZZLE_128B K-major descriptor whose start address is shifted by whole 128-byte
//     rows (base offset 0: the swizzle XOR is taken from the absolute shared-memory address, tools/probe_umma_shift.cu).  The two
//     positions per row with w' >= W and the positions past the last row produce garbage accumulators that are never stored
//     (8/9 of the issued MMA rows are real at W = 32).  Activation bytes into shared memory: 61 KB per tile instead of 288 KB.
//   * resident weights: each CTA of the pair keeps its 64 output channels of all 18-19 k-blocks (147-156 KB, KT layout:
//     conv_umma.cuh) for its lifetime; tcgen05.mma.cta_group::2 reads both halves.
//   * a pair takes tile t of images (2j, 2j + 1): both CTAs then use IDENTICAL descriptor offsets, which cta_group::2 requires
//     (one descriptor addresses both CTAs' shared memory).
//   * warps: 0 = producer (patches), 1 = MMA issuer of the pair leader, 2..9 = epilogue (umma_epilogue.cuh, complete 128-byte
//     lines), accumulators double-buffered in TMEM.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "conv_umma.cuh"
#include "tma_host.h"
#include "umma_epilogue.cuh"

namespace pb200 {

namespace {

constexpr int kHaloThreads = 320;
constexpr int kPatchRows = 7;   // rows of the halo box: the tile's <= 5 image rows + one above and below

struct HaloParams {
  int B, H, W, Wp;            // Wp = W + 2
  int Cout;                   // = BLOCK_N = 128
  int cblocks;                // Cin / 64
  int ds_cblocks;             // fused projection: 64-channel blocks of the block input (0 = none)
  int tiles_per_img;          // ceil(H * Wp / 128)
  int num_pair_tiles;         // ceil(B / 2) * tiles_per_img
  int patch_bytes;            // 64 ch * 2 B * Wp * kPatchRows
  int ds_rows;                // rows of the projection's box
  int stage_bytes;            // patch stage size (multiple of 1024)
  int num_kblocks;            // 9 * cblocks + ds_cblocks
  int relu;
  int reverse;                // walk the pair tiles from the last to the first (L2 reuse across launches: launch_conv_p8)
  const float* bias;
  const void* residual;       // NHWC [B][H][W][Cout] or null
  void* out;
};

// output row (tile-local position) -> pixel index of the NHWC output, or -1 for the padded columns / rows past the image
struct PaddedRowMap {
  int H, W, Wp, q0;
  long long img_base;  // n * H * W, or < 0: no image (odd batch: the pair's second image does not exist)
  __device__ __forceinline__ long long operator()(int row) const {
    const int q = q0 + row, h = q / Wp, w = q - h * Wp;
    return (img_base >= 0 && h < H && w < W) ? img_base + h * W + w : -1;
  }
};

__global__ void __launch_bounds__(kHaloThreads, 1)
conv_halo_kernel(const __grid_constant__ UmmaConvMaps maps, const HaloParams p) {
  constexpr int BLOCK_N = 128;
  constexpr int kBHalf = BLOCK_N / 2 * 128;  // this CTA's 64 rows of one k-block
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[2];        // leader only: both CTAs' patches of a stage have landed
  __shared__ __align__(8) uint64_t empty_bar[2];       // per CTA: the stage has been read by the MMAs
  __shared__ __align__(8) uint64_t tmem_full_bar[2];   // per CTA
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];  // leader only: 8 epilogue warps x 2 CTAs
  __shared__ __align__(8) uint64_t resb_bar;           // leader only: both weight halves are resident
  __shared__ uint32_t tmem_base_slot;
  __shared__ float s_bias[BLOCK_N];

  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* tiles = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  uint8_t* smem_a = tiles;                        // 2 patch stages
  uint8_t* smem_b = tiles + 2 * p.stage_bytes;    // resident weights: k-block kb at + kb * kBHalf

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int crank = static_cast<int>(cluster_ctarank());

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&maps.a[0]);
    tma_prefetch_desc(&maps.b);
    mbar_init(&full_bar[0], 2);
    mbar_init(&full_bar[1], 2);
    mbar_init(&empty_bar[0], 1);
    mbar_init(&empty_bar[1], 1);
    mbar_init(&tmem_full_bar[0], 1);
    mbar_init(&tmem_full_bar[1], 1);
    mbar_init(&tmem_empty_bar[0], 16);
    mbar_init(&tmem_empty_bar[1], 16);
    mbar_init(&resb_bar, 2);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc2(&tmem_base_slot, 2 * BLOCK_N);
    tmem_relinquish2();
  }
  if (threadIdx.x < BLOCK_N) s_bias[threadIdx.x] = __ldg(p.bias + threadIdx.x);  // plan constant
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  const int pair_id = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;
  pdl_launch_dependents();

  if (warp == 0) {
    // ================================ producer (both CTAs) ========================
    const uint32_t lfull0 = mapa_u32(&full_bar[0], 0);
    if (pair_id < p.num_pair_tiles && elect_one()) {
      // the weights are plan constants: fetched while the previous kernel drains (before the PDL wait)
      const uint32_t lres = mapa_u32(&resb_bar, 0);
      mbar_expect_tx_cluster(lres, static_cast<uint32_t>(p.num_kblocks) * kBHalf);
      for (int kb = 0; kb < p.num_kblocks; ++kb)
        tma_load_2d_2sm(smem_b + kb * kBHalf, &maps.b, lres, 0, (kb * p.Cout + crank * (BLOCK_N / 2)) >> 2);
    }
    __syncwarp();
    pdl_wait();  // the activations are the previous kernel's output
    int stage = 0;
    uint32_t phase = 0;
    for (int pti = pair_id; pti < p.num_pair_tiles; pti += num_pairs) {
      const int pt = p.reverse ? p.num_pair_tiles - 1 - pti : pti;
      const int jp = pt / p.tiles_per_img;
      const int t = pt - jp * p.tiles_per_img;
      int n = 2 * jp + crank;
      if (n >= p.B) n = p.B - 1;  // odd batch: the peer re-reads the last image (its results are not stored)
      const int h0 = (t * 128) / p.Wp;
      for (int cb = 0; cb < p.cblocks + p.ds_cblocks; ++cb) {
        mbar_wait(&empty_bar[stage], phase ^ 1u);
        if (elect_one()) {
          const uint32_t lfull = lfull0 + stage * 8;
          if (cb < p.cblocks) {
            mbar_expect_tx_cluster(lfull, p.patch_bytes);
            tma_load_4d_2sm(smem_a + stage * p.stage_bytes, &maps.a[0], lfull, cb * 64, -1, h0 - 1, n);
          } else {  // the projection's input pixel (2h, 2w') of the block input, same padded-row order, no halo
            mbar_expect_tx_cluster(lfull, 128 * p.Wp * p.ds_rows);
            tma_load_4d_2sm(smem_a + stage * p.stage_bytes, &maps.ds, lfull, (cb - p.cblocks) * 64, 0, h0, n);
          }
        }
        __syncwarp();
        stage ^= 1;
        if (stage == 0) phase ^= 1u;
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (pair leader only) ===============
    if (crank == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(256, BLOCK_N);
      const uint32_t tm = warp_uniform(tmem_base);
      const uint32_t d_hi = desc_hi_swizzled(128);
      const uint32_t b_lo0 = desc_lo(smem_u32(smem_b), 16);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      if (pair_id < p.num_pair_tiles) mbar_wait(&resb_bar, 0);
      for (int pti = pair_id; pti < p.num_pair_tiles; pti += num_pairs, ++it) {
        const int pt = p.reverse ? p.num_pair_tiles - 1 - pti : pti;
        const int t = pt % p.tiles_per_img;
        const int q0 = t * 128;
        const uint32_t pi0 = static_cast<uint32_t>(q0 - (q0 / p.Wp) * p.Wp);  // first A row of tap (0, 0) in the patch
        const int acc = it & 1;
        mbar_wait(&tmem_empty_bar[acc], ((it >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tm + acc * BLOCK_N;
        for (int cb = 0; cb < p.cblocks + p.ds_cblocks; ++cb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_base = desc_lo(smem_u32(smem_a + stage * p.stage_bytes), 16) + pi0 * 8;  // 128 B = 8 x 16 B
            if (cb < p.cblocks) {
#pragma unroll
              for (int tap = 0; tap < 9; ++tap) {
                const int r = tap / 3, s = tap - 3 * r;
                const uint32_t a_lo = a_base + static_cast<uint32_t>(r * p.Wp + s) * 8;
                const uint32_t b_lo = b_lo0 + static_cast<uint32_t>(tap * p.cblocks + cb) * (kBHalf >> 4);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  umma2_bf16_split(d_tmem, a_lo + 2 * k, d_hi, b_lo + 2 * k, d_hi, idesc, (cb | tap | k) != 0 ? 1u : 0u);
              }
            } else {
              const uint32_t b_lo = b_lo0 + static_cast<uint32_t>(9 * p.cblocks + (cb - p.cblocks)) * (kBHalf >> 4);
#pragma unroll
              for (int k = 0; k < 4; ++k) umma2_bf16_split(d_tmem, a_base + 2 * k, d_hi, b_lo + 2 * k, d_hi, idesc, 1u);
            }
            umma2_commit_mcast(&empty_bar[stage], 3);
            if (cb == p.cblocks + p.ds_cblocks - 1) umma2_commit_mcast(&tmem_full_bar[acc], 3);
          }
          __syncwarp();
          stage ^= 1;
          if (stage == 0) phase ^= 1u;
        }
      }
    }
  } else {
    // ================================ epilogue (both CTAs) =========================
    pdl_wait();  // residual reads / output writes must follow the previous kernel
    constexpr int kCols = BLOCK_N / 2;  // 64 channels per thread
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int l8 = lane & 7;
    const bool has_res = p.residual != nullptr;
    const __nv_bfloat16* res = static_cast<const __nv_bfloat16*>(p.residual);
    __nv_bfloat16* out = static_cast<__nv_bfloat16*>(p.out);
    const int col0 = half * kCols;
    int it = 0;
    for (int pti = pair_id; pti < p.num_pair_tiles; pti += num_pairs, ++it) {
      const int pt = p.reverse ? p.num_pair_tiles - 1 - pti : pti;
      const int jp = pt / p.tiles_per_img;
      const int t = pt - jp * p.tiles_per_img;
      const int n = 2 * jp + crank;
      const PaddedRowMap rm = {p.H, p.W, p.Wp, t * 128, n < p.B ? static_cast<long long>(n) * p.H * p.W : -1};
      const int acc = it & 1;
      // this lane's 8-row group: rows (q * 32 + (lane & 24)) + j; after the 8x8 transposes lane l8 moves chunk l8 of row j
      const int row0 = q * 32 + (lane & 24);
      long long pix[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) pix[j] = rm(row0 + j);
      uint4 rv[8];
      if (has_res) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          rv[j] = pix[j] >= 0 ? __ldg(reinterpret_cast<const uint4*>(res + pix[j] * p.Cout + col0) + l8) : make_uint4(0, 0, 0, 0);
        transpose8x8_u4(rv, l8);  // rv[c] = chunk c of THIS lane's row
      }
      mbar_wait(&tmem_full_bar[acc], (it >> 1) & 1u);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BLOCK_N + col0;
      uint4 o[8];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t v[32];
        tmem_ld_32x32(t_addr + h * 32, v);
        tmem_ld_wait_regs(v);
        if (h == 1) {  // accumulator stage fully read by this warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (crank == 0) mbar_arrive(&tmem_empty_bar[acc]);
            else mbar_arrive_cluster(mapa_u32(&tmem_empty_bar[acc], 0));
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[8 * j + e]) + s_bias[col0 + h * 32 + 8 * j + e];
          if (has_res) {
            const uint4 r4 = rv[h * 4 + j];
            f[0] += bf16_lo(r4.x); f[1] += bf16_hi(r4.x);
            f[2] += bf16_lo(r4.y); f[3] += bf16_hi(r4.y);
            f[4] += bf16_lo(r4.z); f[5] += bf16_hi(r4.z);
            f[6] += bf16_lo(r4.w); f[7] += bf16_hi(r4.w);
          }
          if (p.relu) {
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
          }
          o[h * 4 + j] = make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                                    pack_bf16x2(f[6], f[7]));
        }
      }
      transpose8x8_u4(o, l8);  // o[j] = chunk l8 of the group's row j: one store instruction = 4 complete 128-byte lines
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (pix[j] >= 0) reinterpret_cast<uint4*>(out + pix[j] * p.Cout + col0)[l8] = o[j];
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer's TMEM / barriers stay alive until the leader's last MMA and commit are done
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, 2 * BLOCK_N);
  }
}

constexpr int kHaloMaxSmem = 222 * 1024;

}  // namespace

bool conv_halo_supported(int B, int H, int W, int Cin, int Cout, int R, int stride, int pad, int ds_Cin) {
  if (R != 3 || stride != 1 || pad != 1 || Cout != 128 || Cin % 64 != 0 || ds_Cin % 64 != 0 || B < 2) return false;
  const int Wp = W + 2;
  if (Wp > 256 || 3 * Wp + 130 > kPatchRows * Wp) return false;  // the tile's nine shifted reads must stay inside the box
  const int nkb = 9 * (Cin / 64) + ds_Cin / 64;
  const int stage = (128 * Wp * kPatchRows + 1023) / 1024 * 1024;
  return 2 * stage + nkb * 64 * 128 + 1024 <= kHaloMaxSmem;
}

int build_conv_halo(HaloLaunch* L, const void* in, int B, int H, int W, int Cin, const void* wpacked_kt, const float* bias,
                    const void* residual, void* out, int Cout, int relu, const void* ds_in, int ds_Cin, int reverse) {
  PB_CHECK(conv_halo_supported(B, H, W, Cin, Cout, 3, 1, 1, ds_in ? ds_Cin : 0), "conv_halo: unsupported shape");
  memset(L, 0, sizeof(*L));
  const int Wp = W + 2;
  HaloParams p = {};
  p.B = B; p.H = H; p.W = W; p.Wp = Wp; p.Cout = Cout;
  p.reverse = reverse ? 1 : 0;
  p.cblocks = Cin / 64;
  p.ds_cblocks = ds_in ? ds_Cin / 64 : 0;
  p.tiles_per_img = (H * Wp + 127) / 128;
  p.num_pair_tiles = (B + 1) / 2 * p.tiles_per_img;
  p.patch_bytes = 128 * Wp * kPatchRows;
  p.ds_rows = (Wp - 1 + 127) / Wp + 1;  // rows touched by 128 consecutive positions starting anywhere in a row
  p.stage_bytes = (p.patch_bytes + 1023) / 1024 * 1024;
  p.num_kblocks = 9 * p.cblocks + p.ds_cblocks;
  p.relu = relu; p.bias = bias; p.residual = residual; p.out = out;
  static_assert(sizeof(HaloParams) <= sizeof(L->params), "HaloLaunch::params too small");
  memcpy(L->params, &p, sizeof(p));
  L->smem = 2 * p.stage_bytes + p.num_kblocks * 64 * 128 + 1024;
  L->flops = 2.0 * B * H * W * Cout * (9.0 * Cin + (ds_in ? ds_Cin : 0));
  // activations: {64 ch, Wp columns, kPatchRows rows, 1 image}, SWIZZLE_128B, zero fill outside the image
  const uint64_t e = 2, C = static_cast<uint64_t>(Cin);
  {
    const uint64_t dims[4] = {C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint64_t str[3] = {C * e, (uint64_t)W * C * e, (uint64_t)H * W * C * e};
    const uint32_t box[4] = {64u, (uint32_t)Wp, (uint32_t)kPatchRows, 1u};
    if (int rc = encode_tma_bf16(&L->maps.a[0], in, 4, dims, str, box, 128)) return rc;
    for (int i = 1; i < 4; ++i) L->maps.a[i] = L->maps.a[0];
  }
  L->maps.ds = L->maps.a[0];
  if (ds_in != nullptr) {  // parity (0, 0) of the [B][2H][2W][ds_Cin] block input, padded-row order (columns 0 .. Wp-1)
    const uint64_t dC = static_cast<uint64_t>(ds_Cin);
    const uint64_t dims[4] = {dC, (uint64_t)W, (uint64_t)H, (uint64_t)B};
    const uint64_t str[3] = {2 * dC * e, 2 * (uint64_t)(2 * W) * dC * e, (uint64_t)(2 * H) * (2 * W) * dC * e};
    const uint32_t box[4] = {64u, (uint32_t)Wp, (uint32_t)p.ds_rows, 1u};
    if (int rc = encode_tma_bf16(&L->maps.ds, ds_in, 4, dims, str, box, 128)) return rc;
  }
  {  // KT-layout weights as 512-byte rows; this CTA's 64 rows of one k-block = 16 rows
    const uint64_t total = static_cast<uint64_t>(p.num_kblocks) * Cout * 128;
    const uint64_t dims[2] = {256, total / 512};
    const uint64_t strides[1] = {512};
    const uint32_t box[2] = {256u, 16u};
    if (int rc = encode_tma_bf16(&L->maps.b, wpacked_kt, 2, dims, strides, box, 0)) return rc;
  }
  // persistent pairs
  static int max_pairs[kMaxDevices] = {};
  const int dev = current_device();
  if (max_pairs[dev] == 0) {
    PB_CUDA(cudaFuncSetAttribute(conv_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kHaloMaxSmem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(num_sms() / 2 * 2);
    cfg.blockDim = dim3(kHaloThreads);
    cfg.dynamicSmemBytes = kHaloMaxSmem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    PB_CUDA(cudaOccupancyMaxActiveClusters(&n, conv_halo_kernel, &cfg));
    PB_CHECK(n > 0, "conv_halo: no resident CTA pair");
    max_pairs[dev] = n;
  }
  L->grid = 2 * (p.num_pair_tiles < max_pairs[dev] ? p.num_pair_tiles : max_pairs[dev]);
  return PB200_OK;
}

int launch_conv_halo(const HaloLaunch& L, cudaStream_t st) {
  HaloParams p;
  memcpy(&p, L.params, sizeof(p));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(L.grid);
  cfg.blockDim = dim3(kHaloThreads);
  cfg.dynamicSmemBytes = L.smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1 + pdl_launch_attrs(attr + 1);
  PB_CUDA(cudaLaunchKernelEx(&cfg, conv_halo_kernel, L.maps, p));
  return PB200_OK;
}

}  // namespace pb200
