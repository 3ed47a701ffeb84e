// Convolution as a TMA-fed tcgen05 implicit GEMM (sm_100a), BN folded, bias + residual + ReLU in
// the epilogue.  Replaces cuDNN fprop + BN + ReLU (+ add) of torchvision's ResNet-18
// (resnet.py:89-105, 268-276) that perseus/detector/models.py:34-40 runs.
//
// GEMM view: D[M = B*Ho*Wo][N = Cout] = A[M][K] * W[N][K]^T, K = (r, s, c).
//   * A is never materialised.  An M tile is 128 consecutive output pixels = a box of
//     (bh rows x Wo cols x imgs) of the NHWC input; for filter tap (r, s) the box is simply
//     shifted by (r - pad, s - pad), so ONE 4-D tiled TMA load per (tap, 64-channel block)
//     lands a 128 x 64 bf16 K-major SWIZZLE_128B operand tile in shared memory; out-of-image
//     taps are zero-filled by TMA's bounds check (the conv's zero padding).  Stride-2 layers use
//     four parity-split views of the input (even/odd rows x even/odd cols), one per tap parity.
//   * (the stem and layer1 do not use this kernel: stem_pool.cu / conv_p8.cu)
//   * warp roles (320 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane) + TMEM
//     owner, warps 2..9 = epilogue (TMEM -> registers -> bias/residual/ReLU -> bf16 NHWC).
//   * persistent CTAs (grid = #SMs), static tile round-robin, smem ring of kStages stages,
//     two TMEM accumulators so the epilogue of tile i overlaps the MMAs of tile i+1.
#include <stdlib.h>

#include "common.cuh"
#include "conv_umma.cuh"
#include "tma_host.h"
#include "umma_epilogue.cuh"

namespace pb200 {

constexpr int kUmmaThreads = 352;  // 2 TMA warps + MMA warp + 8 epilogue warps
constexpr int kTileM = 128;

// MODE: 0 = bf16 single pass; 1 = split bf16 ("bf16x3": [hi|lo] activations AND weights, 3 k-blocks per (tap, channel
// block)); 2 = "fp16x2": fp16 activations, weights = fp16 hi + fp16 lo -- every stage carries the hi AND the lo weight
// tile of its k-block and both are multiplied with the SAME activation tile (2 MMAs per K step, A bytes unchanged).
template <int BLOCK_N, int SW, int MODE = 0>
struct UmmaCfg {
  static constexpr int kABytes = kTileM * SW;
  static constexpr int kBBytes = BLOCK_N * SW * (MODE == 2 ? 2 : 1);
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (192 * 1024 / kStageBytes) > 8 ? 8 : (192 * 1024 / kStageBytes);
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024;  // + alignment slack
  static constexpr int kBlockK = SW / 2;                             // bf16 elements per k-block
  static constexpr int kKSteps = kBlockK / 16;                       // UMMA K = 16
  static constexpr int kTmemCols = 2 * BLOCK_N;                      // double-buffered accumulator
};

template <int BLOCK_N, int SW, int MODE>
__global__ void __launch_bounds__(kUmmaThreads, 1)
conv_umma_kernel(const __grid_constant__ UmmaConvMaps maps, const UmmaConvParams p) {
  using Cfg = UmmaCfg<BLOCK_N, SW, MODE>;
  constexpr bool SPLIT = MODE == 1;
  constexpr bool X2 = MODE == 2;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kStages];
  __shared__ __align__(8) uint64_t empty_bar[kStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ float s_bias[512];
  __shared__ float s_stat[2 * 512];  // training forward: per-CTA partial sum / sum of squares per output channel  // folded-BN shift of every output channel (Cout <= 512)

  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t tiles_addr = (raw_addr + 1023u) & ~1023u;
  uint8_t* tiles = smem_raw + (tiles_addr - raw_addr);
  uint8_t* smem_a = tiles;
  uint8_t* smem_b = tiles + kStages * Cfg::kABytes;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.a[0]);
    tma_prefetch_desc(&maps.b);
    if (p.ds_cblocks > 0) tma_prefetch_desc(&maps.ds);
#pragma unroll
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], p.cl);  // every CTA of the cluster releases a stage (its B slice is multicast)
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
  // PDL: everything above overlaps the previous kernel's tail; nothing below may run before it has completed (its
  // output is this kernel's input; in the training step the bias / packed weights are written by earlier kernels too)
  pdl_launch_dependents();
  pdl_wait();
  for (int i = threadIdx.x; i < p.Cout; i += kUmmaThreads) s_bias[i] = __ldg(p.bias + i);
  if (p.stats != nullptr)
    for (int i = threadIdx.x; i < 2 * p.Cout; i += kUmmaThreads) s_stat[i] = 0.f;
  tc_fence_before();
  __syncthreads();
  if (p.cl > 1) cluster_sync_all();  // peers' barriers are initialised before anyone multicasts into them
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  // cluster geometry: the CL CTAs of a cluster take CL consecutive M tiles of the same N tile and share its
  // weight tile: each CTA loads 1/CL of it and TMA-multicasts the slice into every CTA's stage buffer.
  const int cl = p.cl;
  const int crank = cl > 1 ? static_cast<int>(cluster_ctarank()) : 0;
  const int cluster_id = blockIdx.x / cl;
  const int num_clusters = gridDim.x / cl;
  const uint16_t cmask = static_cast<uint16_t>((1u << cl) - 1u);

  if (warp == 0 || warp == 10) {
    // two producer warps take alternate k-blocks: one warp's issue loop (barrier wait + expect_tx + 2 TMA issues)
    // costs about as many cycles as a 256-cycle k-block of MMAs (profiles/README.md r01)
    const int pi = warp == 0 ? 0 : 1;
    uint32_t g = 0;  // k-block counter of this CTA across all its tiles
    // ================================ TMA producer ================================
    // the whole warp runs the loop so every operand stays warp-uniform; one elected lane issues
    int stage = 0;
    uint32_t phase = 0;
    const int slice_rows = BLOCK_N / cl;
    for (int st = cluster_id; st < p.num_super; st += num_clusters) {
      const int m_super = st / p.num_n_tiles;
      const int n_tile = st - m_super * p.num_n_tiles;
      const int m_tile = m_super * cl + crank;
      int n0, ho0;
      if (p.tiles_per_img > 0) {
        n0 = m_tile / p.tiles_per_img;
        ho0 = (m_tile - n0 * p.tiles_per_img) * p.bh;
      } else {
        n0 = m_tile * p.imgs_per_tile;
        ho0 = 0;
      }
      // the issue loop is kept as light as possible: one warp must feed a k-block every 128-512 MMA cycles
      // K order: 64-channel block outermost, filter taps inside (the order csrc/conv_halo.cu walks its patches in)
      int kcol = 0;
      for (int cb = 0; cb < p.cblocks; ++cb) {
        for (int tap = 0; tap < p.ntaps; ++tap) {
          const int dw = p.tap_dw[tap];
          const CUtensorMap* amap = &maps.a[p.tap_mi[tap]];
          const int hrow = ho0 + p.tap_dh[tap];
          kcol = p.tap_k[tap] + cb * Cfg::kBlockK;
          // SPLIT ("bf16x3"): a = a_hi + a_lo (channels [0,Cin) and [Cin,2Cin) of the activation tensor),
          // w = w_hi + w_lo (columns [0,Ktot) and [Ktot,2Ktot) of the packed weights); three k-blocks per
          // (tap, channel block): a_hi*w_hi, a_lo*w_hi, a_hi*w_lo (lo*lo is below fp32 resolution)
#pragma unroll
          for (int ps = 0; ps < (SPLIT ? 3 : 1); ++ps) {
          if ((g++ & 1u) == static_cast<uint32_t>(pi)) {
          const int a_ch = cb * 64 + ((SPLIT && ps == 1) ? p.cin : 0);
          const int b_col = kcol + ((SPLIT && ps == 2) ? p.ktot : 0);
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
            tma_load_4d(smem_a + stage * Cfg::kABytes, amap, &full_bar[stage], a_ch, dw, hrow, n0);
#pragma unroll
            for (int hl = 0; hl < (X2 ? 2 : 1); ++hl) {  // fp16x2: the hi tile, then the lo tile (columns + ktot)
              uint8_t* bdst = smem_b + stage * Cfg::kBBytes + hl * BLOCK_N * SW;
              if (cl == 1 && p.b_tiled)  // KT layout (conv_umma.cuh): one contiguous run of 512-byte rows
                tma_load_2d(bdst, &maps.b, &full_bar[stage], 0,
                            (((hl * p.kb_total + (b_col >> 6)) * p.Cout) + n_tile * BLOCK_N) >> 2);
              else if (cl == 1)
                tma_load_2d(bdst, &maps.b, &full_bar[stage], b_col + hl * p.ktot, n_tile * BLOCK_N);
              else
                tma_load_2d_mcast(bdst + crank * slice_rows * SW, &maps.b, &full_bar[stage], b_col + hl * p.ktot,
                                  n_tile * BLOCK_N + crank * slice_rows, cmask);
            }
          }
          __syncwarp();
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1u;
          }
          }
        }
      }
      // fused 1x1/2 projection of the block input: extra k-blocks whose A tile is the (2 ho, 2 wo) pixel of ds_in
      kcol = p.ntaps * p.cin;  // the projection's columns follow the R*S*Cin columns of the conv
      for (int cb = 0; cb < p.ds_cblocks; ++cb, kcol += Cfg::kBlockK) {
        if ((g++ & 1u) == static_cast<uint32_t>(pi)) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          if (elect_one()) {
            mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
            tma_load_4d(smem_a + stage * Cfg::kABytes, &maps.ds, &full_bar[stage], cb * 64, 0, ho0, n0);
#pragma unroll
            for (int hl = 0; hl < (X2 ? 2 : 1); ++hl) {
              uint8_t* bdst = smem_b + stage * Cfg::kBBytes + hl * BLOCK_N * SW;
              if (cl == 1 && p.b_tiled)
                tma_load_2d(bdst, &maps.b, &full_bar[stage], 0,
                            (((hl * p.kb_total + (kcol >> 6)) * p.Cout) + n_tile * BLOCK_N) >> 2);
              else if (cl == 1)
                tma_load_2d(bdst, &maps.b, &full_bar[stage], kcol + hl * p.ktot, n_tile * BLOCK_N);
              else
                tma_load_2d_mcast(bdst + crank * slice_rows * SW, &maps.b, &full_bar[stage], kcol + hl * p.ktot,
                                  n_tile * BLOCK_N + crank * slice_rows, cmask);
            }
          }
          __syncwarp();
        }
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    constexpr uint32_t idesc = umma_idesc_16<X2>(kTileM, BLOCK_N);
    const uint32_t tm = warp_uniform(tmem_base);
    const uint32_t d_hi = desc_hi_swizzled(SW);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int st = cluster_id; st < p.num_super; st += num_clusters, ++it) {
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tm + acc * BLOCK_N;
      for (int kb = 0; kb < p.num_kblocks; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_lo = desc_lo(smem_u32(smem_a + stage * Cfg::kABytes), 16);
          const uint32_t b_lo = desc_lo(smem_u32(smem_b + stage * Cfg::kBBytes), 16);
#pragma unroll
          for (int k = 0; k < Cfg::kKSteps; ++k) {
            // advancing K by 16 elements = 32 bytes inside the swizzle span: +2 in the address field
            umma_bf16_split(d_tmem, a_lo + 2 * k, d_hi, b_lo + 2 * k, d_hi, idesc, (kb | k) != 0 ? 1u : 0u);
            if (X2)  // same activations x the lo weight tile (BLOCK_N rows further in the stage)
              umma_bf16_split(d_tmem, a_lo + 2 * k, d_hi, b_lo + 2 * k + (BLOCK_N * SW >> 4), d_hi, idesc, 1u);
          }
          // frees the smem stage (in every CTA of the cluster) once these MMAs have read it
          if (cl == 1) umma_commit(&empty_bar[stage]);
          else umma_commit_mcast(&empty_bar[stage], cmask);
          if (kb == p.num_kblocks - 1) umma_commit(&tmem_full_bar[acc]);
        }
        __syncwarp();
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else {
    // ================================ epilogue ====================================
    // 8 warps: warp -> (TMEM lane quarter q, column half).  Thread = one output pixel x BLOCK_N/2 channels.
    // The residual row of the tile (all 16-byte loads of this thread) is requested BEFORE the accumulator
    // wait so its L2/HBM latency hides behind the MMAs; the bias sits in shared memory.
    constexpr int kCols = BLOCK_N / 2;            // columns per epilogue thread
    constexpr int kChunks = kCols / 32;           // 32-column TMEM loads per thread
    constexpr int kVec = kCols / 8;               // 16-byte vectors of bf16 per thread
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const RowMap rmap = {p.up2, p.up_ph, p.up_pw, p.up_lh, p.up_lw};
    int it = 0;
    for (int st = cluster_id; st < p.num_super; st += num_clusters, ++it) {
      const int m_super = st / p.num_n_tiles;
      const int n_tile = st - m_super * p.num_n_tiles;
      const int m_tile = m_super * cl + crank;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      const long long gm = static_cast<long long>(m_tile) * kTileM + row;
      const bool valid = gm < p.M_total;
      const int col0 = n_tile * BLOCK_N + half * kCols;
      const bool has_res = p.residual != nullptr;
      if constexpr (SPLIT) {
        // rows of the split tensors are [hi(Cout) | lo(Cout)]; the residual is hi + lo; the result is stored as
        // hi = bf16(f), lo = bf16(f - hi)
        mbar_wait(&tmem_full_bar[acc], acc_phase);
        tc_fence_after();
        const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BLOCK_N + half * kCols;
        const long long rowoff = gm * (2LL * p.Cout) + col0;
        uint4* out_hi = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + rowoff);
        uint4* out_lo = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + rowoff + p.Cout);
        const uint4* res_hi = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.residual) + rowoff);
        const uint4* res_lo = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.residual) + rowoff + p.Cout);
        const float* bias = s_bias + col0;
#pragma unroll 1
        for (int chunk = 0; chunk < kChunks; ++chunk) {
          uint32_t v[32];
          tmem_ld_32x32(t_addr + chunk * 32, v);
          tmem_ld_wait_regs(v);
          if (chunk == kChunks - 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float f[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[8 * j + e]) + bias[chunk * 32 + 8 * j + e];
            if (has_res && valid) {
              const uint4 h4 = __ldg(res_hi + chunk * 4 + j), l4 = __ldg(res_lo + chunk * 4 + j);
              f[0] += bf16_lo(h4.x) + bf16_lo(l4.x); f[1] += bf16_hi(h4.x) + bf16_hi(l4.x);
              f[2] += bf16_lo(h4.y) + bf16_lo(l4.y); f[3] += bf16_hi(h4.y) + bf16_hi(l4.y);
              f[4] += bf16_lo(h4.z) + bf16_lo(l4.z); f[5] += bf16_hi(h4.z) + bf16_hi(l4.z);
              f[6] += bf16_lo(h4.w) + bf16_lo(l4.w); f[7] += bf16_hi(h4.w) + bf16_hi(l4.w);
            }
            if (p.relu) {
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
            }
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              hi[e] = pack_bf16x2(f[2 * e], f[2 * e + 1]);
              lo[e] = pack_bf16x2(f[2 * e] - bf16_lo(hi[e]), f[2 * e + 1] - bf16_hi(hi[e]));
            }
            if (valid) {
              out_hi[chunk * 4 + j] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
              out_lo[chunk * 4 + j] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
          }
        }
        continue;
      }
      uint4 rv[kVec];
      if constexpr (kCols % 64 == 0) {
        // line-granular path (umma_epilogue.cuh): residual and output move as complete 128-byte lines
        const long long gm_group = static_cast<long long>(m_tile) * kTileM + q * 32;
        if (has_res) {
          load_residual_lines<kCols>(static_cast<const __nv_bfloat16*>(p.residual), gm_group, p.M_total, p.Cout, col0, lane, rv, rmap);
        } else {
#pragma unroll
          for (int j = 0; j < kVec; ++j) rv[j] = make_uint4(0, 0, 0, 0);
        }
        mbar_wait(&tmem_full_bar[acc], acc_phase);
        tc_fence_after();
        const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BLOCK_N + half * kCols;
        epilogue_lines<kCols, X2>(t_addr, s_bias + col0, has_res, rv, static_cast<__nv_bfloat16*>(p.out), gm_group, p.M_total,
                              p.Cout, col0, p.relu, lane, [&]() {
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
                              }, p.stats != nullptr ? s_stat : nullptr, rmap);
        continue;
      }
      if (has_res && valid) {
        const uint4* res = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.residual) + rmap(gm) * p.Cout + col0);
#pragma unroll
        for (int j = 0; j < kVec; ++j) rv[j] = __ldg(res + j);
      } else {
#pragma unroll
        for (int j = 0; j < kVec; ++j) rv[j] = make_uint4(0, 0, 0, 0);
      }
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BLOCK_N + half * kCols;
      uint4* out_row = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + rmap(gm) * p.Cout + col0);
      const float* bias = s_bias + col0;
#pragma unroll
      for (int chunk = 0; chunk < kChunks; ++chunk) {
        uint4 oc[4];
        uint32_t v[32];
        tmem_ld_32x32(t_addr + chunk * 32, v);
        tmem_ld_wait_regs(v);
        if (chunk == kChunks - 1) {  // accumulator fully read: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[8 * j + e]) + bias[chunk * 32 + 8 * j + e];
          if (has_res) {
            const uint4 r4 = rv[chunk * 4 + j];
            f[0] += lo16<X2>(r4.x); f[1] += hi16<X2>(r4.x);
            f[2] += lo16<X2>(r4.y); f[3] += hi16<X2>(r4.y);
            f[4] += lo16<X2>(r4.z); f[5] += hi16<X2>(r4.z);
            f[6] += lo16<X2>(r4.w); f[7] += hi16<X2>(r4.w);
          }
          if (p.relu) {
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
          }
          const uint4 ov = make_uint4(pack16x2<X2>(f[0], f[1]), pack16x2<X2>(f[2], f[3]), pack16x2<X2>(f[4], f[5]),
                                      pack16x2<X2>(f[6], f[7]));
          if (valid) out_row[chunk * 4 + j] = ov;
          oc[j] = ov;
        }
        if (p.stats != nullptr) accumulate_stats32(oc, valid, s_stat, p.Cout, col0 + chunk * 32, lane);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (p.stats != nullptr)  // fused batch statistics: one fp64 atomic per channel and CTA
    for (int i = threadIdx.x; i < 2 * p.Cout; i += kUmmaThreads)
      if (s_stat[i] != 0.f) atomicAdd(&p.stats[i], static_cast<double>(s_stat[i]));
  if (cl > 1) cluster_sync_all();  // no CTA leaves while a peer may still multicast into it or signal its barriers
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
template <int BLOCK_N, int SW>
static int set_attr() {
  PB_CUDA(cudaFuncSetAttribute(conv_umma_kernel<BLOCK_N, SW, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               UmmaCfg<BLOCK_N, SW, 0>::kSmemBytes));
  PB_CUDA(cudaFuncSetAttribute(conv_umma_kernel<BLOCK_N, SW, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               UmmaCfg<BLOCK_N, SW, 1>::kSmemBytes));
  PB_CUDA(cudaFuncSetAttribute(conv_umma_kernel<BLOCK_N, SW, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               UmmaCfg<BLOCK_N, SW, 2>::kSmemBytes));
  return PB200_OK;
}

int umma_conv_init() {
  static bool done[kMaxDevices] = {};  // function attributes belong to the device's context
  const int dev = current_device();
  if (done[dev]) return PB200_OK;
  if (int rc = set_attr<64, 128>()) return rc;
  if (int rc = set_attr<128, 128>()) return rc;
  if (int rc = set_attr<256, 128>()) return rc;
  done[dev] = true;
  return PB200_OK;
}

static int encode_weight_map(CUtensorMap* m, const void* w, int Cout, int Ktot, int block_n, int sw) {
  const uint64_t dims[2] = {static_cast<uint64_t>(Ktot), static_cast<uint64_t>(Cout)};
  const uint64_t strides[1] = {static_cast<uint64_t>(Ktot) * 2};
  const uint32_t box[2] = {static_cast<uint32_t>(sw / 2), static_cast<uint32_t>(block_n)};
  return encode_tma_bf16(m, w, 2, dims, strides, box, sw);
}

// persistent grid = as many clusters as can be co-resident (clusters of 4 do not tile all 148 SMs)
template <int BLOCK_N>
static int query_clusters(int cl) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_sms() / cl * cl);
  cfg.blockDim = dim3(kUmmaThreads);
  cfg.dynamicSmemBytes = UmmaCfg<BLOCK_N, 128>::kSmemBytes;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cl;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, conv_umma_kernel<BLOCK_N, 128, 0>, &cfg) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  return n;
}
static int max_active_clusters(int block_n, int cl) {
  if (cl == 1) return num_sms();
  static int cache[kMaxDevices][3][5] = {};
  const int dev = current_device();
  const int bi = block_n == 64 ? 0 : (block_n == 128 ? 1 : 2);
  if (cache[dev][bi][cl] == 0)
    cache[dev][bi][cl] = block_n == 64 ? query_clusters<64>(cl) : (block_n == 128 ? query_clusters<128>(cl) : query_clusters<256>(cl));
  return cache[dev][bi][cl];
}

// Tile choice of a TMA-im2col conv: BLOCK_N in {64, 128, 256} x {single CTA, CTA pair}.  Cost model (cycles per 64-deep
// k-block and per CTA, measured with tools/mma_bench.cu and the r01/r02 per-op profiles): persistent CTAs take
// ceil(items / units) equal work items; a k-block costs max(MMA issue time, shared-memory traffic / bandwidth), where
//   MMA     = 4 K-steps x passes x max(48, N/2) cycles (an M128 x N x K16 MMA; passes = 2 in the fp16x2 mode),
//   traffic = what the MMAs READ from shared memory (4 x passes x (4 KB of A + 32 B per weight row held by this CTA))
//             + what TMA WRITES into it (16 KB of A + 128 B per weight row and pass),
// at 128 B/clk x 0.85.  A pair holds half of the weight rows per CTA, which is what makes the wide tiles pay: the r01
// choice (pair<128> for layer3, single<256> for layer4) sat at 452 / 903 cycles per k-block where pair<256> needs 602
// for twice / the same the work (measured 473 / 810).  cluster: 1 = single CTA, -2 = CTA pair.
void umma_choose_tile(long long m_tiles, int Cout, int passes, int* block_n, int* cluster) {
  double best_cost = -1.0;
  *block_n = 64;
  *cluster = 1;
  for (int bn = 64; bn <= 256 && bn <= Cout; bn *= 2) {
    if (Cout % bn) continue;
    for (int pair = 0; pair < 2; ++pair) {
      if (pair && m_tiles < 2) continue;
      const long long n_tiles = Cout / bn;
      const long long items = (pair ? (m_tiles + 1) / 2 : m_tiles) * n_tiles;
      const long long units = pair ? num_sms() / 2 : num_sms();
      const long long rounds = (items + units - 1) / units;
      const int brows = pair ? bn / 2 : bn;
      const double mma = 4.0 * passes * (bn / 2 < 48 ? 48 : bn / 2);
      const double smem = 4.0 * passes * (4096 + brows * 32) + 16384 + passes * brows * 128.0;
      const double t = mma > smem / (128.0 * 0.85) ? mma : smem / (128.0 * 0.85);
      const double cost = static_cast<double>(rounds) * t;
      if (best_cost < 0 || cost < best_cost * 0.999) {
        best_cost = cost;
        *block_n = bn;
        *cluster = pair ? -2 : 1;
      }
    }
  }
}

int build_umma_conv(UmmaConvLaunch* L, const void* in, int B, int Hin, int Win, int Cin,
                    const void* wpacked, const float* bias, const void* residual, void* out, int Cout,
                    int R, int stride, int pad, int relu, int block_n, int cluster, int split, const void* ds_in,
                    int ds_Hin, int ds_Win, int ds_Cin, double* stats, int b_tiled) {
  PB_CHECK(stride == 1 || stride == 2, "umma conv: stride %d unsupported", stride);
  PB_CHECK(Cin % 64 == 0, "umma conv: Cin %d not a multiple of 64", Cin);
  PB_CHECK(block_n == 64 || block_n == 128 || block_n == 256, "umma conv: block_n %d", block_n);
  PB_CHECK(Cout % block_n == 0, "umma conv: Cout %d %% block_n %d", Cout, block_n);
  PB_CHECK(Cout <= 512, "umma conv: Cout %d > 512 (bias staging)", Cout);
  const int Ho = (Hin + 2 * pad - R) / stride + 1;
  const int Wo = (Win + 2 * pad - R) / stride + 1;
  PB_CHECK(stride == 1 || (Hin % 2 == 0 && Win % 2 == 0), "umma conv: stride 2 needs even H, W");
  PB_CHECK(Wo <= kTileM && kTileM % Wo == 0, "umma conv: Wo=%d must divide 128", Wo);
  const int P = Ho * Wo;
  int bh, imgs, tiles_per_img;
  if (P >= kTileM) {
    PB_CHECK(P % kTileM == 0, "umma conv: Ho*Wo=%d must be a multiple of 128", P);
    bh = kTileM / Wo;
    imgs = 1;
    tiles_per_img = P / kTileM;
  } else {
    PB_CHECK(kTileM % P == 0, "umma conv: Ho*Wo=%d must divide 128", P);
    bh = Ho;
    imgs = kTileM / P;
    tiles_per_img = 0;
  }
  const long long M = static_cast<long long>(B) * P;
  memset(L, 0, sizeof(*L));
  UmmaConvParams& p = L->p;
  p.M_total = static_cast<int>(M);
  p.Cout = Cout;
  p.stem = 0;
  p.Ho = Ho;
  p.Wo = Wo;
  p.tiles_per_img = tiles_per_img;
  p.imgs_per_tile = imgs;
  p.bh = bh;
  p.stride = stride;
  p.pad = pad;
  p.S = R;
  p.cblocks = Cin / 64;
  p.num_kblocks = R * R * p.cblocks * (split == 1 ? 3 : 1);
  p.ntaps = R * R;
  for (int t = 0; t < R * R; ++t) {
    const int r = t / R, sx = t - r * R;
    int dh = r - pad, dw = sx - pad, mi = 0;
    if (stride == 2) {
      const int hp = dh & 1, wp = dw & 1;
      dh = (dh - hp) >> 1;
      dw = (dw - wp) >> 1;
      mi = hp * 2 + wp;
    }
    p.tap_dh[t] = dh;
    p.tap_dw[t] = dw;
    p.tap_mi[t] = mi;
    p.tap_k[t] = t * Cin;
  }
  p.up2 = p.up_ph = p.up_pw = p.up_lh = p.up_lw = 0;
  p.ds_cblocks = 0;
  PB_CHECK(stats == nullptr || !split, "umma conv: fused statistics are not available in the split / fp16x2 modes");
  if (ds_in != nullptr) {
    PB_CHECK(split != 1 && ds_Cin % 64 == 0 && ds_Hin == 2 * Ho && ds_Win == 2 * Wo,
             "umma conv: fused projection needs a bf16 [B][2Ho][2Wo][64k] input (got %dx%dx%d)", ds_Hin, ds_Win, ds_Cin);
    p.ds_cblocks = ds_Cin / 64;
    p.num_kblocks += p.ds_cblocks;
  }
  p.cin = Cin;
  // column offset of the lo halves in the packed weight rows [hi K | lo K] (K includes a fused projection's columns)
  p.ktot = R * R * Cin + ((split == 2 && ds_in != nullptr) ? ds_Cin : 0);
  L->split = split;
  if (split == 1) cluster = 1;  // the split-bf16 mode runs on the single-CTA kernel
  p.num_n_tiles = Cout / block_n;
  const int m_tiles = static_cast<int>((M + kTileM - 1) / kTileM);
  p.num_tiles = m_tiles * p.num_n_tiles;
  p.relu = relu;
  p.bias = bias;
  p.residual = residual;
  p.out = out;
  p.stats = stats;
  L->block_n = block_n;
  L->sw = 128;
  // cluster size: share the weight tile among up to 4 CTAs when there are enough M tiles to keep every SM busy
  int cl = cluster;
  if (cl == 0) {
    const char* force = getenv("PB200_FORCE_CLUSTER");  // tests: exercise the cluster paths on small shapes
    if (force != nullptr && force[1] == 0) {
      if (force[0] == '1' || force[0] == '2' || force[0] == '4') cl = force[0] - '0';
      if (force[0] == 'p') cl = -2;
    }
  }
  // default: CTA pairs (halves the weight bytes landing in each SM) whenever there is a whole wave of pair tiles;
  // plain multicast clusters measured no gain (profiles/README.md) and stay opt-in
  // (fp16x2: a pair halves the doubled weight bytes per SM, so it pays from one full wave on)
  if (cl == 0)
    cl = (static_cast<long long>(m_tiles) * p.num_n_tiles >= (split == 2 ? 1LL : 2LL) * num_sms()) ? -2 : 1;
  PB_CHECK(cl == 1 || cl == 2 || cl == 4 || cl == -2, "umma conv: cluster %d", cl);
  L->pair = cl == -2;
  if (L->pair) cl = 2;
  p.cl = cl;
  p.num_super = (m_tiles + cl - 1) / cl * p.num_n_tiles;
  int max_clusters;
  L->resb = 0;
  p.a_stages = 0;
  // OFF by default: measured on layer2 (r02, profiles/README.md) it is 5 % SLOWER than streaming the weight tile with the
  // activations (0.274 vs 0.261 ms for the four layer2 convs) -- the layer is not bound by the bytes of its weight tile,
  // and the resident image leaves 4 activation stages instead of 8.  PB200_RESB=1 enables it for A/B runs.
  if (L->pair && split == 0 && p.num_n_tiles == 1 && stats == nullptr && getenv("PB200_RESB")) {
    // resident weights (conv_umma2.cu RESB): this CTA's half of all k-blocks + at least 3 activation stages must fit
    const int resb_bytes = p.num_kblocks * (block_n / 2) * 128;
    const int budget = 216 * 1024;  // + 1 KB alignment slack + ~7.4 KB static shared memory < 227 KB
    int stages = (budget - resb_bytes) / (128 * 128);
    if (stages > 8) stages = 8;
    if (stages >= 3) {
      L->resb = 1;
      p.a_stages = stages;
      L->smem = stages * 128 * 128 + resb_bytes + 1024;
    }
  }
  if (L->pair) {
    if (int rc = umma_pair_init()) return rc;
    max_clusters = umma_pair_max_pairs(block_n, split == 2);
  } else {
    max_clusters = max_active_clusters(block_n, cl);
  }
  PB_CHECK(max_clusters > 0, "umma conv: no resident cluster of %d CTAs", cl);
  L->grid = (p.num_super < max_clusters ? p.num_super : max_clusters) * cl;
  L->flops = 2.0 * static_cast<double>(M) * Cout * (R * R * Cin + (ds_in != nullptr ? ds_Cin : 0));

  const uint32_t box[4] = {64u, static_cast<uint32_t>(Wo), static_cast<uint32_t>(bh),
                           static_cast<uint32_t>(imgs)};
  const uint64_t e = 2;  // bytes per element
  const uint64_t Cs = static_cast<uint64_t>(Cin) * (split == 1 ? 2 : 1);  // stored channels per pixel: [hi | lo] when split
  if (stride == 1) {
    const uint64_t dims[4] = {Cs, (uint64_t)Win, (uint64_t)Hin, (uint64_t)B};
    const uint64_t str[3] = {Cs * e, (uint64_t)Win * Cs * e, (uint64_t)Hin * Win * Cs * e};
    if (int rc = encode_tma_bf16(&L->maps.a[0], in, 4, dims, str, box, 128)) return rc;
    for (int i = 1; i < 4; ++i) L->maps.a[i] = L->maps.a[0];
  } else {
    for (int hp = 0; hp < 2; ++hp)
      for (int wp = 0; wp < 2; ++wp) {
        const uint8_t* base = static_cast<const uint8_t*>(in) + (static_cast<uint64_t>(hp) * Win + wp) * Cs * e;
        const uint64_t dims[4] = {Cs, (uint64_t)Win / 2, (uint64_t)Hin / 2, (uint64_t)B};
        const uint64_t str[3] = {2 * Cs * e, 2 * (uint64_t)Win * Cs * e, (uint64_t)Hin * Win * Cs * e};
        if (int rc = encode_tma_bf16(&L->maps.a[hp * 2 + wp], base, 4, dims, str, box, 128)) return rc;
      }
  }
  L->maps.ds = L->maps.a[0];
  if (ds_in != nullptr) {  // parity (0, 0) of the block input: pixel (2 ho, 2 wo), same box as the conv's own tiles
    const uint64_t dC = static_cast<uint64_t>(ds_Cin);
    const uint64_t dims[4] = {dC, (uint64_t)ds_Win / 2, (uint64_t)ds_Hin / 2, (uint64_t)B};
    const uint64_t str[3] = {2 * dC * e, 2 * (uint64_t)ds_Win * dC * e, (uint64_t)ds_Hin * ds_Win * dC * e};
    if (int rc = encode_tma_bf16(&L->maps.ds, ds_in, 4, dims, str, box, 128)) return rc;
  }
  const int Krow = R * R * Cin + (ds_in != nullptr ? ds_Cin : 0);
  p.b_tiled = b_tiled ? 1 : 0;
  p.reverse = 0;
  p.kb_total = Krow / 64;
  if (b_tiled) {
    PB_CHECK(split != 1 && (L->pair || p.cl == 1), "umma conv: KT-layout weights need the single-CTA or the pair kernel");
    // the packed buffer as 512-byte rows; a CTA's tile of one k-block = (rows held by the CTA) / 4 consecutive rows
    const int rows_cta = block_n / p.cl;
    const uint64_t total = static_cast<uint64_t>(split == 2 ? 2 : 1) * p.kb_total * Cout * 128;
    const uint64_t dims[2] = {256, total / 512};
    const uint64_t strides[1] = {512};
    const uint32_t box[2] = {256u, static_cast<uint32_t>(rows_cta / 4)};
    return encode_tma_bf16(&L->maps.b, wpacked, 2, dims, strides, box, 0);
  }
  return encode_weight_map(&L->maps.b, wpacked, Cout, Krow * (split ? 2 : 1), block_n / p.cl, 128);
}

int build_umma_dgrad_s2(UmmaConvLaunch* L, const void* dz, int B, int Ho, int Wo, int Cout_f, const void* w_dgrad,
                        const float* zero_bias, const void* residual_or_null, void* dx, int Cin_f, int R, int ph, int pw,
                        int block_n) {
  PB_CHECK(R == 3 || R == 1, "dgrad_s2: R = %d (3x3 pad 1 or 1x1 pad 0)", R);
  PB_CHECK(ph >= 0 && ph < 2 && pw >= 0 && pw < 2, "dgrad_s2: parity (%d, %d)", ph, pw);
  PB_CHECK((Ho & (Ho - 1)) == 0 && (Wo & (Wo - 1)) == 0, "dgrad_s2: Ho x Wo = %d x %d must be powers of two", Ho, Wo);
  // the tile grid is the LOW-resolution dz grid; A boxes come from dz at stride 1 (zero fill outside = the taps that
  // fall off the bottom / right edge), B from the full data-gradient operand at the selected taps' k columns
  if (int rc = build_umma_conv(L, dz, B, Ho, Wo, Cout_f, w_dgrad, zero_bias, residual_or_null, dx, Cin_f, R, 1, R / 2, 0,
                               block_n))
    return rc;
  UmmaConvParams& p = L->p;
  p.reverse = 0;
  L->resb = 0;  // the tap table below re-orders the k columns: the resident-weights walk (columns kb * 64) does not apply
  // forward: z[i] reads x[2i + r - pad].  For x row u = 2a + ph the forward taps are r = ph + pad (mod 2):
  // R = 3: ph = 0 -> r = 1 (i = a); ph = 1 -> r = 0 (i = a + 1), r = 2 (i = a).  R = 1: ph = 0 -> r = 0 (i = a); ph = 1 -> none
  int rr[2], ri[2], nr = 0, cc[2], ci[2], nc = 0;
  if (R == 3) {
    if (ph == 0) { rr[nr] = 1; ri[nr++] = 0; } else { rr[nr] = 0; ri[nr++] = 1; rr[nr] = 2; ri[nr++] = 0; }
    if (pw == 0) { cc[nc] = 1; ci[nc++] = 0; } else { cc[nc] = 0; ci[nc++] = 1; cc[nc] = 2; ci[nc++] = 0; }
  } else {
    if (ph == 0) { rr[nr] = 0; ri[nr++] = 0; }
    if (pw == 0) { cc[nc] = 0; ci[nc++] = 0; }
  }
  p.ntaps = 0;
  for (int a = 0; a < nr; ++a)
    for (int b = 0; b < nc; ++b) {
      const int t = p.ntaps++;
      p.tap_dh[t] = ri[a];
      p.tap_dw[t] = ci[b];
      p.tap_mi[t] = 0;
      p.tap_k[t] = ((R - 1 - rr[a]) * R + (R - 1 - cc[b])) * Cout_f;  // pack mode 1 stores w[r][s] at tap (R-1-r, R-1-s)
    }
  p.num_kblocks = p.ntaps * p.cblocks;
  if (p.ntaps == 0) p.num_tiles = 0;  // nothing to launch: this parity receives no gradient through the conv
  p.up2 = 1;
  p.up_ph = ph;
  p.up_pw = pw;
  p.up_lh = 0;
  while ((1 << p.up_lh) < Ho) ++p.up_lh;
  p.up_lw = 0;
  while ((1 << p.up_lw) < Wo) ++p.up_lw;
  L->flops = 2.0 * static_cast<double>(p.M_total) * Cin_f * p.ntaps * Cout_f;
  return PB200_OK;
}

template <int BLOCK_N, int MODE>
static int launch_one(const UmmaConvLaunch& L, cudaStream_t st) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(L.grid);
  cfg.blockDim = dim3(kUmmaThreads);
  cfg.dynamicSmemBytes = UmmaCfg<BLOCK_N, 128, MODE>::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = L.p.cl;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1 + pdl_launch_attrs(attr + 1);
  PB_CUDA(cudaLaunchKernelEx(&cfg, conv_umma_kernel<BLOCK_N, 128, MODE>, L.maps, L.p));
  return PB200_OK;
}

int launch_umma_conv(const UmmaConvLaunch& L, cudaStream_t st) {
  if (L.p.num_tiles == 0) return PB200_OK;
  if (L.pair) return launch_umma_pair(L, st);
  if (L.split == 1) {
    if (L.block_n == 64) return launch_one<64, 1>(L, st);
    if (L.block_n == 128) return launch_one<128, 1>(L, st);
    return launch_one<256, 1>(L, st);
  }
  if (L.split == 2) {
    if (L.block_n == 64) return launch_one<64, 2>(L, st);
    if (L.block_n == 128) return launch_one<128, 2>(L, st);
    return launch_one<256, 2>(L, st);
  }
  if (L.block_n == 64) return launch_one<64, 0>(L, st);
  if (L.block_n == 128) return launch_one<128, 0>(L, st);
  return launch_one<256, 0>(L, st);
}

}  // namespace pb200
