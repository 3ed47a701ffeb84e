// 3x3 stride-1 convolution 64 -> 64 channels (+ folded BN, residual add, ReLU) on the P8 layout:
// layer1 of the detector's ResNet-18 (torchvision BasicBlock.forward, resnet.py:89-105, reached from
// perseus/detector/models.py:34-40).  tcgen05 implicit GEMM in which NO operand is ever re-read:
//
//   * P8 = [n][C/8][H+2][W+2][8] bf16 with a zero border (layout.h).  Over the flattened padded plane
//     a 3x3 conv is out[q] = sum_{r,s} in[q + r*(W+2) + s] . w[r][s], so one shared-memory patch of
//     128 + 2(W+2) + 2 positions per 8-channel chunk serves all nine taps of a 128-position M tile: each tap
//     is the same patch read through a no-swizzle K-major descriptor (row pitch 16 B, LBO = chunk
//     pitch) whose start address is shifted by (r*(W+2) + s) * 16 bytes.  The two positions per row that
//     fall on the border produce garbage accumulators and are stored as the zero border itself.
//   * the whole folded weight tensor (64 x 576 bf16 = 72 KB) stays resident in shared memory.
//   * L2 -> SM traffic per tile is 33.5 KB (one patch) instead of 9 x (16 KB + 8 KB) for a TMA-im2col
//     implicit GEMM, which at N = 64 is L2-bandwidth-bound (profiles/README.md, r01 v0).
//   * warp 0: producer (8 cp.async.bulk per patch, no tensor map); warp 1: 36 x tcgen05.mma
//     (M128 N64 K16) per tile into one of two TMEM accumulators; warps 2..5: epilogue
//     (TMEM -> +bias (+residual) -> ReLU -> bf16 -> 16-byte stores, lanes = consecutive positions so
//     every store instruction of a warp is one contiguous 512-byte run in P8).
//   * PAIR (default): two CTAs of a cluster work on tiles 2j and 2j+1 with ONE tcgen05.mma.cta_group::2 per K step
//     (M = 256: 128 patch rows from each CTA; each CTA holds HALF of the weight rows: output channels 32c..32c+31 in
//     bf16, the W_hi or the W_lo image in fp16x2).  The N = 64 MMA is bound by its operand fetch (6 KB per K step from one
//     SM's shared memory for 32 cycles of math); in a pair every SM fetches 4 KB of patch + half of the weights.
#include <stdlib.h>

#include "common.cuh"
#include "layout.h"

namespace pb200 {

namespace {

constexpr int p8_threads(int epi_warps) { return 64 + 32 * epi_warps; }
constexpr int kChunks = 8;                       // Cin / 8
constexpr int kWBytes = 64 * 576 * 2;            // 73728
constexpr int kWSbo = 72 * 128;                  // 9216: bytes per 8-output-channel group

struct ConvP8Params {
  const uint8_t* in;        // P8 [B][8][H+2][W+2][8]
  const uint8_t* wimg;      // 73728-byte shared-memory image
  const float* bias;        // [64]
  const uint8_t* residual;  // P8 or null
  uint8_t* out;             // P8, or NHWC [B][H][W][64] when out_nhwc
  int H, W, Wp;
  int tiles_per_img, num_tiles;
  int npos;                 // patch positions per chunk = 128 + 2*Wp + 2
  int patch_pitch;          // bytes between chunk planes of a patch in shared memory (multiple of 128)
  int relu, out_nhwc;
  int reverse;              // 1: the tile list is walked from the LAST tile to the first (see launch_conv_p8)
  long long img_bytes;      // bytes per P8 image
  long long plane_bytes;    // bytes per P8 chunk plane
};

// kEpiWarps = 4 (thread = position x 64 channels) or 8 (x 32 channels).  With 4 the epilogue has ONE warp per scheduler
// and its ~350 dependent instructions per tile take longer than the tile's 36 MMAs (1728 cycles): latency-bound.
//
// X2 = the "fp16x2" mode: fp16 activations, weights = fp16 hi + fp16 lo.  The lo image sits right behind the hi image
// in shared memory, so ONE N = 128 descriptor (8-channel groups kWSbo apart) reads [W_hi ; W_lo]: accumulator columns
// [0, 64) = A.W_hi and [64, 128) = A.W_lo from a single read of the A rows -- 64 cycles per K step instead of the
// 2 x 48 of two N = 64 MMAs (the N = 64 MMA is bound by its operand fetch, tools/mma_bench.cu).  The epilogue adds the
// two column groups.  144 KB of resident weights leave room for two patch stages.
template <int kEpiWarps, bool X2, bool PAIR = false>
__device__ __forceinline__ void conv_p8_body(const ConvP8Params& p) {
  // X2: the 144 KB of weights leave 67 KB for patches = two whole-tile stages, i.e. ONE tile of prefetch (1.2 us of
  // MMAs) against ~2 us of patch latency.  The ring is therefore cut into HALF-tile stages (4 of the 8 channel chunks
  // each, 4 stages): the producer runs three half-stages ahead of the MMAs.
  // PAIR: each CTA keeps half of the weight rows (36 / 72 KB): room for 5 / 4 WHOLE-tile stages.
  constexpr int kSplit = (X2 && !PAIR) ? 2 : 1;    // stages per tile
  constexpr int kStageChunks = kChunks / kSplit;   // channel chunks per stage
  constexpr int kStages = (PAIR && !X2) ? 5 : 4;
  constexpr int kWAll = (X2 ? 2 * kWBytes : kWBytes) / (PAIR ? 2 : 1);  // resident weight bytes of THIS CTA
  constexpr int kAccCols = X2 ? 128 : 64;
  constexpr int kNch = 8 / (kEpiWarps / 4);  // 8-channel chunks (P8 planes) per epilogue thread
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kStages];
  __shared__ __align__(8) uint64_t empty_bar[kStages];
  __shared__ __align__(8) uint64_t w_bar;
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ float s_bias[64];

  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 127u) & ~127u) - raw_addr);
  uint8_t* s_w = smem;
  uint8_t* s_patch = smem + kWAll;
  const int stage_bytes = kStageChunks * p.patch_pitch;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int crank = PAIR ? static_cast<int>(cluster_ctarank()) : 0;
  const bool leader = crank == 0;
  // PAIR: blockIdx.x = 2 * pair + crank, so "tile = blockIdx.x + i * gridDim.x" hands the pair tiles 2j and 2j+1; both
  // CTAs run as long as the LEADER's tile exists (the peer of an odd last tile re-reads it and stores nothing)
  auto alive = [&](int tile) { return tile - crank < p.num_tiles; };

  if (threadIdx.x == 0) {
    // PAIR: the leader's full / weight barriers also take one arrival relayed from the peer (its patch / weights landed);
    // the leader's accumulator-free barriers take both CTAs' epilogue warps
#pragma unroll
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], (PAIR && leader) ? 2 : 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(&w_bar, (PAIR && leader) ? 2 : 1);
    mbar_init(&tmem_full_bar[0], 1);
    mbar_init(&tmem_full_bar[1], 1);
    mbar_init(&tmem_empty_bar[0], PAIR ? 2 * kEpiWarps : kEpiWarps);
    mbar_init(&tmem_empty_bar[1], PAIR ? 2 * kEpiWarps : kEpiWarps);
    fence_barrier_init();
  }
  if (warp == 1) {
    if constexpr (PAIR) {
      tmem_alloc2(&tmem_base_slot, 2 * kAccCols);
      tmem_relinquish2();
    } else {
      tmem_alloc(&tmem_base_slot, 2 * kAccCols);
      tmem_relinquish();
    }
  }
  if (threadIdx.x >= 64 && threadIdx.x < 128) s_bias[threadIdx.x - 64] = __ldg(p.bias + threadIdx.x - 64);
  tc_fence_before();
  __syncthreads();
  if constexpr (PAIR) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  pdl_launch_dependents();

  if (warp == 0) {
    // ================================ producer ====================================
    // the whole warp runs the loop (operands stay warp-uniform); one elected lane issues the copies
    if (elect_one()) {
      mbar_expect_tx(&w_bar, kWAll);
      // plan constant: fetched while the previous kernel drains (PDL).  PAIR: this CTA's half of the rows
      bulk_load_1d(s_w, p.wimg + crank * kWAll, kWAll, &w_bar);
    }
    pdl_wait();  // the input patches are the previous kernel's output
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t chunk_bytes = static_cast<uint32_t>(p.npos) * 16u;
    for (int tile = blockIdx.x; alive(tile); tile += gridDim.x) {
      int tl = tile < p.num_tiles ? tile : p.num_tiles - 1;
      if (p.reverse) tl = p.num_tiles - 1 - tl;
      const int n = tl / p.tiles_per_img;
      const int q0 = (tl - n * p.tiles_per_img) * 128;
      const uint8_t* src = p.in + n * p.img_bytes + static_cast<long long>(q0) * 16;
#pragma unroll
      for (int h = 0; h < kSplit; ++h) {
        mbar_wait(&empty_bar[stage], phase ^ 1u);
        if (elect_one()) {
          mbar_expect_tx(&full_bar[stage], chunk_bytes * kStageChunks);
          uint8_t* dst = s_patch + stage * stage_bytes;
#pragma unroll
          for (int j = 0; j < kStageChunks; ++j)
            bulk_load_1d(dst + j * p.patch_pitch, src + (h * kStageChunks + j) * p.plane_bytes, chunk_bytes, &full_bar[stage]);
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
    if (PAIR && !leader) {
      // the peer relays "my weights / my patch of this stage have landed" to the leader's barriers: cp.async.bulk can
      // only credit a barrier of the CTA it writes to, and the leader's MMAs read both CTAs' shared memory
      const uint32_t lw = mapa_u32(&w_bar, 0);
      const uint32_t lfull0 = mapa_u32(&full_bar[0], 0);
      mbar_wait(&w_bar, 0);
      if (lane == 0) mbar_arrive_cluster(lw);
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; alive(tile); tile += gridDim.x) {
        mbar_wait(&full_bar[stage], phase);
        if (lane == 0) mbar_arrive_cluster(lfull0 + stage * 8);
        __syncwarp();
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1u;
        }
      }
    } else {
      constexpr uint32_t idesc = umma_idesc_16<X2>(PAIR ? 256 : 128, kAccCols);
      const uint32_t tm = warp_uniform(tmem_base);
      const uint32_t a_hi = desc_hi_interleaved(128);
      const uint32_t b_hi = desc_hi_interleaved(kWSbo);
      const uint32_t b_lo0 = desc_lo(smem_u32(s_w), 128);
      const uint32_t pitch16 = static_cast<uint32_t>(p.patch_pitch) >> 4;
      mbar_wait(&w_bar, 0);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        mbar_wait(&tmem_empty_bar[acc], ((it >> 1) & 1u) ^ 1u);
#pragma unroll
        for (int h = 0; h < kSplit; ++h) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t d_tmem = tm + acc * kAccCols;
            const uint32_t a_lo0 = desc_lo(smem_u32(s_patch) + stage * stage_bytes, p.patch_pitch);
#pragma unroll
            for (int t = 0; t < 9; ++t) {
              const int r = t / 3, s = t - 3 * r;
              const uint32_t a_tap = a_lo0 + static_cast<uint32_t>(r * p.Wp + s);  // (r*Wp + s) * 16 B >> 4
#pragma unroll
              for (int kk = 0; kk < kStageChunks / 2; ++kk) {  // K = 16 = two channel chunks of this stage
                const uint32_t b_lo = b_lo0 + (t * 8 + h * kStageChunks + 2 * kk) * 8;
                const uint32_t accum = (h | t | kk) != 0 ? 1u : 0u;
                if constexpr (PAIR) umma2_bf16_split(d_tmem, a_tap + 2 * kk * pitch16, a_hi, b_lo, b_hi, idesc, accum);
                else umma_bf16_split(d_tmem, a_tap + 2 * kk * pitch16, a_hi, b_lo, b_hi, idesc, accum);
              }
            }
            if constexpr (PAIR) {
              umma2_commit_mcast(&empty_bar[stage], 3);  // both CTAs may refill this stage
              if (h == kSplit - 1) umma2_commit_mcast(&tmem_full_bar[acc], 3);
            } else {
              umma_commit(&empty_bar[stage]);
              if (h == kSplit - 1) umma_commit(&tmem_full_bar[acc]);
            }
          }
          __syncwarp();
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else {
    // ================================ epilogue ====================================
    pdl_wait();  // residual reads / output writes must follow the previous kernel (the output buffer may be its input)
    const int q = warp & 3;
    const int j0 = ((warp - 2) >> 2) * kNch;  // first chunk plane of this thread
    const int m = q * 32 + lane;
    const long long plane16 = p.plane_bytes / 16;
    int it = 0;
    const bool has_res = p.residual != nullptr;
    // residual tiles of THIS thread's row: loaded TWO tiles ahead into two register sets (one tile of lead, ~1.3 us,
    // still left the epilogue waiting on HBM for the residual at the top of every tile)
    uint4 rv0[kNch], rv1[kNch];
    auto load_residual = [&](int tile, uint4 (&rv)[kNch]) {
      bool ok = false;
      if (has_res && tile < p.num_tiles) {  // (a peer tile past the end loads nothing)
        const int tl = p.reverse ? p.num_tiles - 1 - tile : tile;
        const int n = tl / p.tiles_per_img;
        const int q0 = (tl - n * p.tiles_per_img) * 128 + m;
        const int h = q0 / p.Wp;
        if (q0 < p.H * p.Wp && q0 - h * p.Wp < p.W) {
          const uint4* res = reinterpret_cast<const uint4*>(p.residual + n * p.img_bytes) + q0 + p.Wp + 1;
#pragma unroll
          for (int j = 0; j < kNch; ++j) rv[j] = __ldg(res + (j0 + j) * plane16);
          ok = true;
        }
      }
      if (!ok) {
#pragma unroll
        for (int j = 0; j < kNch; ++j) rv[j] = make_uint4(0, 0, 0, 0);
      }
    };
    load_residual(blockIdx.x, rv0);
    load_residual(blockIdx.x + gridDim.x, rv1);
    float bias[8 * kNch];
#pragma unroll
    for (int e = 0; e < 8 * kNch; ++e) bias[e] = s_bias[8 * j0 + e];
    const uint32_t lempty0 = PAIR ? mapa_u32(&tmem_empty_bar[0], 0) : 0u;
    auto process = [&](int tile, uint4 (&rv)[kNch]) {
      const bool live = tile < p.num_tiles;  // false only for the peer of an odd last tile
      const int tl = (p.reverse && live) ? p.num_tiles - 1 - tile : tile;
      const int n = tl / p.tiles_per_img;
      const int ti = tl - n * p.tiles_per_img;
      const int q0 = ti * 128 + m;
      const int h = q0 / p.Wp;
      const int w = q0 - h * p.Wp;
      const bool in_plane = live && q0 < p.H * p.Wp;
      const bool valid = in_plane && w < p.W;
      const int a = it & 1;
      const long long opos = static_cast<long long>(q0) + p.Wp + 1;  // padded position of this output
      mbar_wait(&tmem_full_bar[a], (it >> 1) & 1u);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + a * kAccCols + 8 * j0;
      uint32_t x[kNch / 4][32];
#pragma unroll
      for (int c = 0; c < kNch / 4; ++c) tmem_ld_32x32(t_addr + 32 * c, x[c]);
#pragma unroll
      for (int c = 0; c < kNch / 4; ++c) tmem_ld_wait_regs(x[c]);
      if constexpr (X2) {  // + the A.W_lo partial sums (accumulator columns 64 further)
        uint32_t xl[kNch / 4][32];
#pragma unroll
        for (int c = 0; c < kNch / 4; ++c) tmem_ld_32x32(t_addr + 64 + 32 * c, xl[c]);
#pragma unroll
        for (int c = 0; c < kNch / 4; ++c) tmem_ld_wait_regs(xl[c]);
#pragma unroll
        for (int c = 0; c < kNch / 4; ++c)
#pragma unroll
          for (int e = 0; e < 32; ++e) x[c][e] = __float_as_uint(__uint_as_float(x[c][e]) + __uint_as_float(xl[c][e]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(&tmem_empty_bar[a]);
        else mbar_arrive_cluster(lempty0 + a * 8);
      }

      uint4 o[kNch];
#pragma unroll
      for (int j = 0; j < kNch; ++j) {
        const uint32_t* xj = &x[j / 4][8 * (j % 4)];
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(xj[e]) + bias[8 * j + e];
        if (has_res) {
          f[0] += lo16<X2>(rv[j].x); f[1] += hi16<X2>(rv[j].x);
          f[2] += lo16<X2>(rv[j].y); f[3] += hi16<X2>(rv[j].y);
          f[4] += lo16<X2>(rv[j].z); f[5] += hi16<X2>(rv[j].z);
          f[6] += lo16<X2>(rv[j].w); f[7] += hi16<X2>(rv[j].w);
        }
        if (p.relu) {
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
        }
        o[j] = valid ? make_uint4(pack16x2<X2>(f[0], f[1]), pack16x2<X2>(f[2], f[3]), pack16x2<X2>(f[4], f[5]),
                                  pack16x2<X2>(f[6], f[7]))
                     : make_uint4(0, 0, 0, 0);
      }
      load_residual(tile + 2 * gridDim.x, rv);  // rv is dead from here on: prefetch the residual of the tile after next
      if (p.out_nhwc) {
        // NHWC pixel = 8 chunks = one 128-byte line.  A lane holds kNch chunks of ITS position; storing them directly
        // makes every store instruction touch 32 different lines with 16 bytes each.  A kNch x kNch transpose inside
        // each group of kNch lanes (butterfly stages of 16-byte shuffles) turns lane l into the owner of chunk
        // j0 + l % kNch of the group's positions, so one instruction writes complete 16 * kNch-byte runs.
        const int ln = lane & (kNch - 1);
#pragma unroll
        for (int sft = kNch / 2; sft >= 1; sft >>= 1) {
          const bool up = (ln & sft) != 0;
#pragma unroll
          for (int j = 0; j < kNch; ++j) {
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
        // pixel index of this lane's own position (-1: not an output); position g*kNch + j lives in lane g*kNch + j
        const long long own = valid ? (static_cast<long long>(n) * p.H + h) * p.W + w : -1;
#pragma unroll
        for (int j = 0; j < kNch; ++j) {
          const long long pix = __shfl_sync(0xffffffffu, own, (lane & ~(kNch - 1)) + j);
          if (pix >= 0) reinterpret_cast<uint4*>(p.out)[pix * 8 + j0 + ln] = o[j];
        }
      } else {
        uint4* oimg = reinterpret_cast<uint4*>(p.out + n * p.img_bytes);
        if (in_plane) {
#pragma unroll
          for (int j = 0; j < kNch; ++j) oimg[(j0 + j) * plane16 + opos] = o[j];
        }
        // the zero border rows the flattened tiles do not reach
        if (live && ti == 0 && m < p.Wp + 1) {
#pragma unroll
          for (int j = 0; j < kNch; ++j) oimg[(j0 + j) * plane16 + m] = make_uint4(0, 0, 0, 0);
        }
        if (live && ti == p.tiles_per_img - 1 && m < p.Wp - 1) {
          const long long bpos = static_cast<long long>(p.H + 1) * p.Wp + 1 + m;
#pragma unroll
          for (int j = 0; j < kNch; ++j) oimg[(j0 + j) * plane16 + bpos] = make_uint4(0, 0, 0, 0);
        }
      }
      ++it;
    };
    for (int tile = blockIdx.x; alive(tile); tile += 2 * gridDim.x) {
      process(tile, rv0);
      if (alive(tile + static_cast<int>(gridDim.x))) process(tile + gridDim.x, rv1);
    }
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (PAIR) cluster_sync_all();  // the peer's TMEM / barriers stay alive until the leader's last MMA and commit
  if (warp == 1) {
    tc_fence_after();
    if constexpr (PAIR) tmem_dealloc2(tmem_base, 2 * kAccCols);
    else tmem_dealloc(tmem_base, 2 * kAccCols);
  }
}

__global__ void __launch_bounds__(p8_threads(8), 1) conv_p8_kernel(const ConvP8Params p) { conv_p8_body<8, false>(p); }
__global__ void __launch_bounds__(p8_threads(4), 1) conv_p8_kernel4(const ConvP8Params p) { conv_p8_body<4, false>(p); }
__global__ void __launch_bounds__(p8_threads(8), 1) conv_p8_x2_kernel(const ConvP8Params p) { conv_p8_body<8, true>(p); }
__global__ void __launch_bounds__(p8_threads(8), 1) conv_p8_pair_kernel(const ConvP8Params p) { conv_p8_body<8, false, true>(p); }
__global__ void __launch_bounds__(p8_threads(8), 1) conv_p8_x2_pair_kernel(const ConvP8Params p) { conv_p8_body<8, true, true>(p); }

constexpr int kMaxDynSmem = 227 * 1024 - 1024;  // the kernel also has ~400 B of static shared memory
inline int patch_pitch_for(int Wp) { return ((128 + 2 * Wp + 2) * 16 + 127) / 128 * 128; }
inline int smem_bytes_for(int Wp, bool x2, bool pair = false) {
  // single CTA: 4 stages of 8 (whole tile) / 4 (half tile, X2) chunk planes; pair: half the weights, 5 / 4 whole-tile stages
  if (pair) return (x2 ? 2 : 1) * kWBytes / 2 + (x2 ? 4 : 5) * kChunks * patch_pitch_for(Wp) + 128;
  return (x2 ? 2 : 1) * kWBytes + 4 * (x2 ? kChunks / 2 : kChunks) * patch_pitch_for(Wp) + 128;
}

static int g_p8_max_pairs[kMaxDevices][2] = {};

template <typename K>
int p8_pair_occupancy(K kernel, int smem, int* out) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_sms() / 2 * 2);
  cfg.blockDim = dim3(p8_threads(8));
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  PB_CUDA(cudaOccupancyMaxActiveClusters(&n, kernel, &cfg));
  *out = n;
  return PB200_OK;
}

}  // namespace

int conv_p8_init() {
  static bool done[kMaxDevices] = {};  // function attributes belong to the device's context
  const int dev = current_device();
  if (done[dev]) return PB200_OK;
  PB_CUDA(cudaFuncSetAttribute(conv_p8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
  PB_CUDA(cudaFuncSetAttribute(conv_p8_kernel4, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
  PB_CUDA(cudaFuncSetAttribute(conv_p8_x2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
  PB_CUDA(cudaFuncSetAttribute(conv_p8_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
  PB_CUDA(cudaFuncSetAttribute(conv_p8_x2_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
  // co-resident CTA pairs at the largest shared-memory footprint (the detector's 64 x 64 planes use less)
  if (int rc = p8_pair_occupancy(conv_p8_pair_kernel, kMaxDynSmem, &g_p8_max_pairs[dev][0])) return rc;
  if (int rc = p8_pair_occupancy(conv_p8_x2_pair_kernel, kMaxDynSmem, &g_p8_max_pairs[dev][1])) return rc;
  done[dev] = true;
  return PB200_OK;
}

// reverse = 1 walks the tile list back to front.  The layer1 tensors (140 MB each) are larger than L2, and every kernel
// streams through them: when a launch ends, L2 holds the part of its input / output it touched LAST.  A consumer that
// starts there (a backward walk after a forward producer, and vice versa) finds the first ~60 MB of its input -- and of
// its residual, which the previous launch read in the same direction -- in L2 instead of HBM.
int launch_conv_p8(const void* in_p8, const void* wimg, const float* bias, const void* residual_p8, void* out,
                   int out_nhwc, int relu, int B, int H, int W, cudaStream_t st, int x2, int reverse) {
  const int Wp = W + 2;
  // the CTA-pair kernels are the default where a pair fits (PB200_P8_NOPAIR=1: the single-CTA kernels)
  const int max_pairs = g_p8_max_pairs[current_device()][x2 ? 1 : 0];
  const bool pair = max_pairs > 0 && smem_bytes_for(Wp, x2 != 0, true) <= kMaxDynSmem && getenv("PB200_P8_NOPAIR") == nullptr &&
                    getenv("PB200_P8_EPI4") == nullptr;
  PB_CHECK(smem_bytes_for(Wp, x2 != 0, pair) <= kMaxDynSmem, "conv_p8: W = %d too wide for the resident-weight kernel", W);
  PB_CHECK(Wp - 1 <= 128, "conv_p8: W = %d too wide for the border writer", W);
  if (B == 0) return PB200_OK;
  ConvP8Params p;
  p.in = static_cast<const uint8_t*>(in_p8);
  p.wimg = static_cast<const uint8_t*>(wimg);
  p.bias = bias;
  p.residual = static_cast<const uint8_t*>(residual_p8);
  p.out = static_cast<uint8_t*>(out);
  p.H = H;
  p.W = W;
  p.Wp = Wp;
  p.tiles_per_img = (H * Wp + 127) / 128;
  p.num_tiles = B * p.tiles_per_img;
  p.npos = 128 + 2 * Wp + 2;
  p.patch_pitch = patch_pitch_for(Wp);
  p.relu = relu;
  p.out_nhwc = out_nhwc;
  p.reverse = reverse;
  p.plane_bytes = static_cast<long long>(p8_plane_elems(H, W)) * 16;
  p.img_bytes = p.plane_bytes * 8;
  int grid = p.num_tiles < num_sms() ? p.num_tiles : num_sms();
  if (pair) {
    const int tile_pairs = (p.num_tiles + 1) / 2;
    grid = 2 * (tile_pairs < max_pairs ? tile_pairs : max_pairs);
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  const bool narrow = !x2 && getenv("PB200_P8_EPI4") != nullptr;
  cfg.blockDim = dim3(p8_threads(narrow ? 4 : 8));
  cfg.dynamicSmemBytes = smem_bytes_for(Wp, x2 != 0, pair);
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  cfg.attrs = attr;
  cfg.numAttrs = 0;
  if (pair) {
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.numAttrs = 1;
  }
  cfg.numAttrs += pdl_launch_attrs(attr + cfg.numAttrs);
  if (pair && x2) PB_CUDA(cudaLaunchKernelEx(&cfg, conv_p8_x2_pair_kernel, p));
  else if (pair) PB_CUDA(cudaLaunchKernelEx(&cfg, conv_p8_pair_kernel, p));
  else if (x2) PB_CUDA(cudaLaunchKernelEx(&cfg, conv_p8_x2_kernel, p));
  else if (narrow) PB_CUDA(cudaLaunchKernelEx(&cfg, conv_p8_kernel4, p));
  else PB_CUDA(cudaLaunchKernelEx(&cfg, conv_p8_kernel, p));
  return PB200_OK;
}

}  // namespace pb200
