// Fused stem of the detector: conv1 7x7/2 (+ folded BN) + ReLU + maxpool 3x3/2 in ONE tcgen05 kernel.
// Replaces resnet.conv1 / bn1 / relu / maxpool (torchvision resnet.py:268-271; RGBD conv1 at
// perseus/detector/models.py:27-28) of KeypointCNN.forward (models.py:34-40).
//
// GEMM view per conv output row: D[128 pixels][64 ch] = A[128][224] * Wt[64][224]^T with
// K = (filter row r) x (8 taps) x (4 frame channels).  The frames buffer is [B][H+6][W+8][4] bf16 with
// zero margins, so for filter row r the A row of output pixel wo is the 64 contiguous bytes of frames
// row 2*ho + r starting at pixel 2*wo: consecutive output pixels are 16 B apart and OVERLAP.  That is
// exactly a no-swizzle K-major UMMA operand whose core matrices (8 rows x 16 B, contiguous) are LBO = 16 B
// apart along K and SBO = 128 B apart along M: the raw frame row in shared memory IS the im2col
// operand; nothing is expanded or re-read.
//
//   * warp 0 (one elected lane): streams frame row-PAIRS (4224 contiguous bytes) into a 16-slot ring with
//     cp.async.bulk (no tensor map).
//   * warp 1 (one elected lane): per iteration 10 tcgen05.mma M128 N128 K16 + 8 M128 N64 K16 = the conv-row pair
//     (2k, 2k+1) into one of two 128-column TMEM accumulator stages; the weights (28 KB image) stay
//     resident in shared memory for the CTA's lifetime.
//   * warps 2..17: epilogue.  Thread = conv column x 16 channels.  TMEM -> vertical running 3-max in fp32
//     registers on the raw accumulators; once per pooled row: +bias, ReLU, bf16x2 pack -> horizontal 3-max: the two
//     lanes of a column pair split the channel words of their pool centre, so one pair exchange + one shift by two
//     lanes per word do it (the epilogue is bound by the shuffle / shared-memory pipe, not by issue slots or latency:
//     16 warps x 16 channels measured the same as 8 x 32; halving the shuffles and vectorising the warp-boundary
//     hand-off took 0.03 ms off) -> both lanes store their half of the pooled pixel in the P8 layout (layout.h),
//     borders included.  Post-ReLU values are >= 0, so the pool's -inf padding can be 0 horizontally.
//   * work unit = (image, band of PB pooled rows); a band below the top recomputes one conv-row pair.
#include <stdlib.h>

#include "common.cuh"
#include "layout.h"

namespace pb200 {

namespace {

// epilogue warps: 4 TMEM lane quarters x (kEpiWarps / 4) channel groups; threads = producer warp + MMA warp + epilogue
constexpr int stem_threads(int epi_warps) { return 64 + 32 * epi_warps; }
constexpr int kU8Producers = 3;  // U8 kernels: warp 0 and the two warps behind the epilogue warps convert row-pairs (640 threads x 96 registers; 672 do not launch)
constexpr int stem_threads_u8(int epi_warps) { return stem_threads(epi_warps) + 32 * (kU8Producers - 1); }
constexpr int kRowBytes = 2112;           // (256 + 8) px * 4 ch * 2 B
constexpr int kPairBytes = 2 * kRowBytes; // 4224
// TMEM accumulator stages of 128 columns (power of two).  Four stages (all 512 columns) measured the same as two (0.1123 vs
// 0.1118 ms, r02): the pooling epilogue's throughput, not the MMA warp's look-ahead, bounds this kernel.
constexpr int kAcc = 2;
constexpr int kAccShift = 1;
constexpr int kRing = 16;                 // row-pair slots
constexpr int kBSbo = kStemK / 8 * 128;   // 3584: 8-channel groups of the weight image
constexpr int kSmemW = 0;
// Second copy of the weight image, 8 K-chunks (= two filter rows) further than a whole image: read through ONE N = 128
// descriptor (8-channel groups kBSbo apart), rows 64..127 of B are then the SAME channels two filter rows earlier.
constexpr int kSmemW2 = kStemWimgBytes + 8 * 128;             // 29696
constexpr int kSmemWSet = kSmemW2 + kStemWimgBytes;           // 58368: one weight set = image + shifted copy
// X2 ("fp16x2" mode: fp16 frames, weights = fp16 hi + fp16 lo): a second weight set (the lo image + its shifted copy)
// follows the first; every MMA is issued twice, once per set, into the same accumulator.
__host__ __device__ constexpr int smem_ring(bool x2) { return (x2 ? 2 : 1) * kSmemWSet; }
__host__ __device__ constexpr int smem_xchg(bool x2) { return smem_ring(x2) + kRing * kPairBytes; }  // + 67584
__host__ __device__ constexpr int smem_bias(bool x2) { return smem_xchg(x2) + 2 * 4 * 128; }         // + 1024 ([2][4 quarters][64 channels] 16-bit)
__host__ __device__ constexpr int smem_total(bool x2) { return smem_bias(x2) + 64 * 4 + 128; }       // + alignment slack

struct StemPoolParams {
  // U8 kernels: the stem reads the tight uint8 [B][H][256][3] batch itself (rgb8 != null, frames unused)
  const uint8_t* rgb8;
  int bgr;                  // U8: source pixels are B, G, R (scripts/streaming.py:69 reverses them)
  int reverse;              // walk the units from the last to the first (L2 reuse across launches: launch_conv_p8)
  int H;                    // U8: image rows
  const uint8_t* frames;
  const uint8_t* wimg;
  const float* bias;
  uint8_t* out;
  int num_units, bands_per_img, PB, Hpool;  // Hpool = pooled rows per image (64)
  long long frame_bytes, out_img_bytes;
};

// U8 = the ingest step fused into the producer (RGB uint8 batches, the BASELINE workload): warp 0 loads the image rows
// itself -- a lane takes 16 pixels = 48 contiguous bytes, the warp one frame row-PAIR per pass --, converts them to the
// frames encoding (bf16 k, or fp16 k * 2^-12: both exact, bit-identical to csrc/frames.cu::ingest_kernel), writes them
// into the ring slot and arrives on the slot's barrier.  The 141 MB frames tensor is neither written nor read and the
// ingest launch disappears.  The ring's left / right margins are zeroed once, margin ROWS are written as zeros.
// MEASURED (r02, bs 256, tools/gpu/r02_i.sh): bit-identical, 2-3 % faster than ingest + forward under the power cap (1.14
// vs 1.18 ms) but 8-12 % slower at full clocks (1.11-1.16 vs 1.03 ms) in every variant tried (one / three converting
// warps; 16-pixel lanes with conflicted stores, with a shuffle transpose, 2-pixel chunks without either): the pooling
// epilogue already bounds this kernel and the conversion competes with it.  Hence opt-in (pb200_forward_rgb8).
template <int kEpiWarps, bool X2, bool U8 = false>
__device__ __forceinline__ void stem_pool_body(const StemPoolParams& p) {
  constexpr int kCh = 64 / (kEpiWarps / 4);  // channels per epilogue thread
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kRing];
  __shared__ __align__(8) uint64_t empty_bar[kRing];
  __shared__ __align__(8) uint64_t w_bar;
  __shared__ __align__(8) uint64_t tmem_full_bar[kAcc];
  __shared__ __align__(8) uint64_t tmem_empty_bar[kAcc];
  __shared__ uint32_t tmem_base_slot;

  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 127u) & ~127u) - raw_addr);
  uint8_t* s_w = smem + kSmemW;
  uint8_t* s_ring = smem + smem_ring(X2);
  uint32_t* s_xchg = reinterpret_cast<uint32_t*>(smem + smem_xchg(X2));  // [2][4 lane quarters][4 channel quarters][8]
  float* s_bias = reinterpret_cast<float*>(smem + smem_bias(X2));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < kRing; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(&w_bar, 1);
#pragma unroll
    for (int i = 0; i < kAcc; ++i) {
      mbar_init(&tmem_full_bar[i], 1);
      mbar_init(&tmem_empty_bar[i], kEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(&tmem_base_slot, 128 * kAcc);
    tmem_relinquish();
  }
  if (threadIdx.x >= 64 && threadIdx.x < 128) s_bias[threadIdx.x - 64] = __ldg(p.bias + threadIdx.x - 64);
  if constexpr (U8) {
    // margins of every ring slot (4 pixels = 32 bytes left and right of both rows): zero for the CTA's lifetime (the
    // converting warps write the pixel area only)
    for (int i = threadIdx.x; i < kRing * 2 * 2 * 2; i += blockDim.x) {
      const int slot = i >> 3, r = (i >> 2) & 1, side = (i >> 1) & 1, h = i & 1;
      *reinterpret_cast<uint4*>(s_ring + slot * kPairBytes + r * kRowBytes + (side ? kRowBytes - 32 : 0) + h * 16) =
          make_uint4(0, 0, 0, 0);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  pdl_launch_dependents();  // layer1's first kernel may start its prologue on SMs this grid has left
  pdl_wait();               // the frames come from the ingest kernel (weights / bias above are plan constants)

  // Work decomposition shared by the three roles.  A unit = (image n, band of PB pooled rows).  It runs
  // `n_it` iterations; iteration k (k0 <= k < k0 + n_it) computes the conv-row pair (2k, 2k+1).  Pooled row k =
  // max(conv rows 2k-1, 2k, 2k+1), so a band that does not start at the top of the image runs one extra
  // leading iteration (k0 = p0 - 1) whose only product is conv row 2k+1 (the carried "previous odd row").
  // Conv rows 2k, 2k+1 read frame rows 4k .. 4k+8, i.e. frame row-pairs 2k .. 2k+4; iteration k releases pairs
  // 2k and 2k+1.  Per unit the ring therefore sees 2*n_it + 3 pairs.
  constexpr int kFirstExtra = 2 + kEpiWarps;  // U8: warps kFirstExtra.. are producers 1..kU8Producers-1
  if (warp == 0 || (U8 && warp >= kFirstExtra)) {
    // ================================ producer ====================================
    // whole warp runs the loop (warp-uniform operands); one elected lane issues the copies
    if (warp == 0 && elect_one()) {
      mbar_expect_tx(&w_bar, (X2 ? 4 : 2) * kStemWimgBytes);
      bulk_load_1d(s_w, p.wimg, kStemWimgBytes, &w_bar);
      bulk_load_1d(s_w + kSmemW2, p.wimg, kStemWimgBytes, &w_bar);
      if (X2) {  // the lo image follows the hi image in global memory
        bulk_load_1d(s_w + kSmemWSet, p.wimg + kStemWimgBytes, kStemWimgBytes, &w_bar);
        bulk_load_1d(s_w + kSmemWSet + kSmemW2, p.wimg + kStemWimgBytes, kStemWimgBytes, &w_bar);
      }
    }
    uint32_t G = 0;
    if constexpr (U8) {
      // ---- fused ingest: kU8Producers warps, warp pi converts the row-pairs G = pi, pi + 3, ... ----
      // (ONE warp could not keep up: ~1400 cycles of dependent work per pair against the ~800 the kernel spends on one)
      const int pi = warp == 0 ? 0 : warp - kFirstExtra + 1;
      const int lrow = lane >> 4;   // row of the pair this lane converts
      const int lx = lane & 15;     // its 16-pixel group
      struct Pos {                  // a row-pair of the CTA's work list
        int u, g, n_pairs, y0;      // unit, pair of the unit, pairs of the unit, image row of the unit's first frame row
        uint32_t G;                 // running pair count (ring slot / phase)
        const uint8_t* img;
      };
      auto unit_pos = [&](int u, uint32_t G) {
        Pos q;
        q.u = u;
        q.g = 0;
        q.n_pairs = 0;
        q.y0 = 0;
        q.G = G;
        q.img = p.rgb8;
        if (u < p.num_units) {
          const int ur = p.reverse ? p.num_units - 1 - u : u;  // q.u counts positions of the walk, ur is the unit there
          const int n = ur / p.bands_per_img;
          const int band = ur - n * p.bands_per_img;
          const int k0 = band == 0 ? 0 : band * p.PB - 1;
          q.n_pairs = 2 * (p.PB + (band == 0 ? 0 : 1)) + 3;
          q.y0 = 4 * k0 - kFrameMarginTop;
          q.img = p.rgb8 + static_cast<long long>(n) * p.H * (256 * 3);
        }
        return q;
      };
      auto step = [&](Pos& q, int n) {
        for (int k = 0; k < n && q.u < p.num_units; ++k) {
          ++q.G;
          if (++q.g == q.n_pairs) q = unit_pos(q.u + gridDim.x, q.G);
        }
      };
      // Lane -> row lrow of the pair, and in it the 2-pixel chunks c = lx + 16 j (j = 0..7): for every j the 16 lanes of a
      // row write 256 contiguous bytes of the slot (conflict-free stores without any shuffle; both were tried: a lane owning
      // 16 consecutive pixels stores with 32-way bank conflicts, and transposing through shuffles loads the pipe this
      // kernel's epilogue is bound by).  A chunk's 6 source bytes start at 6 c: 4-byte aligned for even c, 2 mod 4 for odd c,
      // and the parity of c is the lane's -- so every lane issues ONE aligned 32-bit and ONE 16-bit load per chunk.
      const int odd = lx & 1;
      auto load_pair = [&](const Pos& q, uint32_t (&v)[8][2]) {
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) v[jj][0] = v[jj][1] = 0u;
        if (q.u < p.num_units) {
          const int y = q.y0 + 2 * q.g + lrow;
          if (y >= 0 && y < p.H) {
            const uint8_t* row = q.img + static_cast<long long>(y) * (256 * 3) + 6 * lx;
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) {
              const uint8_t* b = row + 96 * jj;
              v[jj][0] = __ldg(reinterpret_cast<const uint32_t*>(b + (odd ? 2 : 0)));
              v[jj][1] = __ldg(reinterpret_cast<const unsigned short*>(b + (odd ? 0 : 4)));
            }
          }
        }
      };
      constexpr int kAhead = 2;     // pairs of THIS warp in flight (6 of the CTA)
      uint32_t buf[kAhead][8][2];
      Pos ld = unit_pos(blockIdx.x, 0);
      step(ld, pi);
      Pos cv = ld;                  // the pair being converted
#pragma unroll
      for (int i = 0; i < kAhead; ++i) {
        load_pair(ld, buf[i]);
        step(ld, kU8Producers);
      }
      while (cv.u < p.num_units) {
        {
          const uint32_t slot = cv.G % kRing, ph = (cv.G / kRing) & 1u;
          uint32_t cur[8][2];
#pragma unroll
          for (int jj = 0; jj < 8; ++jj) {
            cur[jj][0] = buf[0][jj][0];
            cur[jj][1] = buf[0][jj][1];
          }
#pragma unroll
          for (int i = 0; i + 1 < kAhead; ++i)
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) {
              buf[i][jj][0] = buf[i + 1][jj][0];
              buf[i][jj][1] = buf[i + 1][jj][1];
            }
          load_pair(ld, buf[kAhead - 1]);  // this warp's pair kAhead further on: in flight while this one is converted
          step(ld, kU8Producers);
          mbar_wait(&empty_bar[slot], ph ^ 1u);
          tc_fence_after();  // the MMAs that read this slot (async proxy) are complete
          uint4* dst = reinterpret_cast<uint4*>(s_ring + slot * kPairBytes + lrow * kRowBytes + kFrameMarginLeft * 8) + lx;
          // Exact uint8 -> frames encoding at full instruction rate (I2F runs on a quarter-rate pipe):
          //   bf16: as_float(0x4B000000 | k) - 2^23 = float(k) exactly; two of them -> cvt.rn.bf16x2 (exact: k < 256)
          //   fp16: as_half2(0x6400 | k) = 1024 + k; one HFMA2 x * 2^-12 - 0.25 = k * 2^-12 exactly
#pragma unroll
          for (int jj = 0; jj < 8; ++jj) {
            // bytes r0 g0 b0 r1 | g1 b1 of the chunk
            const uint32_t w0 = odd ? (cur[jj][1] | (cur[jj][0] << 16)) : cur[jj][0];
            const uint32_t w1 = odd ? (cur[jj][0] >> 16) : cur[jj][1];
            uint32_t o[4];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              // pixel h: bytes 3h, 3h+1, 3h+2 of (w0, w1); selectors 0..3 pick from w0, 4..7 from w1
              const int b0 = 3 * h;
              const int sr = p.bgr ? b0 + 2 : b0, sg = b0 + 1, sb = p.bgr ? b0 : b0 + 2;
              auto byte_of = [&](int sel) -> uint32_t { return __byte_perm(w0, w1, 0x4440 | sel) & 0xffu; };
              if constexpr (X2) {
                const uint32_t rg = 0x64006400u | byte_of(sr) | (byte_of(sg) << 16);
                const uint32_t bz = 0x64006400u | byte_of(sb);
                const __half2 sc = __float2half2_rn(kF16RgbScale), off = __float2half2_rn(-1024.0f * kF16RgbScale);
                const __half2 vrg = __hfma2(*reinterpret_cast<const __half2*>(&rg), sc, off);
                const __half2 vbz = __hfma2(*reinterpret_cast<const __half2*>(&bz), sc, off);
                o[2 * h] = *reinterpret_cast<const uint32_t*>(&vrg);
                o[2 * h + 1] = *reinterpret_cast<const uint32_t*>(&vbz);
              } else {
                const float fr = __uint_as_float(0x4B000000u | byte_of(sr)) - 8388608.0f;
                const float fg = __uint_as_float(0x4B000000u | byte_of(sg)) - 8388608.0f;
                const float fb = __uint_as_float(0x4B000000u | byte_of(sb)) - 8388608.0f;
                o[2 * h] = pack_bf16x2(fr, fg);
                o[2 * h + 1] = pack_bf16x2(fb, 0.f);
              }
            }
            dst[16 * jj] = make_uint4(o[0], o[1], o[2], o[3]);
          }
          // generic-proxy writes -> visible to the tensor core's (async proxy) operand reads, then the slot is full
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(&full_bar[slot]);
        }
        step(cv, kU8Producers);
      }
    } else
    for (int ui = blockIdx.x; ui < p.num_units; ui += gridDim.x) {
      const int u = p.reverse ? p.num_units - 1 - ui : ui;
      const int n = u / p.bands_per_img;
      const int band = u - n * p.bands_per_img;
      const int k0 = band == 0 ? 0 : band * p.PB - 1;
      const int n_pairs = 2 * (p.PB + (band == 0 ? 0 : 1)) + 3;
      const uint8_t* src = p.frames + n * p.frame_bytes + static_cast<long long>(4 * k0) * kRowBytes;
      for (int g = 0; g < n_pairs; ++g, ++G) {
        const uint32_t slot = G % kRing, ph = (G / kRing) & 1u;
        mbar_wait(&empty_bar[slot], ph ^ 1u);
        if (elect_one()) {
          mbar_expect_tx(&full_bar[slot], kPairBytes);
          bulk_load_1d(s_ring + slot * kPairBytes, src + static_cast<long long>(g) * kPairBytes, kPairBytes,
                       &full_bar[slot]);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    // Frame row fr of the iteration's nine (relative to 4k) feeds conv row 2k through filter row fr and conv row 2k+1
    // through filter row fr - 2.  For fr = 2..6 both exist: ONE N = 128 MMA (B = [W_fr ; W_fr-2] via the shifted second
    // weight copy) fills accumulator columns [0,64) and [64,128) from a single read of the A rows -- the N = 64 MMA is
    // bound by its shared-memory operand fetch (~85 cycles measured for the unaligned no-swizzle rows), not by math.
    // Those five go first (the first one initialises all 128 columns); fr = 0, 1 (conv row 2k only) and fr = 7, 8
    // (conv row 2k+1 only) follow as N = 64 accumulations: 18 MMAs per pair instead of 28.
    constexpr uint32_t idesc64 = umma_idesc_16<X2>(128, 64);
    constexpr uint32_t idesc128 = umma_idesc_16<X2>(128, 128);
    const uint32_t tm = warp_uniform(tmem_base);
    const uint32_t a_hi = desc_hi_interleaved(128);
    const uint32_t b_hi = desc_hi_interleaved(kBSbo);
    const uint32_t b_lo0 = desc_lo(smem_u32(s_w), 128);
    const uint32_t ring_addr = smem_u32(s_ring);
    mbar_wait(&w_bar, 0);
    uint32_t G = 0;
    int it = 0;
    for (int ui = blockIdx.x; ui < p.num_units; ui += gridDim.x) {
      const int u = p.reverse ? p.num_units - 1 - ui : ui;
      const int band = u % p.bands_per_img;
      const int n_it = p.PB + (band == 0 ? 0 : 1);
      for (uint32_t j = 0; j < 3; ++j) mbar_wait(&full_bar[(G + j) % kRing], ((G + j) / kRing) & 1u);
      for (int i = 0; i < n_it; ++i, ++it) {
        const uint32_t Gi = G + 2 * i;
        mbar_wait(&full_bar[(Gi + 3) % kRing], ((Gi + 3) / kRing) & 1u);
        mbar_wait(&full_bar[(Gi + 4) % kRing], ((Gi + 4) / kRing) & 1u);
        const int acc = it & (kAcc - 1);
        mbar_wait(&tmem_empty_bar[acc], ((it >> kAccShift) & 1u) ^ 1u);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t d_tmem = tm + acc * 128;
          {
#pragma unroll
            for (int j = 0; j < 9; ++j) {
              const int fr = j < 5 ? j + 2 : (j < 7 ? j - 5 : j);  // 2 3 4 5 6 | 0 1 | 7 8
              const uint32_t row = ring_addr + ((Gi + (fr >> 1)) % kRing) * kPairBytes + (fr & 1) * kRowBytes;
              const uint32_t a_lo = desc_lo(row, 16);
              const int r = fr <= 6 ? fr : fr - 2;               // filter row of the FIRST 64 B rows
              const uint32_t d = d_tmem + (fr > 6 ? 64 : 0);
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                umma_bf16_split(d, a_lo + 2 * h, a_hi, b_lo0 + (r * 4 + h * 2) * 8, b_hi, j < 5 ? idesc128 : idesc64,
                                (j | h) != 0 ? 1u : 0u);
                if (X2)  // the same frame rows x the lo weight set
                  umma_bf16_split(d, a_lo + 2 * h, a_hi, b_lo0 + (kSmemWSet >> 4) + (r * 4 + h * 2) * 8, b_hi,
                                  j < 5 ? idesc128 : idesc64, 1u);
              }
            }
          }
          umma_commit(&tmem_full_bar[acc]);
          umma_commit(&empty_bar[Gi % kRing]);
          umma_commit(&empty_bar[(Gi + 1) % kRing]);
        }
        __syncwarp();
      }
      if (elect_one()) {
        for (uint32_t j = 0; j < 3; ++j) umma_commit(&empty_bar[(G + 2 * n_it + j) % kRing]);
      }
      __syncwarp();
      G += 2 * n_it + 3;
    }
  } else if (warp < kFirstExtra) {
    // ================================ epilogue ====================================
    // 16 warps: warp -> (TMEM lane quarter q, channel quarter cq).  Thread = conv column x 16 channels.
    // Per iteration: both conv rows of the pair come out of TMEM, the vertical 3-max (with the carried odd
    // row) runs in fp32 on the RAW accumulators, then bias, ReLU, the bf16 rounding and the horizontal 3-max
    // happen once per POOLED row.  Exact: x -> bf16(relu(x + b)) is monotone, so it commutes with max.
    const int q = warp & 3;                // TMEM lane quarter
    const int cq = (warp - 2) >> 2;        // channels [kCh*cq, kCh*cq + kCh)
    const int wo = q * 32 + lane;          // conv output column
    constexpr int Wp8 = 66;                // padded pooled width
    constexpr int kPlanes = kCh / 8;       // P8 chunk planes per thread
    const long long plane = static_cast<long long>(p.Hpool + 2) * Wp8;  // elements (16 B) per chunk plane
    const float* bias = s_bias + kCh * cq;
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + cq * kCh;
    int it = 0;
    uint32_t fin = 0;
    for (int ui = blockIdx.x; ui < p.num_units; ui += gridDim.x) {
      const int u = p.reverse ? p.num_units - 1 - ui : ui;
      const int n = u / p.bands_per_img;
      const int band = u - n * p.bands_per_img;
      const int k0 = band == 0 ? 0 : band * p.PB - 1;
      const int n_it = p.PB + (band == 0 ? 0 : 1);
      uint4* out_img = reinterpret_cast<uint4*>(p.out + n * p.out_img_bytes) + kPlanes * cq * plane;
      // top / bottom zero border rows of the P8 tensor
      if (band == 0 && wo < Wp8) {
#pragma unroll
        for (int j = 0; j < kPlanes; ++j) out_img[j * plane + wo] = make_uint4(0, 0, 0, 0);
      }
      if (band == p.bands_per_img - 1 && wo < Wp8) {
#pragma unroll
        for (int j = 0; j < kPlanes; ++j)
          out_img[j * plane + static_cast<long long>(p.Hpool + 1) * Wp8 + wo] = make_uint4(0, 0, 0, 0);
      }
      float podd[kCh];
#pragma unroll
      for (int k = 0; k < kCh; ++k) podd[k] = -INFINITY;
      for (int i = 0; i < n_it; ++i, ++it) {
        const int a = it & (kAcc - 1);
        mbar_wait(&tmem_full_bar[a], (it >> kAccShift) & 1u);
        tc_fence_after();
        uint32_t xe[kCh], xo[kCh];
        if constexpr (kCh == 16) {
          tmem_ld_32x16(t_lane + a * 128, xe);
          tmem_ld_32x16(t_lane + a * 128 + 64, xo);
          tmem_ld_wait_regs16(xe);
          tmem_ld_wait_regs16(xo);
        } else {
          tmem_ld_32x32(t_lane + a * 128, xe);
          tmem_ld_32x32(t_lane + a * 128 + 64, xo);
          tmem_ld_wait_regs(xe);
          tmem_ld_wait_regs(xo);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty_bar[a]);
        const bool owned = i > 0 || band == 0;  // the leading iteration of a lower band only produces podd
        uint32_t v[kCh / 2];
        if (owned) {
#pragma unroll
          for (int k = 0; k < kCh / 2; ++k) {
            const float2 b = *reinterpret_cast<const float2*>(bias + 2 * k);
            const float m0 = fmaxf(fmaxf(podd[2 * k], __uint_as_float(xe[2 * k])), __uint_as_float(xo[2 * k]));
            const float m1 = fmaxf(fmaxf(podd[2 * k + 1], __uint_as_float(xe[2 * k + 1])), __uint_as_float(xo[2 * k + 1]));
            v[k] = pack16x2<X2>(fmaxf(m0 + b.x, 0.f), fmaxf(m1 + b.y, 0.f));
          }
        }
#pragma unroll
        for (int k = 0; k < kCh; ++k) podd[k] = __uint_as_float(xo[k]);
        if (!owned) continue;
        // horizontal 3-max.  Pool centres are the EVEN columns; the two lanes of a column pair (c, c+1) share the
        // work of centre c: the even lane finishes the first half of this thread's channel words, the odd lane the
        // second half, so one pair exchange + one shift by two lanes per word replace a shuffle up and a shuffle down
        // of every word (the epilogue was bound by the shuffle / shared-memory pipe).  The warp's last column goes to
        // the next warp through shared memory (vector accesses by the two lanes that need it).
        constexpr int kW = kCh / 2, kHW = kW / 2;  // bf16x2 words per thread / per finishing lane
        uint32_t* xw = s_xchg + (((fin & 1) * 4 + q) * (kEpiWarps / 4) + cq) * kW;
        if (lane == 31) {
#pragma unroll
          for (int k = 0; k < kW / 4; ++k)
            reinterpret_cast<uint4*>(xw)[k] = make_uint4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
        }
        const bool odd = (lane & 1) != 0;
        uint32_t mine[kHW], part[kHW];
#pragma unroll
        for (int j = 0; j < kHW; ++j) {
          const uint32_t send = odd ? v[j] : v[kHW + j];  // what the pair partner finishes
          part[j] = __shfl_xor_sync(0xffffffffu, send, 1);
          mine[j] = odd ? v[kHW + j] : v[j];
        }
        // one barrier over ALL epilogue warps, although only the four lane-quarter warps of a channel group exchange
        // columns: per-group 128-thread barriers measured 13 % SLOWER (0.1265 vs 0.1118 ms, r02) -- the groups drift apart and
        // the accumulator stage is released by the slowest one
        named_bar_sync(1, 32 * kEpiWarps);
        uint32_t edge[kHW];
#pragma unroll
        for (int j = 0; j < kHW; ++j) edge[j] = 0u;  // column -1 is the pool's padding (values are >= 0)
        if (lane < 2 && q > 0) {
          const uint4* xr = reinterpret_cast<const uint4*>(s_xchg + (((fin & 1) * 4 + q - 1) * (kEpiWarps / 4) + cq) * kW + (odd ? kHW : 0));
#pragma unroll
          for (int k = 0; k < kHW / 4; ++k) {
            const uint4 t = xr[k];
            edge[4 * k] = t.x; edge[4 * k + 1] = t.y; edge[4 * k + 2] = t.z; edge[4 * k + 3] = t.w;
          }
        }
        ++fin;
        // column c-1 of the words a lane finishes sits two lanes to the left: as `part` of an even lane, `mine` of an odd
#pragma unroll
        for (int j = 0; j < kHW; ++j) {
          uint32_t left = __shfl_up_sync(0xffffffffu, odd ? mine[j] : part[j], 2);
          if (lane < 2) left = edge[j];
          mine[j] = max16x2<X2>(max16x2<X2>(left, mine[j]), part[j]);
        }
        const int po = k0 + i;
        uint4* orow = out_img + static_cast<long long>(po + 1) * Wp8;
        const int pw = wo >> 1;
        constexpr int kLanePlanes = kHW / 4;  // P8 chunk planes a lane stores
#pragma unroll
        for (int j = 0; j < kLanePlanes; ++j)
          orow[((odd ? kLanePlanes : 0) + j) * plane + pw + 1] =
              make_uint4(mine[4 * j], mine[4 * j + 1], mine[4 * j + 2], mine[4 * j + 3]);
        if (wo == 1 || wo == 127) {
          const int col = wo == 1 ? 0 : Wp8 - 1;
#pragma unroll
          for (int j = 0; j < kPlanes; ++j) orow[j * plane + col] = make_uint4(0, 0, 0, 0);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128 * kAcc);
  }
}

// 16 epilogue warps x 16 channels (18 warps: 5 on one scheduler, so at most 96 registers per thread) or 8 x 32
__global__ void __maxnreg__(96) stem_pool_kernel16(const StemPoolParams p) { stem_pool_body<16, false>(p); }
__global__ void __launch_bounds__(stem_threads(8), 1) stem_pool_kernel8(const StemPoolParams p) { stem_pool_body<8, false>(p); }
__global__ void __maxnreg__(96) stem_pool_x2_kernel16(const StemPoolParams p) { stem_pool_body<16, true>(p); }
__global__ void __maxnreg__(96) stem_pool_u8_kernel16(const StemPoolParams p) { stem_pool_body<16, false, true>(p); }
__global__ void __maxnreg__(96) stem_pool_u8_x2_kernel16(const StemPoolParams p) { stem_pool_body<16, true, true>(p); }

}  // namespace

int stem_pool_init() {
  static bool done[kMaxDevices] = {};  // function attributes belong to the device's context
  const int dev = current_device();
  if (done[dev]) return PB200_OK;
  PB_CUDA(cudaFuncSetAttribute(stem_pool_kernel16, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_total(false)));
  PB_CUDA(cudaFuncSetAttribute(stem_pool_kernel8, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_total(false)));
  PB_CUDA(cudaFuncSetAttribute(stem_pool_x2_kernel16, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_total(true)));
  PB_CUDA(cudaFuncSetAttribute(stem_pool_u8_kernel16, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_total(false)));
  PB_CUDA(cudaFuncSetAttribute(stem_pool_u8_x2_kernel16, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_total(true)));
  done[dev] = true;
  return PB200_OK;
}

int launch_stem_pool(const void* frames, const void* wimg, const float* bias, void* out_p8, int B, int H,
                     int W, cudaStream_t st, int x2, const void* rgb8, int bgr, int reverse) {
  PB_CHECK(W == 256 && H % 4 == 0 && H >= 8, "stem_pool: needs W = 256 and H %% 4 == 0 (got %dx%d)", H, W);
  PB_CHECK(rgb8 == nullptr || (reinterpret_cast<uintptr_t>(rgb8) & 15) == 0, "stem_pool: the uint8 batch must be 16-byte aligned");
  if (B == 0) return PB200_OK;
  const int Hpool = H / 4;
  // pooled rows per band: the largest power of two that still gives every SM ~2 units
  int PB = 16;
  while (PB > 1 && (Hpool % PB != 0 || static_cast<long long>(B) * (Hpool / PB) < 2LL * num_sms())) PB >>= 1;
  StemPoolParams p;
  p.rgb8 = static_cast<const uint8_t*>(rgb8);
  p.bgr = bgr;
  p.reverse = reverse ? 1 : 0;
  p.H = H;
  p.frames = static_cast<const uint8_t*>(frames);
  p.wimg = static_cast<const uint8_t*>(wimg);
  p.bias = bias;
  p.out = static_cast<uint8_t*>(out_p8);
  p.PB = PB;
  p.Hpool = Hpool;
  p.bands_per_img = Hpool / PB;
  p.num_units = B * p.bands_per_img;
  p.frame_bytes = static_cast<long long>(H + 6) * kRowBytes;
  p.out_img_bytes = static_cast<long long>(p8_image_bytes(64, Hpool, 64));
  const int grid = p.num_units < num_sms() ? p.num_units : num_sms();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  const bool wide = !x2 && rgb8 == nullptr && getenv("PB200_STEM_EPI8") != nullptr;
  cfg.blockDim = dim3(rgb8 != nullptr ? stem_threads_u8(16) : stem_threads(wide ? 8 : 16));
  cfg.dynamicSmemBytes = smem_total(x2 != 0);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  cfg.attrs = attr;
  cfg.numAttrs = pdl_launch_attrs(attr);
  if (rgb8 != nullptr && x2) PB_CUDA(cudaLaunchKernelEx(&cfg, stem_pool_u8_x2_kernel16, p));
  else if (rgb8 != nullptr) PB_CUDA(cudaLaunchKernelEx(&cfg, stem_pool_u8_kernel16, p));
  else if (x2) PB_CUDA(cudaLaunchKernelEx(&cfg, stem_pool_x2_kernel16, p));
  else if (wide) PB_CUDA(cudaLaunchKernelEx(&cfg, stem_pool_kernel8, p));
  else PB_CUDA(cudaLaunchKernelEx(&cfg, stem_pool_kernel16, p));
  return PB200_OK;
}

}  // namespace pb200
