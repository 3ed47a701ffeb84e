// CTA-pair (tcgen05 cta_group::2) variant of the TMA-im2col implicit-GEMM convolution of conv_umma.cu.
// Same GEMM view, tensor maps and epilogue; what changes is who holds what:
//   * a cluster of two CTAs computes a 256 x BLOCK_N tile.  Each CTA TMA-loads the activation tile of ITS 128
//     output pixels and only HALF of the weight tile (BLOCK_N/2 rows) into its own shared memory; one
//     tcgen05.mma.cta_group::2 issued by the pair leader reads A and the B halves from both CTAs and writes
//     the accumulator rows of each CTA into that CTA's TMEM.
//   * why: these layers are bound by the bytes that must LAND in each SM's shared memory per MMA cycle
//     (smem capacity / L2 latency, tools/tma_bench.cu): per 64-deep k-block a CTA now receives 16 KB + BLOCK_N/2 rows
//     instead of 16 KB + BLOCK_N rows, and the smaller stage buys two more pipeline stages.
//   * barriers: both producers credit the LEADER's full barrier (remote arrive.expect_tx + TMA completion bytes);
//     the leader's tcgen05.commit multicasts the "stage free" / "accumulator ready" arrivals to both CTAs; the
//     peer's epilogue warps release the accumulator on the leader's barrier through its cluster address.
#include <stdlib.h>

#include "common.cuh"
#include "conv_umma.cuh"
#include "umma_epilogue.cuh"

namespace pb200 {

constexpr int kPairThreads = 352;  // 2 TMA warps + MMA warp + 8 epilogue warps (per CTA)

// X2 = the "fp16x2" mode (conv_umma.cu MODE 2): fp16 operands, every stage carries this CTA's half of the hi AND of the
// lo weight tile; both are multiplied with the same activation tile.
template <int BLOCK_N, bool X2 = false>
struct PairCfg {
  static constexpr int kABytes = 128 * 128;
  static constexpr int kBHalf = BLOCK_N / 2 * 128;  // this CTA's half of one weight tile
  static constexpr int kBBytes = kBHalf * (X2 ? 2 : 1);
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (192 * 1024 / kStageBytes) > 8 ? 8 : (192 * 1024 / kStageBytes);
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024;
  static constexpr int kTmemCols = 2 * BLOCK_N;
};

// RESB = resident weights: when the conv has ONE N tile (BLOCK_N = Cout: layer2) and this CTA's half of the whole packed
// weight matrix fits next to an activation ring (18-19 k-blocks x 8 KB at Cout = 128), the producers load it ONCE per
// CTA and afterwards only activation tiles move: per 64-deep k-block 16 KB land in shared memory instead of 24 KB, on
// layers that are bound by shared-memory traffic (MMA operand reads + TMA fills), not by MMA issue.
//
// CHAIN = several convolutions in ONE persistent launch (conv_chain_kernel: the eight convs of layers 3-4).  The work list
// is the concatenation of the convs' tile lists, handed to the pairs round-robin: the tail of one conv fills up with the
// head of the next (3.46 / 1.73 waves per conv become 20.8 rounds for the whole list) and six launch ramps / tails go.
// A tile of conv c reads whole images of conv dep[c]'s output (Ho x Wo <= 256: an M-super = whole images); every
// epilogue warp bumps the counter of its (conv, M-super) after its stores, the producers of a dependent tile spin on it
// (same launch, all pairs co-resident, dependencies point backwards in the list: no deadlock), and the last CTA to
// finish zeroes the counters for the next launch.
constexpr int kChainBias = 3072;  // floats of bias of all convs of a chain (4 x 256 + 4 x 512)

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// (the form without a state space compiles to MEMBAR.ALL.GPU + FENCE.VIEW.ASYNC: in a TMA producer loop the MEMBAR waits for
// the warp's own bulk copies in flight and serialises the pipeline; .global is the FENCE.VIEW.ASYNC.G alone)
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async.global;" ::: "memory"); }

template <int BLOCK_N, bool X2, bool RESB, bool CHAIN>
__device__ __forceinline__ void pair_body(const UmmaConvMaps* __restrict__ mapsv, const UmmaConvParams* __restrict__ pv,
                                          const ChainCtl* __restrict__ ctl) {
  using Cfg = PairCfg<BLOCK_N, X2>;
  const UmmaConvParams& p = pv[0];   // prologue / RESB: the (first) conv
  const UmmaConvMaps& maps = mapsv[0];
  const int total_items = CHAIN ? ctl->total_items : p.num_super;
  // item -> (conv c, work item st of that conv)
  auto decode = [&](int item, int& c, int& st) {
    c = 0;
    st = (!CHAIN && p.reverse) ? p.num_super - 1 - item : item;
    if constexpr (CHAIN) {
      while (st >= pv[c].num_super) {
        st -= pv[c].num_super;
        ++c;
      }
    }
  };
  constexpr int kStages = RESB ? 8 : Cfg::kStages;  // RESB: p.a_stages (<= 8) activation stages, set by the host
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kStages];   // used in the leader CTA only
  __shared__ __align__(8) uint64_t empty_bar[kStages];  // per CTA: its stage buffers are free again
  __shared__ __align__(8) uint64_t tmem_full_bar[2];    // per CTA: accumulator stage complete
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];   // leader only: both CTAs' epilogues drained the stage
  __shared__ __align__(8) uint64_t resb_bar;            // leader only: both CTAs' resident weight halves have landed
  __shared__ uint32_t tmem_base_slot;
  __shared__ float s_bias[CHAIN ? kChainBias : 512];
  __shared__ float s_stat[CHAIN ? 2 : 2 * 512];  // training forward: per-CTA partial sum / sum of squares per output channel

  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* tiles = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  const int nst = RESB ? p.a_stages : kStages;  // stages of the ring
  const int stage_b = RESB ? 0 : Cfg::kBBytes;  // bytes of weights per ring stage
  uint8_t* smem_a = tiles;
  uint8_t* smem_b = tiles + nst * Cfg::kABytes;  // RESB: the resident weights, k-block kb at + kb * kBHalf

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int crank = static_cast<int>(cluster_ctarank());

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&maps.a[0]);
    tma_prefetch_desc(&maps.b);
#pragma unroll
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 2);   // one arrive.expect_tx per producer of the pair
      mbar_init(&empty_bar[i], 1);  // the leader's multicast commit
    }
    mbar_init(&tmem_full_bar[0], 1);
    mbar_init(&tmem_full_bar[1], 1);
    mbar_init(&tmem_empty_bar[0], 16);  // 8 epilogue warps x 2 CTAs
    mbar_init(&tmem_empty_bar[1], 16);
    mbar_init(&resb_bar, 2);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc2(&tmem_base_slot, Cfg::kTmemCols);
    tmem_relinquish2();
  }
  // PDL: the prologue above overlaps the previous kernel's tail; nothing below may run before it has completed
  pdl_launch_dependents();
  pdl_wait();
  if constexpr (CHAIN) {
    for (int c = 0; c < ctl->nconv; ++c)
      for (int i = threadIdx.x; i < pv[c].Cout; i += kPairThreads) s_bias[ctl->bias_off[c] + i] = __ldg(pv[c].bias + i);
  } else {
    for (int i = threadIdx.x; i < p.Cout; i += kPairThreads) s_bias[i] = __ldg(p.bias + i);
    if (p.stats != nullptr)
      for (int i = threadIdx.x; i < 2 * p.Cout; i += kPairThreads) s_stat[i] = 0.f;
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  const int pair_id = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;

  if (warp == 0 || warp == 10) {
    // two producer warps take alternate k-blocks: one warp's issue loop (barrier wait + expect_tx + 2 TMA issues)
    // costs about as many cycles as a 256-cycle k-block of MMAs (profiles/README.md r01)
    const int pi = warp == 0 ? 0 : 1;
    uint32_t g = 0;  // k-block counter of this CTA across all its tiles
    // ================================ TMA producer (both CTAs) ====================
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t lfull0 = mapa_u32(&full_bar[0], 0);
    if (RESB && pi == 0 && pair_id < p.num_super) {
      // this CTA's half (rows crank * BLOCK_N/2 ..) of EVERY k-block of the packed weights, once; k-blocks are stored in
      // the producers' walking order (taps x channel blocks, then the fused projection's blocks): columns kb * 64
      if (elect_one()) {
        const uint32_t lres = mapa_u32(&resb_bar, 0);
        mbar_expect_tx_cluster(lres, static_cast<uint32_t>(p.num_kblocks) * Cfg::kBHalf);
        for (int kb = 0; kb < p.num_kblocks; ++kb)
          if (p.b_tiled) tma_load_2d_2sm(smem_b + kb * Cfg::kBHalf, &maps.b, lres, 0, (kb * p.Cout + crank * (BLOCK_N / 2)) >> 2);
          else tma_load_2d_2sm(smem_b + kb * Cfg::kBHalf, &maps.b, lres, kb * 64, crank * (BLOCK_N / 2));
      }
      __syncwarp();
    }
    for (int item = pair_id; item < total_items; item += num_pairs) {
      int c, st;
      decode(item, c, st);
      const UmmaConvParams& p = pv[c];
      const UmmaConvMaps& maps = mapsv[c];
      const int m_super = st / p.num_n_tiles;
      const int n_tile = st - m_super * p.num_n_tiles;
      const int m_tile = m_super * 2 + crank;
      if constexpr (CHAIN) {
        // the images this tile reads must have been stored completely by conv dep[c] (same launch, other pairs)
        const int d = ctl->dep[c];
        if (d >= 0) {
          const int r = ctl->dep_ratio[c];
          const int nd = ctl->cnt_off[d + 1] - ctl->cnt_off[d];
          const unsigned want = static_cast<unsigned>(ctl->target[d]);
          for (int j = m_super * r; j < (m_super + 1) * r && j < nd; ++j) {
            const unsigned* cp = ctl->cnt + ctl->cnt_off[d] + j;
            unsigned spins = 0;
            while (ld_acquire_gpu(cp) < want) {
              __nanosleep(40);
              if (++spins > (1u << 24)) __trap();  // ~1 s: a lost arrival must not hang the GPU
            }
          }
          fence_proxy_async_all();  // the TMA loads below (async proxy) read what those generic-proxy stores wrote
        }
      }
      int n0, ho0;
      if (p.tiles_per_img > 0) {
        n0 = m_tile / p.tiles_per_img;
        ho0 = (m_tile - n0 * p.tiles_per_img) * p.bh;
      } else {
        n0 = m_tile * p.imgs_per_tile;
        ho0 = 0;
      }
      // the issue loop is kept as light as possible: one warp must feed a k-block every 256-512 MMA cycles
      const int n_row = n_tile * BLOCK_N + crank * (BLOCK_N / 2);
      // K order: 64-channel block outermost, filter taps inside (the same order csrc/conv_halo.cu walks its patches in, so
      // that a frame's result does not depend on which of the two kernels its batch size selects)
      int kcol = 0;
      for (int cb = 0; cb < p.cblocks; ++cb) {
        for (int tap = 0; tap < p.ntaps; ++tap) {
          const int dw = p.tap_dw[tap];
          const CUtensorMap* amap = &maps.a[p.tap_mi[tap]];
          const int hrow = ho0 + p.tap_dh[tap];
          kcol = p.tap_k[tap] + cb * 64;
          if ((g++ & 1u) == static_cast<uint32_t>(pi)) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          if (elect_one()) {
            const uint32_t lfull = lfull0 + stage * 8;  // the leader's full barrier of this stage
            mbar_expect_tx_cluster(lfull, Cfg::kABytes + stage_b);
            tma_load_4d_2sm(smem_a + stage * Cfg::kABytes, amap, lfull, cb * 64, dw, hrow, n0);
            if (!RESB) {
              if (p.b_tiled) {  // KT layout: this CTA's rows of k-block kcol/64 are one contiguous run of 512-byte rows
                const int row = ((kcol >> 6) * p.Cout + n_row) >> 2;
                tma_load_2d_2sm(smem_b + stage * Cfg::kBBytes, &maps.b, lfull, 0, row);
                if (X2) tma_load_2d_2sm(smem_b + stage * Cfg::kBBytes + Cfg::kBHalf, &maps.b, lfull, 0, row + ((p.kb_total * p.Cout) >> 2));
              } else {
                tma_load_2d_2sm(smem_b + stage * Cfg::kBBytes, &maps.b, lfull, kcol, n_row);
                if (X2) tma_load_2d_2sm(smem_b + stage * Cfg::kBBytes + Cfg::kBHalf, &maps.b, lfull, kcol + p.ktot, n_row);
              }
            }
          }
          __syncwarp();
          }
          if (++stage == nst) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
      // fused 1x1/2 projection of the block input: extra k-blocks whose A tile is the (2 ho, 2 wo) pixel of ds_in
      kcol = p.ntaps * p.cin;  // the projection's columns follow the R*S*Cin columns of the conv
      for (int cb = 0; cb < p.ds_cblocks; ++cb, kcol += 64) {
        if ((g++ & 1u) == static_cast<uint32_t>(pi)) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          if (elect_one()) {
            const uint32_t lfull = lfull0 + stage * 8;
            mbar_expect_tx_cluster(lfull, Cfg::kABytes + stage_b);
            tma_load_4d_2sm(smem_a + stage * Cfg::kABytes, &maps.ds, lfull, cb * 64, 0, ho0, n0);
            if (!RESB) {
              if (p.b_tiled) {
                const int row = ((kcol >> 6) * p.Cout + n_row) >> 2;
                tma_load_2d_2sm(smem_b + stage * Cfg::kBBytes, &maps.b, lfull, 0, row);
                if (X2) tma_load_2d_2sm(smem_b + stage * Cfg::kBBytes + Cfg::kBHalf, &maps.b, lfull, 0, row + ((p.kb_total * p.Cout) >> 2));
              } else {
                tma_load_2d_2sm(smem_b + stage * Cfg::kBBytes, &maps.b, lfull, kcol, n_row);
                if (X2) tma_load_2d_2sm(smem_b + stage * Cfg::kBBytes + Cfg::kBHalf, &maps.b, lfull, kcol + p.ktot, n_row);
              }
            }
          }
          __syncwarp();
        }
        if (++stage == nst) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (pair leader only) ===============
    if (crank == 0) {
      constexpr uint32_t idesc = umma_idesc_16<X2>(256, BLOCK_N);
      const uint32_t tm = warp_uniform(tmem_base);
      const uint32_t d_hi = desc_hi_swizzled(128);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      if (RESB && pair_id < p.num_super) mbar_wait(&resb_bar, 0);
      for (int item = pair_id; item < total_items; item += num_pairs, ++it) {
        int c, st;
        decode(item, c, st);
        const UmmaConvParams& p = pv[c];
        const int acc = it & 1;
        mbar_wait(&tmem_empty_bar[acc], ((it >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tm + acc * BLOCK_N;
        for (int kb = 0; kb < p.num_kblocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_lo = desc_lo(smem_u32(smem_a + stage * Cfg::kABytes), 16);
            // RESB: the resident k-blocks sit in weight order (tap-major); the walk is channel-block-major
            const int nreg = p.ntaps * p.cblocks;
            const int kbw = kb < nreg ? (kb % p.ntaps) * p.cblocks + kb / p.ntaps : kb;
            const uint32_t b_lo = desc_lo(smem_u32(smem_b + (RESB ? kbw * Cfg::kBHalf : stage * Cfg::kBBytes)), 16);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              umma2_bf16_split(d_tmem, a_lo + 2 * k, d_hi, b_lo + 2 * k, d_hi, idesc, (kb | k) != 0 ? 1u : 0u);
              if (X2) umma2_bf16_split(d_tmem, a_lo + 2 * k, d_hi, b_lo + 2 * k + (Cfg::kBHalf >> 4), d_hi, idesc, 1u);
            }
            umma2_commit_mcast(&empty_bar[stage], 3);  // both CTAs may refill this stage
            if (kb == p.num_kblocks - 1) umma2_commit_mcast(&tmem_full_bar[acc], 3);
          }
          __syncwarp();
          if (++stage == nst) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else {
    // ================================ epilogue (both CTAs) =========================
    constexpr int kCols = BLOCK_N / 2;
    constexpr int kChunks = kCols / 32;
    constexpr int kVec = kCols / 8;
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    int it = 0;
    for (int item = pair_id; item < total_items; item += num_pairs, ++it) {
      int c, st;
      decode(item, c, st);
      const UmmaConvParams& p = pv[c];
      const float* s_bias_c = CHAIN ? s_bias + ctl->bias_off[c] : s_bias;
      const RowMap rmap = {p.up2, p.up_ph, p.up_pw, p.up_lh, p.up_lw};
      const int m_super = st / p.num_n_tiles;
      const int n_tile = st - m_super * p.num_n_tiles;
      const int m_tile = m_super * 2 + crank;
      const int acc = it & 1;
      const long long gm = static_cast<long long>(m_tile) * 128 + row;
      const bool valid = gm < p.M_total;
      const int col0 = n_tile * BLOCK_N + half * kCols;
      uint4 rv[kVec];
      const bool has_res = p.residual != nullptr;
      if constexpr (kCols % 64 == 0) {
        // line-granular path (umma_epilogue.cuh): residual and output move as complete 128-byte lines
        const long long gm_group = static_cast<long long>(m_tile) * 128 + q * 32;
        if (has_res) {
          // CHAIN: the residual may have been written earlier in this launch by another SM: L2-coherent loads
          load_residual_lines<kCols, CHAIN>(static_cast<const __nv_bfloat16*>(p.residual), gm_group, p.M_total, p.Cout, col0, lane, rv, rmap);
        } else {
#pragma unroll
          for (int j = 0; j < kVec; ++j) rv[j] = make_uint4(0, 0, 0, 0);
        }
        mbar_wait(&tmem_full_bar[acc], (it >> 1) & 1u);
        tc_fence_after();
        const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BLOCK_N + half * kCols;
        epilogue_lines<kCols, X2>(t_addr, s_bias_c + col0, has_res, rv, static_cast<__nv_bfloat16*>(p.out), gm_group, p.M_total,
                              p.Cout, col0, p.relu, lane, [&]() {
                                tc_fence_before();
                                __syncwarp();
                                if (lane == 0) {
                                  if (crank == 0) mbar_arrive(&tmem_empty_bar[acc]);
                                  else mbar_arrive_cluster(mapa_u32(&tmem_empty_bar[acc], 0));
                                }
                              }, p.stats != nullptr ? s_stat : nullptr, rmap);
        if constexpr (CHAIN) {
          if (ctl->target[c] > 0) {  // a later conv of the chain reads this M-super: publish this warp's stores
            __threadfence();
            fence_proxy_async_all();
            __syncwarp();
            if (lane == 0) atomicAdd(ctl->cnt + ctl->cnt_off[c] + m_super, 1u);
          }
        }
        continue;
      }
      static_assert(!CHAIN || kCols % 64 == 0, "chained convs use the line-granular epilogue");
      if (has_res && valid) {
        const uint4* res = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.residual) + rmap(gm) * p.Cout + col0);
#pragma unroll
        for (int j = 0; j < kVec; ++j) rv[j] = __ldg(res + j);
      } else {
#pragma unroll
        for (int j = 0; j < kVec; ++j) rv[j] = make_uint4(0, 0, 0, 0);
      }
      mbar_wait(&tmem_full_bar[acc], (it >> 1) & 1u);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BLOCK_N + half * kCols;
      uint4* out_row = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + rmap(gm) * p.Cout + col0);
      const float* bias = s_bias_c + col0;
#pragma unroll
      for (int chunk = 0; chunk < kChunks; ++chunk) {
        uint4 oc[4];
        uint32_t v[32];
        tmem_ld_32x32(t_addr + chunk * 32, v);
        tmem_ld_wait_regs(v);
        if (chunk == kChunks - 1) {  // accumulator stage fully read by this warp: tell the leader's MMA warp
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
  if constexpr (!CHAIN) {
    if (p.stats != nullptr)  // fused batch statistics: one fp64 atomic per channel and CTA
      for (int i = threadIdx.x; i < 2 * p.Cout; i += kPairThreads)
        if (s_stat[i] != 0.f) atomicAdd(&p.stats[i], static_cast<double>(s_stat[i]));
  }
  cluster_sync_all();  // the peer's TMEM / barriers stay alive until the leader's last MMA and commit are done
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, Cfg::kTmemCols);
  }
  if constexpr (CHAIN) {
    // every wait of this CTA is behind it: the last CTA to get here re-arms the counters for the next launch
    __shared__ unsigned s_ticket;
    unsigned* fin = ctl->cnt + ctl->ncnt;
    if (threadIdx.x == 0) {
      __threadfence();
      s_ticket = atomicAdd(fin, 1u);
    }
    __syncthreads();
    if (s_ticket == gridDim.x - 1) {
      for (int i = threadIdx.x; i < ctl->ncnt; i += kPairThreads) ctl->cnt[i] = 0u;
      if (threadIdx.x == 0) *fin = 0u;
      __threadfence();
    }
  }
}

template <int BLOCK_N, bool X2, bool RESB>
__global__ void __launch_bounds__(kPairThreads, 1)
conv_umma2_kernel(const __grid_constant__ UmmaConvMaps maps, const UmmaConvParams p) {
  pair_body<BLOCK_N, X2, RESB, false>(&maps, &p, nullptr);
}

template <bool X2>
__global__ void __launch_bounds__(kPairThreads, 1) conv_chain_kernel(const __grid_constant__ ChainArgs A) {
  pair_body<256, X2, false, true>(A.maps, A.p, &A.ctl);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
template <int BLOCK_N, bool X2>
static int pair_cfg(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, int grid, cudaStream_t st, bool pdl = false) {
  *cfg = cudaLaunchConfig_t{};
  cfg->gridDim = dim3(grid);
  cfg->blockDim = dim3(kPairThreads);
  cfg->dynamicSmemBytes = PairCfg<BLOCK_N, X2>::kSmemBytes;
  cfg->stream = st;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg->attrs = attr;
  cfg->numAttrs = 1 + (pdl ? pdl_launch_attrs(attr + 1) : 0);
  return PB200_OK;
}

constexpr int kResbMaxSmem = 217 * 1024;  // resident-weights launches (build_umma_conv sizes them to <= 216 KB + 1 KB)

template <int BLOCK_N, bool X2>
static int pair_init_one(int* max_pairs) {
  PB_CUDA(cudaFuncSetAttribute(conv_umma2_kernel<BLOCK_N, X2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               PairCfg<BLOCK_N, X2>::kSmemBytes));
  if (!X2)
    PB_CUDA(cudaFuncSetAttribute(conv_umma2_kernel<BLOCK_N, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 kResbMaxSmem));
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[1];
  pair_cfg<BLOCK_N, X2>(&cfg, attr, num_sms() / 2 * 2, nullptr);
  int n = 0;
  PB_CUDA(cudaOccupancyMaxActiveClusters(&n, conv_umma2_kernel<BLOCK_N, X2, false>, &cfg));
  *max_pairs = n;
  return PB200_OK;
}

// per device (function attributes and occupancy belong to the device's context): [device][x2][BLOCK_N = 64, 128, 256]
static int g_max_pairs[kMaxDevices][2][3] = {};
static bool g_pair_done[kMaxDevices] = {};

int umma_pair_init() {
  const int dev = current_device();
  if (g_pair_done[dev]) return PB200_OK;
  if (int rc = pair_init_one<64, false>(&g_max_pairs[dev][0][0])) return rc;
  if (int rc = pair_init_one<128, false>(&g_max_pairs[dev][0][1])) return rc;
  if (int rc = pair_init_one<256, false>(&g_max_pairs[dev][0][2])) return rc;
  if (int rc = pair_init_one<64, true>(&g_max_pairs[dev][1][0])) return rc;
  if (int rc = pair_init_one<128, true>(&g_max_pairs[dev][1][1])) return rc;
  if (int rc = pair_init_one<256, true>(&g_max_pairs[dev][1][2])) return rc;
  g_pair_done[dev] = true;
  return PB200_OK;
}

static int g_chain_pairs[kMaxDevices][2] = {};

template <bool X2>
static int chain_cfg(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, int grid, cudaStream_t st, bool pdl) {
  pair_cfg<256, X2>(cfg, attr, grid, st, pdl);
  return PB200_OK;
}

template <bool X2>
static int chain_init_one(int* max_pairs) {
  PB_CUDA(cudaFuncSetAttribute(conv_chain_kernel<X2>, cudaFuncAttributeMaxDynamicSharedMemorySize, PairCfg<256, X2>::kSmemBytes));
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[2];
  chain_cfg<X2>(&cfg, attr, num_sms() / 2 * 2, nullptr, false);
  int n = 0;
  PB_CUDA(cudaOccupancyMaxActiveClusters(&n, conv_chain_kernel<X2>, &cfg));
  *max_pairs = n;
  return PB200_OK;
}

int build_umma_chain(ChainLaunch* C, const UmmaConvLaunch* const* convs, const int* dep, int n, unsigned* counters) {
  C->grid = 0;
  if (n < 2 || n > kMaxChain || counters == nullptr) return PB200_OK;
  if (int rc = umma_pair_init()) return rc;
  const int dev = current_device();
  const int x2 = convs[0]->split == 2 ? 1 : 0;
  if (g_chain_pairs[dev][x2] == 0) {
    if (int rc = x2 ? chain_init_one<true>(&g_chain_pairs[dev][1]) : chain_init_one<false>(&g_chain_pairs[dev][0])) return rc;
  }
  ChainCtl& ctl = C->args.ctl;
  ctl = ChainCtl{};
  int items = 0, cnt = 0, bias = 0;
  for (int i = 0; i < n; ++i) {
    const UmmaConvLaunch& L = *convs[i];
    const UmmaConvParams& p = L.p;
    // whole images per M-super, the pair<256> kernel, the line-granular epilogue, no training statistics
    if (!L.pair || L.resb || L.block_n != 256 || (L.split == 2) != (x2 != 0) || L.split == 1 || p.stats != nullptr || p.up2 ||
        p.num_tiles == 0 || 256 % (p.Ho * p.Wo) != 0)
      return PB200_OK;
    C->args.maps[i] = L.maps;
    C->args.p[i] = p;
    ctl.dep[i] = dep[i];
    ctl.cnt_off[i] = cnt;
    ctl.bias_off[i] = bias;
    const int m_supers = p.num_super / p.num_n_tiles;
    if (dep[i] >= 0) {
      if (dep[i] >= i) return PB200_OK;
      const UmmaConvParams& d = convs[dep[i]]->p;
      const int mine = p.Ho * p.Wo, theirs = d.Ho * d.Wo;  // output pixels per image
      if (theirs % mine != 0) return PB200_OK;
      ctl.dep_ratio[i] = theirs / mine;
      ctl.target[dep[i]] = 16 * d.num_n_tiles;
    }
    items += p.num_super;
    cnt += m_supers;
    bias += p.Cout;
  }
  ctl.cnt_off[n] = cnt;
  if (cnt + 1 > kChainCounters || bias > kChainBias) return PB200_OK;
  ctl.nconv = n;
  ctl.total_items = items;
  ctl.ncnt = cnt;
  ctl.cnt = counters;
  const int max_pairs = g_chain_pairs[dev][x2];
  // every pair must be resident at once: the dependency spins rely on it
  C->grid = 2 * (items < max_pairs ? items : max_pairs);
  C->x2 = x2;
  return PB200_OK;
}

int launch_umma_chain(const ChainLaunch& C, cudaStream_t st) {
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[2];
  if (C.x2) {
    chain_cfg<true>(&cfg, attr, C.grid, st, true);
    PB_CUDA(cudaLaunchKernelEx(&cfg, conv_chain_kernel<true>, C.args));
  } else {
    chain_cfg<false>(&cfg, attr, C.grid, st, true);
    PB_CUDA(cudaLaunchKernelEx(&cfg, conv_chain_kernel<false>, C.args));
  }
  return PB200_OK;
}

int umma_pair_max_pairs(int block_n, int x2) {
  return g_max_pairs[current_device()][x2 ? 1 : 0][block_n == 64 ? 0 : (block_n == 128 ? 1 : 2)];
}

template <int BLOCK_N, bool X2>
static int launch_pair_one(const UmmaConvLaunch& L, cudaStream_t st) {
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[2];
  pair_cfg<BLOCK_N, X2>(&cfg, attr, L.grid, st, true);
  if (!X2 && L.resb) {
    cfg.dynamicSmemBytes = L.smem;
    PB_CUDA(cudaLaunchKernelEx(&cfg, conv_umma2_kernel<BLOCK_N, false, true>, L.maps, L.p));
  } else {
    PB_CUDA(cudaLaunchKernelEx(&cfg, conv_umma2_kernel<BLOCK_N, X2, false>, L.maps, L.p));
  }
  return PB200_OK;
}

int launch_umma_pair(const UmmaConvLaunch& L, cudaStream_t st) {
  if (L.p.num_tiles == 0) return PB200_OK;
  if (L.split == 2) {
    if (L.block_n == 64) return launch_pair_one<64, true>(L, st);
    if (L.block_n == 128) return launch_pair_one<128, true>(L, st);
    return launch_pair_one<256, true>(L, st);
  }
  if (L.block_n == 64) return launch_pair_one<64, false>(L, st);
  if (L.block_n == 128) return launch_pair_one<128, false>(L, st);
  return launch_pair_one<256, false>(L, st);
}

}  // namespace pb200
