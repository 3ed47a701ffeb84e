// Host interface of the tcgen05 implicit-GEMM convolution (conv_umma.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace pb200 {

struct alignas(64) UmmaConvMaps {
  CUtensorMap a[4];  // activation maps: [0] for stride 1 / the stem, [hp*2+wp] parity maps for stride 2
  CUtensorMap b;     // packed weights [Cout][Kpad]
  CUtensorMap ds;    // fused 1x1/2 projection of the block input (extra K columns), parity (0,0) map; unused otherwise
};

struct UmmaConvParams {
  int num_tiles, num_n_tiles, num_kblocks;
  int cl;         // thread-block cluster size (1, 2 or 4): CTAs sharing one multicast weight tile
  int num_super;  // ceil(m_tiles / cl) * num_n_tiles: work items per cluster loop
  int M_total;  // B*Ho*Wo output pixels
  int Cout;
  int stem;     // unused (kept 0)
  int cin;      // split-bf16 mode: channel offset of the lo halves in the activation tensor (= Cin)
  int ktot;     // split-bf16 mode: column offset of the lo halves in the packed weights (= R*R*Cin)
  int Ho, Wo;
  int tiles_per_img;  // Ho*Wo/128 when >= 1, else 0
  int imgs_per_tile;  // 128/(Ho*Wo) when the image is smaller than a tile, else 1
  int bh;             // output rows per tile
  int stride, pad, S, cblocks;
  // the filter taps the producers walk, in k order: offset of the tap's A box relative to the tile's first output pixel
  // (in the tap's map: parity-split for stride 2), map index, first k column of the tap's weight block.  Filled by
  // build_umma_conv for an R x R conv; a stride-2 data gradient installs the 1 / 2 / 2 / 4 taps of one output parity.
  int ntaps;
  int tap_dh[9], tap_dw[9], tap_mi[9], tap_k[9];
  // output row mapping: up2 = 1 writes row gm = (n, a, b) of the Ho x Wo tile grid to pixel (2a + up_ph, 2b + up_pw) of a
  // [B][2Ho][2Wo][Cout] tensor (and reads the residual there): one output parity of a stride-2 data gradient
  int up2, up_ph, up_pw, up_lh, up_lw;  // up_lh / up_lw = log2(Ho), log2(Wo)
  int ds_cblocks;  // fused downsample: extra 64-channel k-blocks read through maps.ds after the R*S taps (0 = none)
  int a_stages;    // resident-weights pair kernel (UmmaConvLaunch::resb): activation ring stages (2..8)
  // k-block-tiled weights ("KT" layout, b_tiled = 1): the packed weights are stored as the shared-memory images the
  // kernels consume -- for every (hi/lo half, k-block) a [Cout][128 B] slab whose 16-byte chunks are already XOR-swizzled
  // (chunk ^ (row % 8), the SWIZZLE_128B pattern) -- so a CTA's weight tile of a k-block is ONE contiguous run that TMA
  // moves as a few 512-byte rows (no-swizzle map) instead of one 128-byte row per output channel: the TMA-im2col layers
  // were bound by the TMA engine's row rate (rows per k-block x ~2.3 cycles: 192 rows = 452 cycles on layer2 for 256
  // cycles of MMAs), and the weight rows were a third to a half of them.
  int b_tiled;
  int reverse;     // pair kernel (single conv): walk the work items from the last to the first (L2 reuse across launches)
  int kb_total;    // k-blocks per weight row half (= K / 64), KT layout
  int relu;
  const float* bias;            // [Cout] (folded BN shift)
  const void* residual;         // bf16 [M_total][Cout] or null
  void* out;                    // bf16 [M_total][Cout]
  double* stats;                // training forward: [2][Cout] += per-channel sum / sum of squares of the bf16 outputs, or null
};

struct UmmaConvLaunch {
  UmmaConvMaps maps;
  UmmaConvParams p;
  int block_n;  // 64 / 128 / 256
  int sw;       // swizzle bytes = bytes of K per smem row: 128 (generic) or 64 (stem)
  int grid;
  int split;    // 1: split-bf16 ("bf16x3") operands; 2: "fp16x2" (fp16 activations, weights [hi K | lo K] fp16)
  int pair;     // 1: CTA-pair kernel (conv_umma2.cu, tcgen05 cta_group::2), p.cl == 2
  int resb;     // 1: pair kernel with the whole weight matrix resident in shared memory (single N tile; layer2)
  int smem;     // dynamic shared memory of the resb launch
  double flops;  // algorithmic 2*MAC (unpadded)
};

// Generic conv (R x R, stride 1 or 2, Cin % 64 == 0) on NHWC bf16.  wpacked: bf16 [Cout][R*R*Cin]
// with K order (r, s, c).
int build_umma_conv(UmmaConvLaunch* L, const void* in, int B, int Hin, int Win, int Cin,
                    const void* wpacked, const float* bias, const void* residual, void* out, int Cout,
                    int R, int stride, int pad, int relu, int block_n, int cluster = 0 /* 0 = choose; 1/2/4 = multicast cluster; -2 = CTA pair */,
                    int split = 0 /* 1: activations [hi|lo] (2*Cin channels), weights [Cout][hi K | lo K], out/residual [hi|lo];
                                     2: fp16 activations / output, weights [Cout][hi K | lo K] fp16 (K incl. a fused projection) */,
                    // fused 1x1 stride-2 projection (resnet.py:240-243) of `ds_in` [B][ds_Hin][ds_Win][ds_Cin] accumulated into
                    // the same tile: the packed weight rows carry ds_Cin extra columns after R*R*Cin (K-extension)
                    const void* ds_in = nullptr, int ds_Hin = 0, int ds_Win = 0, int ds_Cin = 0,
                    double* stats = nullptr /* fused batch statistics of the output (non-split) */,
                    int b_tiled = 0 /* wpacked is in the KT layout (umma_pack_kt_offset); single-CTA and pair kernels only */);
// byte offset of weight element (half hl, output channel n, k) in the KT layout; Krow = K per half (multiple of 64)
inline size_t umma_pack_kt_offset(int hl, int n, int k, int Cout, int Krow) {
  const size_t kb = static_cast<size_t>(k) / 64, kk = static_cast<size_t>(k) % 64;
  return ((static_cast<size_t>(hl) * (Krow / 64) + kb) * Cout + n) * 128 + (((kk / 8) ^ (static_cast<size_t>(n) % 8)) * 16) + (kk % 8) * 2;
}
// One output parity (ph, pw) of the data gradient of a stride-2 R x R (R = 3 pad 1, or R = 1 pad 0) convolution:
// dx[2a+ph, 2b+pw] (+= residual there) = sum over the forward taps (r, s) of that parity of dz[a+di, b+dj] . w[r][s]^T.
// dz: bf16 [B][Ho][Wo][Cout_f]; w_dgrad: the data-gradient operand [Cin_f][(R-1-r, R-1-s), Cout_f] (pack mode 1);
// dx / residual: bf16 [B][2Ho][2Wo][Cin_f].  Returns PB200_OK with L->p.num_tiles = 0 when the parity has no tap.
int build_umma_dgrad_s2(UmmaConvLaunch* L, const void* dz, int B, int Ho, int Wo, int Cout_f, const void* w_dgrad,
                        const float* zero_bias, const void* residual_or_null, void* dx, int Cin_f, int R, int ph, int pw,
                        int block_n);
// BLOCK_N and {single CTA (1), CTA pair (-2)} for m_tiles x Cout outputs by the shared-memory-traffic cost model (passes = 2: fp16x2)
void umma_choose_tile(long long m_tiles, int Cout, int passes, int* block_n, int* cluster);
int launch_umma_conv(const UmmaConvLaunch& L, cudaStream_t st);  // dispatches on L.pair

// Several pair<256> convolutions in ONE persistent launch (conv_umma2.cu, CHAIN): the work list is the concatenation of
// the convs' tile lists; tiles of conv c wait for the images they read to be stored by conv dep[c] of the same launch.
constexpr int kMaxChain = 8;
struct ChainCtl {
  int nconv, total_items;
  int dep[kMaxChain];          // conv (index in the chain) whose output conv c reads as its input; -1: an earlier launch
  int dep_ratio[kMaxChain];    // M-supers of conv dep[c] per M-super of conv c (same images)
  int cnt_off[kMaxChain + 1];  // first counter of conv c: one per M-super (256 output pixels = whole images)
  int target[kMaxChain];       // arrivals that complete an M-super of conv c (16 epilogue warps x N tiles); 0: nobody waits
  int bias_off[kMaxChain];     // conv c's bias in the kernel's shared-memory bias array
  int ncnt;                    // counters in use; cnt[ncnt] = finished-CTA ticket
  unsigned* cnt;               // zero before the first launch; every launch leaves it zero
};
struct alignas(64) ChainArgs {
  UmmaConvMaps maps[kMaxChain];
  UmmaConvParams p[kMaxChain];
  ChainCtl ctl;
};
struct ChainLaunch {
  ChainArgs args;
  int grid = 0;  // 0: no chain
  int x2 = 0;
};
constexpr int kChainCounters = 4096;  // unsigned words a plan reserves for a chain (+ 1 ticket)
// Builds a chain from launches prepared by build_umma_conv (all pair<256>, Ho*Wo <= 256 so that an M-super holds whole
// images); dep[i] = index in `convs` of the conv producing conv i's input, or -1.  Returns PB200_OK with C->grid = 0 when
// the convs do not qualify.
int build_umma_chain(ChainLaunch* C, const UmmaConvLaunch* const* convs, const int* dep, int n, unsigned* counters);
int launch_umma_chain(const ChainLaunch& C, cudaStream_t st);
// conv_umma2.cu
int umma_pair_init();
int umma_pair_max_pairs(int block_n, int x2 = 0);
int launch_umma_pair(const UmmaConvLaunch& L, cudaStream_t st);
int umma_conv_init();  // sets the dynamic shared-memory attributes once per process

// conv_halo.cu: 3x3/1 conv with Cout = 128 whose KT-layout weights stay resident in a CTA pair's shared memory and whose
// activations arrive as ONE halo patch per tile and 64-channel block (layer2).  in / out / residual: NHWC bf16;
// wpacked_kt: KT layout (umma_pack_kt_offset) incl. the fused projection's columns; ds_in: [B][2H][2W][ds_Cin] or null.
struct HaloLaunch {
  UmmaConvMaps maps;
  alignas(8) unsigned char params[160];
  int grid, smem;
  double flops;
};
bool conv_halo_supported(int B, int H, int W, int Cin, int Cout, int R, int stride, int pad, int ds_Cin);
int build_conv_halo(HaloLaunch* L, const void* in, int B, int H, int W, int Cin, const void* wpacked_kt, const float* bias,
                    const void* residual, void* out, int Cout, int relu, const void* ds_in, int ds_Cin, int reverse = 0);
int launch_conv_halo(const HaloLaunch& L, cudaStream_t st);

}  // namespace pb200
