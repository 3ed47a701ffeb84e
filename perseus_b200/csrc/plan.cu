// The detector plan: weight folding/packing, buffer layout, and the forward launch sequence.
// Mirrors KeypointCNN (perseus/detector/models.py:6-40) = torchvision resnet18 with conv1 swapped
// for RGBD and fc -> Linear(512, 2*n_keypoints); forward = ResNet._forward_impl (resnet.py:266-282).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include <cuda_fp16.h>

#include "common.cuh"
#include "conv_umma.cuh"
#include "layout.h"

namespace pb200 {

struct ConvSpec {
  std::string name;     // "conv1", "layer1.0.conv1", "layer2.0.downsample.0", ...
  std::string bn;       // matching BN prefix
  int Cin, Cout, R, stride, pad;
  int Hin, Win, Ho, Wo;
  bool stem;
  bool p8 = false;      // bf16 modes: weights packed as the resident shared-memory image of conv_p8.cu
  int ds_conv = -1;     // bf16 mode, conv2 of a projection block: index of the 1x1/2 downsample conv fused into this one as
                        // extra K columns (its own op is not launched)
  // device storage (offsets into the weight arena)
  size_t w_off = 0;     // packed weights
  size_t b_off = 0;     // fp32 bias [Cout]
  size_t w_bytes = 0;
};

enum OpKind { OP_REPACK, OP_CONV, OP_MAXPOOL, OP_HEAD, OP_STEM_POOL, OP_CONV_P8 };

struct Op {
  OpKind kind;
  std::string name;
  int conv = -1;   // index into convs
  int in_buf = -1, out_buf = -1, res_buf = -1;  // activation buffer ids
  int relu = 0;
  int out_nhwc = 0;  // OP_CONV_P8: write NHWC (hand-over to the TMA-im2col layers) instead of P8
  int ds_buf = -1;   // OP_CONV with a fused downsample: the block input the projection reads
  int reverse = 0;   // OP_CONV_P8: walk the tiles back to front (alternates launch by launch: launch_conv_p8)
};

// activation buffer ids
// BUF_S0..: layers 3-4 of the 16-bit plans run as ONE chained launch (conv_umma2.cu, CHAIN): every conv output of those
// layers gets a buffer of its own, so that no store of the launch lands in a buffer an earlier conv of the same launch
// may still be reading for other images
enum { BUF_FRAMES = 0, BUF_STEM = 1, BUF_R0 = 2, BUF_R1 = 3, BUF_R2 = 4, BUF_R3 = 5, BUF_P8A = 6, BUF_P8B = 7,
       BUF_P8C = 8, BUF_S0 = 9, BUF_COUNT = 17 };

}  // namespace pb200

struct pb200_plan {
  int device = 0, max_batch = 0, C = 3, H = 256, W = 256, n_kp = 8, precision = PB200_PREC_BF16;
  bool bf16 = true;    // the fused 16-bit tcgen05 path (fused stem, P8 layer1, TMA-im2col layers 2-4): bf16 AND fp16x2 modes
  bool f16 = false;    // "fp16x2" mode of that path: fp16 activations, weights = fp16 hi + fp16 lo (two MMAs per K step)
  bool split = false;  // split-bf16 "bf16x3" mode: fp32 SIMT stem + maxpool, every later conv on tcgen05 with [hi|lo] operands
  size_t elt = 2;
  std::vector<pb200::ConvSpec> convs;
  std::vector<pb200::Op> ops;
  size_t buf_per_frame[pb200::BUF_COUNT] = {0};  // bytes per frame
  size_t buf_off[pb200::BUF_COUNT] = {0};        // offsets in the workspace (for max_batch)
  size_t workspace_bytes = 0;
  // weights
  uint8_t* arena = nullptr;
  size_t arena_bytes = 0;
  size_t fc_w_off = 0, fc_b_off = 0;
  bool loaded = false;
  // cached launch descriptors (TMA maps bake in pointers and the batch size)
  const void* cache_frames = nullptr;
  void* cache_ws = nullptr;
  int cache_B = -1;
  std::vector<pb200::UmmaConvLaunch> umma;  // per conv
  std::vector<pb200::HaloLaunch> halo;      // per conv (grid == 0: the conv does not run on conv_halo.cu)
  bool use_halo = true;                     // bf16 mode, KT weights: layer2's stride-1 convs on the halo-patch kernel
  // layers 3-4 (every run of consecutive pair<256> convs) as ONE persistent launch with in-kernel dependencies
  // (conv_umma2.cu CHAIN); PB200_NO_CHAIN=1: one launch per conv.  chain_at[i] = chain starting at op i (-1: none),
  // chain_len[i] = ops it covers
  // set for the duration of pb200_forward_rgb8: the fused stem reads this uint8 batch instead of the frames buffer
  const void* rgb8_src = nullptr;
  int rgb8_bgr = 0;
  bool use_chain = true;
  size_t chain_off = 0;
  std::vector<pb200::ChainLaunch> chains;
  std::vector<int> chain_at, chain_len;
  std::vector<std::string> op_names;
  std::vector<std::string> op_kernels;  // kernel each op launches (known after the first forward at a batch size)
  // L2 chunking of the front of the network (fused stem + the four layer1 convs, P8 layout): these five kernels move
  // 10 x 140 MB of activations per 256 frames; run on sub-batches of l1_chunk frames their three P8 buffers stay
  // L2-resident and only the frames and the NHWC hand-over to layer2 touch HBM.  MEASURED (r02, profiles/README.md): a
  // loss -- every extra launch of these persistent kernels costs ~9 us of pipeline fill / drain / tail imbalance, more
  // than the HBM time it saves (stem + layer1 0.43 ms unchunked, 0.55 / 0.58 / 0.72 ms at 86 / 64 / 40 frames per
  // chunk).  OFF by default; PB200_L1_CHUNK=<frames> enables it for A/B runs.
  int l1_chunk = 0;
  // weights of the TMA-im2col layers in the k-block-tiled "KT" layout (conv_umma.cuh): the TMA engine moves a CTA's
  // weight tile as a few 512-byte rows instead of one 128-byte row per output channel.  PB200_NO_BTILED=1: row-major.
  bool b_tiled = true;
  int n_front = 0;  // leading ops that are chunked (OP_STEM_POOL + OP_CONV_P8 x 4)
};

namespace pb200 {

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40);  // NaN
  const uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return static_cast<uint16_t>(u >> 16);
}

static uint16_t f32_to_f16_rne(float f) { return __half_as_ushort(__float2half_rn(f)); }
static float f16_bits_to_f32(uint16_t h) { return __half2float(__ushort_as_half(h)); }

// restores the caller's current device when a C-ABI entry has to switch to the plan's
struct DeviceGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != device) err = cudaSetDevice(device);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

static void add_conv(pb200_plan* pl, const std::string& name, const std::string& bn, int Cin, int Cout,
                     int R, int stride, int pad, int Hin, int Win, bool stem = false) {
  ConvSpec c;
  c.name = name;
  c.bn = bn;
  c.Cin = Cin; c.Cout = Cout; c.R = R; c.stride = stride; c.pad = pad;
  c.Hin = Hin; c.Win = Win;
  c.Ho = (Hin + 2 * pad - R) / stride + 1;
  c.Wo = (Win + 2 * pad - R) / stride + 1;
  c.stem = stem;
  pl->convs.push_back(c);
}

static void build_graph(pb200_plan* pl) {
  const int H = pl->H, W = pl->W;
  // stem (resnet.py:268-270; RGBD conv at models.py:27-28)
  add_conv(pl, "conv1", "bn1", pl->C, 64, 7, 2, 3, H, W, true);
  const int H1 = (H + 6 - 7) / 2 + 1, W1 = (W + 6 - 7) / 2 + 1;   // after conv1
  const int H2 = (H1 + 2 - 3) / 2 + 1, W2 = (W1 + 2 - 3) / 2 + 1; // after maxpool
  int cur = 0;  // rotating buffer index holding the block input
  int h = H2, w = W2, cin = 64;
  const int planes[4] = {64, 128, 256, 512};
  int first_layer = 0;
  if (pl->bf16) {
    // fused stem (conv1+bn1+relu+maxpool) and layer1 on the P8 layout (stem_pool.cu, conv_p8.cu)
    {
      Op op;
      op.kind = OP_STEM_POOL; op.name = "conv1+bn1+relu+maxpool"; op.conv = 0; op.in_buf = BUF_FRAMES; op.out_buf = BUF_P8A;
      // the ingest kernel writes the frames front to back: the stem starts at their end (what L2 still holds), and from
      // here on every launch of layers 1-2 walks the other way than the one before it (PB200_P8_NOREVERSE=1: all forward)
      op.reverse = getenv("PB200_P8_NOREVERSE") == nullptr ? 1 : 0;
      pl->ops.push_back(op);
    }
    const int seq_in[4] = {BUF_P8A, BUF_P8B, BUF_P8C, BUF_P8B};
    const int seq_out[4] = {BUF_P8B, BUF_P8C, BUF_P8B, BUF_R0};
    const int seq_res[4] = {-1, BUF_P8A, -1, BUF_P8C};
    for (int i = 0; i < 4; ++i) {
      const std::string pre = "layer1." + std::to_string(i / 2);
      const std::string cn = (i % 2) ? ".conv2" : ".conv1", bn = (i % 2) ? ".bn2" : ".bn1";
      add_conv(pl, pre + cn, pre + bn, 64, 64, 3, 1, 1, H2, W2);
      pl->convs.back().p8 = true;
      Op op;
      op.kind = OP_CONV_P8; op.name = pre + cn + "+bn" + ((i % 2) ? "+add+relu" : "+relu");
      op.conv = static_cast<int>(pl->convs.size()) - 1;
      op.in_buf = seq_in[i]; op.out_buf = seq_out[i]; op.res_buf = seq_res[i]; op.relu = 1; op.out_nhwc = (i == 3);
      // the stem walks backward, so conv 0 starts at the front, conv 1 at the end, ... (PB200_P8_NOREVERSE=1: all forward)
      op.reverse = (i % 2 == 1 && getenv("PB200_P8_NOREVERSE") == nullptr) ? 1 : 0;
      pl->ops.push_back(op);
    }
    first_layer = 1;
  } else {
    {
      Op op;
      op.kind = OP_CONV; op.name = "conv1+bn1+relu"; op.conv = 0; op.in_buf = BUF_FRAMES; op.out_buf = BUF_STEM; op.relu = 1;
      pl->ops.push_back(op);
    }
    {
      Op op;
      op.kind = OP_MAXPOOL; op.name = "maxpool"; op.in_buf = BUF_STEM; op.out_buf = BUF_R0;
      pl->ops.push_back(op);
    }
  }
  int next_fresh = BUF_S0;
  const bool fuse_ds = pl->bf16 && !getenv("PB200_NO_DS_FUSION");
  auto ring = [](int buf, int k) { return BUF_R0 + (buf - BUF_R0 + k) % 4; };
  cur = BUF_R0 + cur;  // from here on `cur` is a buffer id
  for (int L = first_layer; L < 4; ++L) {
    for (int b = 0; b < 2; ++b) {
      const int cout = planes[L];
      const int stride = (L > 0 && b == 0) ? 2 : 1;
      const std::string pre = "layer" + std::to_string(L + 1) + "." + std::to_string(b);
      // layers 3-4 of the 16-bit plans: fresh buffers (see BUF_S0); otherwise the four-buffer ring
      const bool fresh = fuse_ds && L >= 2;
      const int t = fresh ? next_fresh++ : ring(cur, 1);
      const int y = fresh ? next_fresh++ : ring(cur, 2);
      const int ds = fresh ? -1 : ring(cur, 3);
      int res = cur;
      int fused_ds = -1;  // bf16 mode: the projection is accumulated inside conv2 (K-extension), never materialised
      if (stride != 1 || cin != cout) {  // resnet.py:240-243
        add_conv(pl, pre + ".downsample.0", pre + ".downsample.1", cin, cout, 1, stride, 0, h, w);
        if (fuse_ds && stride == 2) {
          fused_ds = static_cast<int>(pl->convs.size()) - 1;
          res = -1;
        } else {
          Op op;
          op.kind = OP_CONV; op.name = pre + ".downsample"; op.conv = static_cast<int>(pl->convs.size()) - 1;
          op.in_buf = cur; op.out_buf = ds; op.relu = 0;
          pl->ops.push_back(op);
          res = ds;
        }
      }
      add_conv(pl, pre + ".conv1", pre + ".bn1", cin, cout, 3, stride, 1, h, w);
      {
        Op op;
        op.kind = OP_CONV; op.name = pre + ".conv1+bn1+relu"; op.conv = static_cast<int>(pl->convs.size()) - 1;
        op.in_buf = cur; op.out_buf = t; op.relu = 1;
        pl->ops.push_back(op);
      }
      const int ho = pl->convs.back().Ho, wo = pl->convs.back().Wo;
      add_conv(pl, pre + ".conv2", pre + ".bn2", cout, cout, 3, 1, 1, ho, wo);
      {
        Op op;
        op.kind = OP_CONV; op.name = pre + (fused_ds >= 0 ? ".conv2+bn2+downsample+add+relu" : ".conv2+bn2+add+relu");
        op.conv = static_cast<int>(pl->convs.size()) - 1;
        op.in_buf = t; op.out_buf = y; op.res_buf = res; op.relu = 1;
        if (fused_ds >= 0) {
          pl->convs.back().ds_conv = fused_ds;
          op.ds_buf = cur;
        }
        pl->ops.push_back(op);
      }
      if (fresh) {
        pl->buf_per_frame[t] = pl->buf_per_frame[y] = static_cast<size_t>(ho) * wo * cout * pl->elt;
      }
      cur = y;
      h = ho; w = wo; cin = cout;
    }
  }
  {
    Op op;
    op.kind = OP_HEAD; op.name = "avgpool+fc+decode"; op.in_buf = cur;
    pl->ops.push_back(op);
  }
  // buffers (bytes per frame)
  const size_t e = pl->elt;
  pl->buf_per_frame[BUF_FRAMES] = static_cast<size_t>(H + 6) * (W + 8) * 4 * e;
  if (pl->bf16) {
    pl->buf_per_frame[BUF_STEM] = 0;  // the stem output never leaves the SM (fused maxpool)
    pl->buf_per_frame[BUF_R0] = static_cast<size_t>(H2) * W2 * 64 * e;
    for (int i = 1; i < 4; ++i) pl->buf_per_frame[BUF_R0 + i] = static_cast<size_t>(H2 / 2) * (W2 / 2) * 128 * e;
    for (int i = 0; i < 3; ++i) pl->buf_per_frame[BUF_P8A + i] = p8_image_bytes(64, H2, W2);
  } else {
    pl->buf_per_frame[BUF_STEM] = static_cast<size_t>(H1) * W1 * 64 * e;
    for (int i = 0; i < 4; ++i) pl->buf_per_frame[BUF_R0 + i] = static_cast<size_t>(H2) * W2 * 64 * e;
  }
  size_t off = 0;
  for (int i = 0; i < BUF_COUNT; ++i) {
    pl->buf_off[i] = off;
    if (pl->buf_per_frame[i] == 0) continue;
    off += align_up(pl->buf_per_frame[i] * pl->max_batch + (i >= BUF_P8A ? kP8SlackBytes : 0), 1024);
  }
  pl->workspace_bytes = off;
  pl->n_front = 0;
  if (pl->bf16)
    while (pl->n_front < static_cast<int>(pl->ops.size()) &&
           (pl->ops[pl->n_front].kind == OP_STEM_POOL || pl->ops[pl->n_front].kind == OP_CONV_P8))
      ++pl->n_front;
  if (const char* e = getenv("PB200_L1_CHUNK")) pl->l1_chunk = atoi(e);
  pl->b_tiled = (pl->bf16 || pl->f16) && getenv("PB200_NO_BTILED") == nullptr;
  pl->use_halo = pl->b_tiled && !pl->f16 && getenv("PB200_NO_HALO") == nullptr;
  pl->use_chain = pl->bf16 && getenv("PB200_NO_CHAIN") == nullptr;
  // weight arena
  size_t woff = 0;
  for (auto& c : pl->convs) {
    const size_t K = (c.stem ? kStemK : static_cast<size_t>(c.R) * c.R * c.Cin) +
                     (c.ds_conv >= 0 ? static_cast<size_t>(pl->convs[c.ds_conv].Cin) : 0);
    c.w_bytes = K * c.Cout * e * (pl->f16 ? 2 : 1);  // fp16x2: hi image / rows followed by the lo ones
    c.w_off = woff;
    woff = align_up(woff + c.w_bytes, 1024);
    c.b_off = woff;
    woff = align_up(woff + sizeof(float) * c.Cout, 1024);
  }
  pl->fc_w_off = woff;
  woff = align_up(woff + sizeof(float) * 512 * 2 * pl->n_kp, 1024);
  pl->fc_b_off = woff;
  woff = align_up(woff + sizeof(float) * 2 * pl->n_kp, 1024);
  // dependency counters of the chained launch (zero between launches; the weight upload rewrites them with zeros)
  pl->chain_off = align_up(woff, 1024);
  woff = pl->chain_off + (kChainCounters + 1) * sizeof(unsigned);
  pl->arena_bytes = woff;
  for (auto& o : pl->ops) pl->op_names.push_back(o.name);
}

}  // namespace pb200

// ------------------------------------------------------------------------------------------
// C ABI: plan lifecycle and weights
// ------------------------------------------------------------------------------------------
using namespace pb200;

extern "C" int pb200_plan_create(pb200_plan** out, int device, int max_batch, int in_channels, int H,
                                 int W, int n_keypoints, int precision) {
  PB_CHECK(out != nullptr, "plan_create: null out");
  *out = nullptr;
  PB_CHECK(max_batch >= 1, "plan_create: max_batch %d", max_batch);
  PB_CHECK(in_channels >= 1 && in_channels <= 4,
           "plan_create: in_channels %d (the frames layout carries up to 4: R,G,B,D)", in_channels);
  PB_CHECK(n_keypoints >= 1, "plan_create: n_keypoints %d", n_keypoints);
  PB_CHECK(precision == PB200_PREC_FP32 || precision == PB200_PREC_BF16 || precision == PB200_PREC_BF16X3 ||
               precision == PB200_PREC_FP16X2,
           "plan_create: precision %d not available (0 = fp32 SIMT, 1 = bf16 tcgen05, 2 = split-bf16 tcgen05, "
           "3 = fp16 x (hi + lo) tcgen05)", precision);
  PB_CHECK(H % 32 == 0 && W % 32 == 0 && H >= 32 && W >= 32, "plan_create: H, W must be multiples of 32");
  if (precision != PB200_PREC_FP32)
    PB_CHECK(H == 256 && W == 256, "plan_create: the tcgen05 path is tiled for 256x256 frames (got %dx%d); "
             "use PB200_PREC_FP32 for other sizes", H, W);
  DeviceGuard guard(device);  // the caller's current device is restored on return
  PB_CUDA(guard.err);
  if (int rc = require_device()) return rc;
  pb200_plan* pl = new pb200_plan();
  pl->device = device;
  pl->max_batch = max_batch;
  pl->C = in_channels;
  pl->H = H;
  pl->W = W;
  pl->n_kp = n_keypoints;
  pl->precision = precision;
  pl->bf16 = precision == PB200_PREC_BF16 || precision == PB200_PREC_FP16X2;
  pl->f16 = precision == PB200_PREC_FP16X2;
  pl->split = precision == PB200_PREC_BF16X3;
  // split mode: fp32 frames / stem output, later activations are [hi | lo] bf16 pairs = 4 bytes per value as well
  pl->elt = pl->bf16 ? 2 : 4;
  build_graph(pl);
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&pl->arena), pl->arena_bytes);
  if (e == cudaSuccess) e = cudaMemset(pl->arena + pl->chain_off, 0, (kChainCounters + 1) * sizeof(unsigned));
  if (e != cudaSuccess) {
    delete pl;
    return cuda_fail(e, "cudaMalloc(weight arena)", __FILE__, __LINE__);
  }
  if (pl->bf16 || pl->split) {
    int rc = umma_conv_init();
    if (rc == PB200_OK && pl->bf16) rc = stem_pool_init();
    if (rc == PB200_OK && pl->bf16) rc = conv_p8_init();
    if (rc) {
      cudaFree(pl->arena);
      delete pl;
      return rc;
    }
  }
  *out = pl;
  return PB200_OK;
}

extern "C" void pb200_plan_destroy(pb200_plan* plan) {
  if (!plan) return;
  if (plan->arena) {
    DeviceGuard guard(plan->device);
    cudaFree(plan->arena);
  }
  delete plan;
}

extern "C" int pb200_plan_workspace_bytes(const pb200_plan* plan, size_t* bytes) {
  PB_CHECK(plan && bytes, "plan_workspace_bytes: null argument");
  *bytes = plan->workspace_bytes;
  return PB200_OK;
}

extern "C" int pb200_plan_frames_bytes(const pb200_plan* plan, int B, size_t* bytes) {
  PB_CHECK(plan && bytes && B >= 0, "plan_frames_bytes: bad argument");
  *bytes = plan->buf_per_frame[BUF_FRAMES] * static_cast<size_t>(B);
  return PB200_OK;
}

namespace pb200 {
struct HostTensor {
  const float* data;
  std::vector<int64_t> shape;
  int64_t numel() const {
    int64_t n = 1;
    for (auto s : shape) n *= s;
    return n;
  }
};
}  // namespace pb200

extern "C" int pb200_plan_load_weights(pb200_plan* plan, const float* const* tensors,
                                       const int64_t* const* shapes, const int* ndims,
                                       const char* const* names, int n, pb200_stream stream) {
  PB_CHECK(plan && tensors && shapes && ndims && names, "load_weights: null argument");
  std::map<std::string, HostTensor> sd;
  for (int i = 0; i < n; ++i) {
    HostTensor t;
    t.data = tensors[i];
    t.shape.assign(shapes[i], shapes[i] + ndims[i]);
    std::string nm(names[i]);
    for (const std::string mod : {"module.", "_orig_mod."}) {  // validate.py:94-96 (DDP); torch.compile saves (train.py:194-196)
      size_t pos;
      while ((pos = nm.find(mod)) != std::string::npos) nm.erase(pos, mod.size());
    }
    sd[nm] = t;
  }
  auto need = [&](const std::string& key, std::vector<int64_t> shape, const HostTensor** out) -> int {
    auto it = sd.find(key);
    PB_CHECK(it != sd.end(), "load_weights: missing tensor '%s'", key.c_str());
    PB_CHECK(it->second.shape == shape, "load_weights: tensor '%s' has the wrong shape", key.c_str());
    PB_CHECK(it->second.data != nullptr, "load_weights: tensor '%s' is null", key.c_str());
    *out = &it->second;
    return PB200_OK;
  };
  std::vector<uint8_t> host(plan->arena_bytes, 0);
  for (const auto& c : plan->convs) {
    const HostTensor *w, *g, *b, *mu, *var;
    const std::string wn = "resnet." + c.name + ".weight", bp = "resnet." + c.bn;
    if (int rc = need(wn, {c.Cout, c.Cin, c.R, c.R}, &w)) return rc;
    if (int rc = need(bp + ".weight", {c.Cout}, &g)) return rc;
    if (int rc = need(bp + ".bias", {c.Cout}, &b)) return rc;
    if (int rc = need(bp + ".running_mean", {c.Cout}, &mu)) return rc;
    if (int rc = need(bp + ".running_var", {c.Cout}, &var)) return rc;
    float* bias = reinterpret_cast<float*>(host.data() + c.b_off);
    const int K = c.stem ? kStemK : c.R * c.R * c.Cin;
    const int Krow = K + (c.ds_conv >= 0 ? plan->convs[c.ds_conv].Cin : 0);  // K-major row length incl. a fused projection
    for (int co = 0; co < c.Cout; ++co) {
      // eval-mode BatchNorm2d, eps = 1e-5: y = (x - mean) / sqrt(var + eps) * gamma + beta
      const double sc = static_cast<double>(g->data[co]) / sqrt(static_cast<double>(var->data[co]) + 1e-5);
      bias[co] = static_cast<float>(static_cast<double>(b->data[co]) - static_cast<double>(mu->data[co]) * sc);
      for (int r = 0; r < c.R; ++r)
        for (int s = 0; s < c.R; ++s)
          for (int ci = 0; ci < c.Cin; ++ci) {
            double v = static_cast<double>(w->data[((static_cast<size_t>(co) * c.Cin + ci) * c.R + r) * c.R + s]) * sc;
            int k;
            if (c.stem) {
              k = r * (kStemTaps * kFrameC) + (s + 1) * kFrameC + ci;
              if (plan->f16) v *= ci < 3 ? (1.0 / 255.0) / kF16RgbScale : 1.0 / kF16DepthScale;  // frames: k * 2^-12, d * 2^-4
              else if (plan->bf16 && ci < 3) v /= 255.0;  // frames carry RGB as integers 0..255
            } else {
              k = (r * c.R + s) * c.Cin + ci;
            }
            if (plan->split && !c.stem) {
              // [Cout][hi K | lo K] bf16 rows: w = hi + lo, hi = bf16(w), lo = bf16(w - hi)
              const float wf = static_cast<float>(v);
              const uint16_t hi = f32_to_bf16_rne(wf);
              uint32_t hbits = static_cast<uint32_t>(hi) << 16;
              float hf;
              memcpy(&hf, &hbits, 4);
              uint16_t* row = reinterpret_cast<uint16_t*>(host.data() + c.w_off) + static_cast<size_t>(co) * 2 * K;
              row[k] = hi;
              row[K + k] = f32_to_bf16_rne(wf - hf);
            } else if (plan->f16) {
              // w = hi + lo, both fp16: hi image / K-row half first, the lo one right behind it
              const float wf = static_cast<float>(v);
              const uint16_t hi = f32_to_f16_rne(wf);
              const uint16_t lo = f32_to_f16_rne(wf - f16_bits_to_f32(hi));
              size_t off = (static_cast<size_t>(co) * 2 * Krow + k) * 2, lo_off = static_cast<size_t>(Krow) * 2;
              if (c.stem) { off = stem_wimg_offset(co, k); lo_off = kStemWimgBytes; }
              else if (c.p8) { off = conv_p8_wimg_offset(co, k); lo_off = 64 * 576 * 2; }
              else if (plan->b_tiled) { off = umma_pack_kt_offset(0, co, k, c.Cout, Krow); lo_off = umma_pack_kt_offset(1, co, k, c.Cout, Krow) - off; }
              *reinterpret_cast<uint16_t*>(host.data() + c.w_off + off) = hi;
              *reinterpret_cast<uint16_t*>(host.data() + c.w_off + off + lo_off) = lo;
            } else if (plan->bf16) {
              size_t off = (static_cast<size_t>(co) * Krow + k) * 2;  // [Cout][K] K-major rows (TMA-im2col kernels)
              if (c.stem) off = stem_wimg_offset(co, k);
              else if (c.p8) off = conv_p8_wimg_offset(co, k);
              else if (plan->b_tiled) off = umma_pack_kt_offset(0, co, k, c.Cout, Krow);
              *reinterpret_cast<uint16_t*>(host.data() + c.w_off + off) = f32_to_bf16_rne(static_cast<float>(v));
            } else {
              reinterpret_cast<float*>(host.data() + c.w_off)[static_cast<size_t>(k) * c.Cout + co] =
                  static_cast<float>(v);
            }
          }
    }
  }
  // fused projections: the downsample's folded weights become columns [R*R*Cin, +Cin_ds) of conv2's rows, its folded
  // shift is added to conv2's (out = conv2(y1) + b2 + ds(x) + b_ds, resnet.py:99-103)
  for (const auto& c : plan->convs) {
    if (c.ds_conv < 0) continue;
    const ConvSpec& d = plan->convs[c.ds_conv];
    const HostTensor *w, *g, *b, *mu, *var;
    const std::string bp = "resnet." + d.bn;
    if (int rc = need("resnet." + d.name + ".weight", {d.Cout, d.Cin, 1, 1}, &w)) return rc;
    if (int rc = need(bp + ".weight", {d.Cout}, &g)) return rc;
    if (int rc = need(bp + ".bias", {d.Cout}, &b)) return rc;
    if (int rc = need(bp + ".running_mean", {d.Cout}, &mu)) return rc;
    if (int rc = need(bp + ".running_var", {d.Cout}, &var)) return rc;
    const int K = c.R * c.R * c.Cin, Krow = K + d.Cin;
    float* bias = reinterpret_cast<float*>(host.data() + c.b_off);
    for (int co = 0; co < c.Cout; ++co) {
      const double sc = static_cast<double>(g->data[co]) / sqrt(static_cast<double>(var->data[co]) + 1e-5);
      bias[co] = static_cast<float>(static_cast<double>(bias[co]) + static_cast<double>(b->data[co]) -
                                    static_cast<double>(mu->data[co]) * sc);
      uint16_t* row = reinterpret_cast<uint16_t*>(host.data() + c.w_off) + static_cast<size_t>(co) * Krow * (plan->f16 ? 2 : 1) + K;
      auto at = [&](int hl, int ci) -> uint16_t& {  // element (half, co, K + ci) of the packed weights
        if (plan->b_tiled) return *reinterpret_cast<uint16_t*>(host.data() + c.w_off + umma_pack_kt_offset(hl, co, K + ci, c.Cout, Krow));
        return row[hl * Krow + ci];
      };
      for (int ci = 0; ci < d.Cin; ++ci) {
        const float wf = static_cast<float>(static_cast<double>(w->data[static_cast<size_t>(co) * d.Cin + ci]) * sc);
        if (plan->f16) {
          const uint16_t hi = f32_to_f16_rne(wf);
          at(0, ci) = hi;
          at(1, ci) = f32_to_f16_rne(wf - f16_bits_to_f32(hi));
        } else {
          at(0, ci) = f32_to_bf16_rne(wf);
        }
      }
    }
  }
  const HostTensor *fw, *fb;
  const int n_out = 2 * plan->n_kp;
  if (int rc = need("resnet.fc.weight", {n_out, 512}, &fw)) return rc;
  if (int rc = need("resnet.fc.bias", {n_out}, &fb)) return rc;
  memcpy(host.data() + plan->fc_w_off, fw->data, sizeof(float) * 512 * n_out);
  memcpy(host.data() + plan->fc_b_off, fb->data, sizeof(float) * n_out);
  DeviceGuard guard(plan->device);
  PB_CUDA(guard.err);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PB_CUDA(cudaMemcpyAsync(plan->arena, host.data(), plan->arena_bytes, cudaMemcpyHostToDevice, st));
  PB_CUDA(cudaStreamSynchronize(st));
  plan->loaded = true;
  return PB200_OK;
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
namespace pb200 {

static inline int frame_elt(const pb200_plan* pl) { return pl->f16 ? FRAME_F16 : (pl->bf16 ? FRAME_BF16 : FRAME_F32); }

// n0 = first frame of an L2 chunk (0 when the op runs on the whole batch): offsets the frames and the NHWC buffers; the
// P8 buffers of a chunked run are chunk-local (always at their base, so they are re-used chunk after chunk in L2)
static inline uint8_t* buf_ptr(const pb200_plan* pl, const void* frames, void* ws, int id, int n0 = 0) {
  if (id == BUF_FRAMES)
    return const_cast<uint8_t*>(static_cast<const uint8_t*>(frames)) + static_cast<size_t>(n0) * pl->buf_per_frame[BUF_FRAMES];
  return static_cast<uint8_t*>(ws) + pl->buf_off[id] + (id >= BUF_P8A ? 0 : static_cast<size_t>(n0) * pl->buf_per_frame[id]);
}

static void choose_tile(const ConvSpec& c, int B, bool f16, int* block_n, int* cluster) {
  const long long m_tiles = (static_cast<long long>(B) * c.Ho * c.Wo + 127) / 128;
  umma_choose_tile(m_tiles, c.Cout, f16 ? 2 : 1, block_n, cluster);  // conv_umma.cu: the shared-memory-traffic cost model
  if (const char* e = getenv("PB200_R01_TILES")) {  // A/B: the round-1 choice (rounds x MMA cycles only, pairs from 2 waves on)
    if (e[0] == '1' && !f16) {
      double bc = -1;
      for (int bn = 64; bn <= 256 && bn <= c.Cout; bn *= 2) {
        if (c.Cout % bn) continue;
        const long long tiles = m_tiles * (c.Cout / bn);
        const double cost = static_cast<double>((tiles + num_sms() - 1) / num_sms()) * (bn / 2 < 48 ? 48 : bn / 2);
        if (bc < 0 || cost <= bc) { bc = cost; *block_n = bn; }
      }
      *cluster = 0;
    }
  }
}

// (re)build the TMA descriptors / launch parameters when a pointer or the batch size changed
static int prepare(pb200_plan* pl, const void* frames, void* ws, int B) {
  if (!pl->bf16 && !pl->split) return PB200_OK;
  // the tensor maps cover workspace buffers only (the stem reads `frames` through a plain pointer argument), so a caller
  // may alternate frames buffers (ingest of batch i+1 under the forward of batch i) without re-encoding anything
  if (pl->cache_ws == ws && pl->cache_B == B) return PB200_OK;
  pl->umma.assign(pl->convs.size(), UmmaConvLaunch());
  pl->halo.assign(pl->convs.size(), HaloLaunch());
  for (const auto& op : pl->ops) {
    if (op.kind != OP_CONV) continue;
    const ConvSpec& c = pl->convs[op.conv];
    if (c.stem) continue;  // bf16 stem = OP_STEM_POOL (no tensor maps); split stem = fp32 SIMT
    UmmaConvLaunch* L = &pl->umma[op.conv];
    const void* in = buf_ptr(pl, frames, ws, op.in_buf);
    void* out = buf_ptr(pl, frames, ws, op.out_buf);
    const void* res = op.res_buf >= 0 ? buf_ptr(pl, frames, ws, op.res_buf) : nullptr;
    const float* bias = reinterpret_cast<const float*>(pl->arena + c.b_off);
    const void* ds_in = op.ds_buf >= 0 ? buf_ptr(pl, frames, ws, op.ds_buf) : nullptr;
    const ConvSpec* d = c.ds_conv >= 0 ? &pl->convs[c.ds_conv] : nullptr;
    if (pl->use_halo && conv_halo_supported(B, c.Hin, c.Win, c.Cin, c.Cout, c.R, c.stride, c.pad, d ? d->Cin : 0)) {
      // alternate the walk direction launch by launch (launch_conv_p8): the halo conv that follows a forward-walking halo
      // conv starts at the end of the tensor
      const size_t idx = static_cast<size_t>(&op - pl->ops.data());
      const bool prev_fwd = idx > 0 && (pl->ops[idx - 1].kind == OP_CONV || pl->ops[idx - 1].kind == OP_CONV_P8) &&
                            !pl->ops[idx - 1].reverse;
      const int rev = prev_fwd && getenv("PB200_P8_NOREVERSE") == nullptr ? 1 : 0;
      const_cast<Op&>(op).reverse = rev;
      int rc = build_conv_halo(&pl->halo[op.conv], in, B, c.Hin, c.Win, c.Cin, pl->arena + c.w_off, bias, res, out, c.Cout,
                               op.relu, d ? ds_in : nullptr, d ? d->Cin : 0, rev);
      if (rc) {
        pl->cache_B = -1;
        return rc;
      }
      continue;
    }
    int bn = 64, cluster = 0;
    choose_tile(c, B, pl->f16, &bn, &cluster);
    if (pl->split) cluster = 1;  // the split-bf16 mode runs on the single-CTA kernel
    int rc = build_umma_conv(L, in, B, c.Hin, c.Win, c.Cin, pl->arena + c.w_off, bias, res, out, c.Cout,
                             c.R, c.stride, c.pad, op.relu, bn, cluster, pl->split ? 1 : (pl->f16 ? 2 : 0),
                             d != nullptr ? ds_in : nullptr, d ? d->Hin : 0, d ? d->Win : 0, d ? d->Cin : 0, nullptr,
                             (pl->b_tiled && !pl->split) ? 1 : 0);
    if (rc) {
      pl->cache_B = -1;
      return rc;
    }
    // layer2's first conv reads the 134 MB hand-over tensor that layer1's last conv has just written front to back: a
    // backward walk starts in the part L2 still holds (csrc/conv_p8.cu: launch_conv_p8).  PB200_P8_NOREVERSE=1: forward.
    if (L->pair && !L->resb && op.in_buf == BUF_R0 && getenv("PB200_P8_NOREVERSE") == nullptr) {
      const size_t idx = static_cast<size_t>(&op - pl->ops.data());
      if (idx > 0 && pl->ops[idx - 1].kind == OP_CONV_P8 && pl->ops[idx - 1].reverse == 0) L->p.reverse = 1;
      const_cast<Op&>(op).reverse = L->p.reverse;  // (the halo conv behind it alternates against this)
    }
  }
  // chains: maximal runs of consecutive convs on the pair<256> kernel whose outputs are whole images per M-super
  pl->chains.clear();
  pl->chain_at.assign(pl->ops.size(), -1);
  pl->chain_len.assign(pl->ops.size(), 0);
  if (pl->use_chain && pl->chains.empty()) {
    auto chainable = [&](size_t i) {
      const Op& op = pl->ops[i];
      if (op.kind != OP_CONV || pl->convs[op.conv].stem || pl->halo[op.conv].grid > 0) return false;
      const UmmaConvLaunch& L = pl->umma[op.conv];
      return L.pair && !L.resb && L.block_n == 256 && L.p.num_tiles > 0 && 256 % (L.p.Ho * L.p.Wo) == 0;
    };
    size_t i = 0;
    while (i < pl->ops.size()) {
      if (!chainable(i)) {
        ++i;
        continue;
      }
      size_t j = i;
      while (j < pl->ops.size() && chainable(j) && j - i < static_cast<size_t>(kMaxChain)) ++j;
      // every output buffer of the run must be new to the run (no store may land where an earlier conv still reads)
      bool safe = true;
      for (size_t k = i; k < j && safe; ++k)
        for (size_t m = i; m <= k && safe; ++m) {
          const Op& a = pl->ops[m];
          const int o = pl->ops[k].out_buf;
          if (a.in_buf == o || a.res_buf == o || a.ds_buf == o || (m < k && a.out_buf == o)) safe = false;
        }
      if (safe && j - i >= 2 && pl->chains.empty()) {  // one chain per plan (one set of counters)
        const UmmaConvLaunch* convs[kMaxChain];
        int dep[kMaxChain];
        for (size_t k = i; k < j; ++k) {
          convs[k - i] = &pl->umma[pl->ops[k].conv];
          dep[k - i] = -1;
          for (size_t m = i; m < k; ++m)
            if (pl->ops[m].out_buf == pl->ops[k].in_buf) dep[k - i] = static_cast<int>(m - i);
        }
        ChainLaunch C;
        int rc = build_umma_chain(&C, convs, dep, static_cast<int>(j - i), reinterpret_cast<unsigned*>(pl->arena + pl->chain_off));
        if (rc) {
          pl->cache_B = -1;
          return rc;
        }
        if (C.grid > 0) {
          pl->chain_at[i] = static_cast<int>(pl->chains.size());
          pl->chain_len[i] = static_cast<int>(j - i);
          pl->chains.push_back(C);
        }
      }
      i = j;
    }
  }
  pl->cache_frames = frames;
  pl->cache_ws = ws;
  pl->cache_B = B;
  return PB200_OK;
}

// the chained launch whose head is op i (chain_at[i] >= 0)
static int launch_chain_at(pb200_plan* pl, size_t i, const void* frames, void* ws, int B, cudaStream_t st) {
  (void)frames; (void)ws; (void)B;
  return launch_umma_chain(pl->chains[pl->chain_at[i]], st);
}

static int run_op(pb200_plan* pl, const Op& op, const void* frames, void* ws, float* out_norm,
                  float* out_px, int B, cudaStream_t st, int n0 = 0) {
  switch (op.kind) {
    case OP_CONV: {
      const ConvSpec& c = pl->convs[op.conv];
      if (pl->bf16 && !c.stem && pl->halo[op.conv].grid > 0) return launch_conv_halo(pl->halo[op.conv], st);
      if (pl->bf16 || (pl->split && !c.stem)) return launch_umma_conv(pl->umma[op.conv], st);
      SimtConv s;
      s.in = reinterpret_cast<const float*>(buf_ptr(pl, frames, ws, op.in_buf));
      s.w = reinterpret_cast<const float*>(pl->arena + c.w_off);
      s.bias = reinterpret_cast<const float*>(pl->arena + c.b_off);
      s.residual = op.res_buf >= 0 ? reinterpret_cast<const float*>(buf_ptr(pl, frames, ws, op.res_buf)) : nullptr;
      s.out = reinterpret_cast<float*>(buf_ptr(pl, frames, ws, op.out_buf));
      s.B = B; s.Ho = c.Ho; s.Wo = c.Wo; s.Cout = c.Cout; s.relu = op.relu;
      if (c.stem) {  // 7 x 8 taps over the zero-margined frames buffer (layout.h)
        s.Hin = pl->H + 6; s.Win = pl->W + 8; s.Cin = kFrameC; s.R = 7; s.S = kStemTaps; s.stride = 2; s.pad = 0;
      } else {
        s.Hin = c.Hin; s.Win = c.Win; s.Cin = c.Cin; s.R = c.R; s.S = c.R; s.stride = c.stride; s.pad = c.pad;
      }
      if (c.stem && stem_f32_supported(s)) return launch_stem_f32(s, pl->C, st);
      return launch_conv_simt(s, st);
    }
    case OP_STEM_POOL: {
      const ConvSpec& c = pl->convs[op.conv];
      return launch_stem_pool(buf_ptr(pl, frames, ws, BUF_FRAMES, n0), pl->arena + c.w_off,
                              reinterpret_cast<const float*>(pl->arena + c.b_off),
                              buf_ptr(pl, frames, ws, op.out_buf, n0), B, pl->H, pl->W, st, pl->f16 ? 1 : 0,
                              pl->rgb8_src == nullptr ? nullptr
                                                      : static_cast<const uint8_t*>(pl->rgb8_src) +
                                                            static_cast<size_t>(n0) * pl->H * pl->W * 3,
                              pl->rgb8_bgr, op.reverse);
    }
    case OP_CONV_P8: {
      const ConvSpec& c = pl->convs[op.conv];
      return launch_conv_p8(buf_ptr(pl, frames, ws, op.in_buf, n0), pl->arena + c.w_off,
                            reinterpret_cast<const float*>(pl->arena + c.b_off),
                            op.res_buf >= 0 ? buf_ptr(pl, frames, ws, op.res_buf, n0) : nullptr,
                            buf_ptr(pl, frames, ws, op.out_buf, n0), op.out_nhwc, op.relu, B, c.Hin, c.Win, st, pl->f16 ? 1 : 0,
                            op.reverse);
    }
    case OP_MAXPOOL: {
      const ConvSpec& c = pl->convs[0];
      if (pl->split)
        return launch_maxpool_split(buf_ptr(pl, frames, ws, op.in_buf), buf_ptr(pl, frames, ws, op.out_buf), B,
                                    c.Ho, c.Wo, c.Cout, st);
      return launch_maxpool(buf_ptr(pl, frames, ws, op.in_buf), buf_ptr(pl, frames, ws, op.out_buf), B,
                            c.Ho, c.Wo, c.Cout, pl->bf16, st);
    }
    case OP_HEAD: {
      const ConvSpec& c = pl->convs.back();
      return launch_head(buf_ptr(pl, frames, ws, op.in_buf),
                         reinterpret_cast<const float*>(pl->arena + pl->fc_w_off),
                         reinterpret_cast<const float*>(pl->arena + pl->fc_b_off), out_norm, out_px, B,
                         c.Ho * c.Wo, c.Cout, 2 * pl->n_kp, pl->H, pl->W, pl->split ? 2 : (pl->f16 ? 3 : (pl->bf16 ? 1 : 0)), st);
    }
    default:
      return PB200_OK;
  }
}

static inline bool front_chunked(const pb200_plan* pl, int B) {
  return pl->n_front > 0 && pl->l1_chunk > 0 && B > pl->l1_chunk;
}

static int check_forward_args(pb200_plan* plan, const void* frames, void* ws, float* out_norm, int B) {
  PB_CHECK(plan && frames && ws && out_norm, "forward: null argument");
  PB_CHECK(B >= 1 && B <= plan->max_batch, "forward: batch %d outside [1, max_batch=%d]", B,
           plan->max_batch);
  if (!plan->loaded) {
    set_error("forward: pb200_plan_load_weights has not been called");
    return PB200_ERR_STATE;
  }
  return PB200_OK;
}

}  // namespace pb200

extern "C" int pb200_forward(pb200_plan* plan, const void* frames, void* workspace, float* out_norm,
                             float* out_px_or_null, int B, pb200_stream stream) {
  if (int rc = check_forward_args(plan, frames, workspace, out_norm, B)) return rc;
  if (int rc = prepare(plan, frames, workspace, B)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  size_t first = 0;
  if (front_chunked(plan, B)) {
    for (int n0 = 0; n0 < B; n0 += plan->l1_chunk) {
      const int Bc = B - n0 < plan->l1_chunk ? B - n0 : plan->l1_chunk;
      for (int i = 0; i < plan->n_front; ++i)
        if (int rc = run_op(plan, plan->ops[i], frames, workspace, out_norm, out_px_or_null, Bc, st, n0)) return rc;
    }
    first = plan->n_front;
  }
  for (size_t i = first; i < plan->ops.size(); ++i) {
    if (!plan->chain_at.empty() && plan->chain_at[i] >= 0) {  // layers 3-4: one persistent launch for the whole run
      if (int rc = launch_chain_at(plan, i, frames, workspace, B, st)) return rc;
      i += plan->chain_len[i] - 1;
      continue;
    }
    if (int rc = run_op(plan, plan->ops[i], frames, workspace, out_norm, out_px_or_null, B, st)) return rc;
  }
  return PB200_OK;
}

extern "C" int pb200_forward_rgb8(pb200_plan* plan, const uint8_t* rgb, int bgr, void* workspace, float* out_norm,
                                  float* out_px_or_null, int B, pb200_stream stream) {
  if (int rc = check_forward_args(plan, rgb, workspace, out_norm, B)) return rc;
  PB_CHECK(plan->bf16 && plan->C == 3, "forward_rgb8: needs a 16-bit tcgen05 plan (bf16 / fp16x2) with 3 input channels; "
           "use pb200_ingest + pb200_forward otherwise");
  PB_CHECK((reinterpret_cast<uintptr_t>(rgb) & 15) == 0, "forward_rgb8: the uint8 batch must be 16-byte aligned");
  plan->rgb8_src = rgb;
  plan->rgb8_bgr = bgr ? 1 : 0;
  // the frames pointer is never dereferenced (the fused stem reads `rgb`); it only has to be non-null
  const int rc = pb200_forward(plan, rgb, workspace, out_norm, out_px_or_null, B, stream);
  plan->rgb8_src = nullptr;
  return rc;
}

extern "C" int pb200_forward_nchw_f32(pb200_plan* plan, const float* x_nchw, void* workspace,
                                      float* out_norm, float* out_px_or_null, int B,
                                      pb200_stream stream) {
  if (int rc = check_forward_args(plan, x_nchw, workspace, out_norm, B)) return rc;
  void* frames = static_cast<uint8_t*>(workspace) + plan->buf_off[BUF_FRAMES];
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (int rc = launch_repack_nchw(x_nchw, plan->C, plan->H, plan->W, B, frames, frame_elt(plan), st)) return rc;
  return pb200_forward(plan, frames, workspace, out_norm, out_px_or_null, B, stream);
}

extern "C" int pb200_plan_num_ops(const pb200_plan* plan) {
  return plan ? static_cast<int>(plan->ops.size()) : 0;
}
extern "C" const char* pb200_plan_op_name(const pb200_plan* plan, int i) {
  if (!plan || i < 0 || i >= static_cast<int>(plan->ops.size())) return "";
  return plan->op_names[i].c_str();
}
// ops launched together starting at op i: 1 = its own launch, k >= 2 = head of a chained launch covering ops i..i+k-1,
// 0 = covered by a chain that started earlier (known once a forward has been prepared at a batch size)
extern "C" int pb200_plan_op_chain_len(const pb200_plan* plan, int i) {
  if (!plan || i < 0 || i >= static_cast<int>(plan->ops.size())) return 0;
  if (plan->chain_at.size() != plan->ops.size()) return 1;
  for (int h = 0; h <= i; ++h)
    if (plan->chain_at[h] >= 0 && i < h + plan->chain_len[h]) return h == i ? plan->chain_len[h] : 0;
  return 1;
}
extern "C" const char* pb200_plan_op_kernel(const pb200_plan* plan, int i) {
  if (!plan || i < 0 || i >= static_cast<int>(plan->ops.size())) return "";
  const Op& op = plan->ops[i];
  pb200_plan* pl = const_cast<pb200_plan*>(plan);
  if (pl->op_kernels.size() != plan->ops.size()) pl->op_kernels.assign(plan->ops.size(), "");
  std::string& k = pl->op_kernels[i];
  switch (op.kind) {
    case OP_STEM_POOL: k = plan->f16 ? "stem_pool_x2_kernel" : "stem_pool_kernel"; break;
    case OP_CONV_P8: k = plan->f16 ? "conv_p8_x2_kernel" : "conv_p8_kernel"; break;
    case OP_MAXPOOL: k = plan->split ? "maxpool_split_kernel" : (plan->bf16 ? "maxpool_bf16_kernel" : "maxpool_f32_kernel"); break;
    case OP_HEAD: k = "head_kernel"; break;
    case OP_CONV:
      if (pb200_plan_op_chain_len(plan, i) != 1) {
        k = plan->f16 ? "conv_chain_kernel<x2>" : "conv_chain_kernel";
        break;
      }
      if (!plan->bf16 && !(plan->split && !plan->convs[op.conv].stem))
        k = (plan->convs[op.conv].stem && plan->H % 16 == 0 && plan->W % 64 == 0) ? "stem_f32_kernel" : "conv_simt_kernel";
      else if (plan->cache_B < 0 || plan->umma.size() != plan->convs.size()) k = "conv_umma*";
      else if (plan->halo.size() == plan->convs.size() && plan->halo[op.conv].grid > 0) k = "conv_halo_kernel";
      else {
        const UmmaConvLaunch& L = plan->umma[op.conv];
        k = std::string(L.pair ? "conv_umma2_kernel<" : "conv_umma_kernel<") + std::to_string(L.block_n) + (L.split == 1 ? ",split>" : (L.split == 2 ? ",x2>" : (L.resb ? ",resB>" : ">")));
      }
      break;
    default: k = ""; break;
  }
  return k.c_str();
}
extern "C" int pb200_plan_launches_per_forward(const pb200_plan* plan) {
  return plan ? static_cast<int>(plan->ops.size()) : 0;  // at batch sizes <= l1_chunk; see pb200_plan_launches_at
}
extern "C" int pb200_plan_launches_at(const pb200_plan* plan, int B) {
  if (!plan) return 0;
  int n = static_cast<int>(plan->ops.size());
  if (plan->cache_B == B)  // chains exist once a forward has been prepared at this batch size
    for (size_t i = 0; i < plan->chain_len.size(); ++i)
      if (plan->chain_at[i] >= 0) n -= plan->chain_len[i] - 1;
  if (!front_chunked(plan, B)) return n;
  const int chunks = (B + plan->l1_chunk - 1) / plan->l1_chunk;
  return n - plan->n_front + chunks * plan->n_front;
}

extern "C" int pb200_plan_op_work(const pb200_plan* plan, int i, int B, double* flops, double* bytes) {
  PB_CHECK(plan && i >= 0 && i < static_cast<int>(plan->ops.size()) && flops && bytes, "op_work: bad argument");
  const Op& op = plan->ops[i];
  const double e = static_cast<double>(plan->elt);
  *flops = 0.0;
  *bytes = 0.0;
  if (op.kind == OP_STEM_POOL) {
    const ConvSpec& c = plan->convs[op.conv];
    *flops = 2.0 * B * c.Ho * c.Wo * c.Cout * c.R * c.R * c.Cin;
    *bytes = e * (static_cast<double>(B) * (plan->H + 6) * (plan->W + 8) * 4 + B * (c.Ho / 2.0) * (c.Wo / 2.0) * c.Cout +
                  static_cast<double>(c.Cout) * c.R * c.R * c.Cin);
  } else if (op.kind == OP_CONV || op.kind == OP_CONV_P8) {
    const ConvSpec& c = plan->convs[op.conv];
    const double M = static_cast<double>(B) * c.Ho * c.Wo;
    const double ds_cin = c.ds_conv >= 0 ? plan->convs[c.ds_conv].Cin : 0.0;  // fused projection: K-extension
    *flops = 2.0 * M * c.Cout * (c.R * c.R * c.Cin + ds_cin);
    const double in_el = c.stem ? static_cast<double>(B) * (plan->H + 6) * (plan->W + 8) * 4
                                : static_cast<double>(B) * c.Hin * c.Win * c.Cin;
    *bytes = e * (in_el + M * ds_cin + M * c.Cout * (op.res_buf >= 0 ? 2.0 : 1.0) +
                  static_cast<double>(c.Cout) * (c.R * c.R * c.Cin + ds_cin));
  } else if (op.kind == OP_MAXPOOL) {
    const ConvSpec& c = plan->convs[0];
    *bytes = e * static_cast<double>(B) * c.Cout * (c.Ho * c.Wo + (c.Ho / 2) * (c.Wo / 2));
  } else if (op.kind == OP_HEAD) {
    const ConvSpec& c = plan->convs.back();
    *flops = 2.0 * B * 512.0 * 2 * plan->n_kp;
    *bytes = e * static_cast<double>(B) * c.Ho * c.Wo * c.Cout + 4.0 * B * 2 * plan->n_kp * 2;
  }
  return PB200_OK;
}

extern "C" int pb200_forward_profile(pb200_plan* plan, const void* frames, void* workspace,
                                     float* out_norm, float* out_px_or_null, int B,
                                     pb200_stream stream, float* ms) {
  if (int rc = check_forward_args(plan, frames, workspace, out_norm, B)) return rc;
  PB_CHECK(ms != nullptr, "forward_profile: null ms");
  if (int rc = prepare(plan, frames, workspace, B)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int n = static_cast<int>(plan->ops.size());
  // launch list in execution order: (op, first frame, frames); the chunked front ops appear once per L2 chunk and their
  // times are summed per op
  struct Launch { int op, n0, b; };
  std::vector<Launch> seq;
  int first = 0;
  if (front_chunked(plan, B)) {
    for (int n0 = 0; n0 < B; n0 += plan->l1_chunk)
      for (int i = 0; i < plan->n_front; ++i) seq.push_back({i, n0, B - n0 < plan->l1_chunk ? B - n0 : plan->l1_chunk});
    first = plan->n_front;
  }
  for (int i = first; i < n; ++i)
    if (pb200_plan_op_chain_len(plan, i) != 0) seq.push_back({i, 0, B});  // a chain is timed as ONE launch at its head op
  std::vector<cudaEvent_t> ev(seq.size() + 1);
  for (auto& e : ev) PB_CUDA(cudaEventCreate(&e));
  int rc = PB200_OK;
  PB_CUDA(cudaEventRecord(ev[0], st));
  for (size_t j = 0; j < seq.size() && rc == PB200_OK; ++j) {
    if (pb200_plan_op_chain_len(plan, seq[j].op) >= 2) rc = launch_chain_at(plan, seq[j].op, frames, workspace, B, st);
    else rc = run_op(plan, plan->ops[seq[j].op], frames, workspace, out_norm, out_px_or_null, seq[j].b, st, seq[j].n0);
    if (rc == PB200_OK && cudaEventRecord(ev[j + 1], st) != cudaSuccess) rc = PB200_ERR_CUDA;
  }
  if (rc == PB200_OK) {
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) rc = cuda_fail(e, "cudaStreamSynchronize", __FILE__, __LINE__);
  }
  if (rc == PB200_OK) {
    for (int i = 0; i < n; ++i) ms[i] = 0.f;
    for (size_t j = 0; j < seq.size(); ++j) {
      float t = 0.f;
      cudaEventElapsedTime(&t, ev[j], ev[j + 1]);
      ms[seq[j].op] += t;
    }
  }
  for (auto& e : ev) cudaEventDestroy(e);
  return rc;
}

// ------------------------------------------------------------------------------------------
// ingest + the single-layer debug entry
// ------------------------------------------------------------------------------------------
extern "C" int pb200_ingest(const pb200_plan* plan, const uint8_t* rgb, int rgb_channels,
                            int64_t rgb_pitch_bytes, int src_h, int src_w, int bgr, const float* depth,
                            float depth_divisor, int clip, float cube_scale, float near_plane,
                            float far_plane, const float* jitter_or_null, void* out_frames, int B,
                            pb200_stream stream) {
  PB_CHECK(plan && rgb && out_frames, "ingest: null argument");
  PB_CHECK(B >= 1, "ingest: B = %d", B);
  PB_CHECK(rgb_channels == 3 || rgb_channels == 4, "ingest: rgb_channels must be 3 or 4, got %d", rgb_channels);
  PB_CHECK(src_h >= plan->H && src_w >= plan->W, "ingest: source %dx%d smaller than the %dx%d window",
           src_h, src_w, plan->H, plan->W);
  PB_CHECK(rgb_pitch_bytes >= static_cast<int64_t>(src_w) * rgb_channels, "ingest: row pitch too small");
  PB_CHECK(depth == nullptr || depth_divisor != 0.f, "ingest: depth_divisor is 0");
  PB_CHECK(!clip || cube_scale != 0.f, "ingest: cube_scale is 0");
  PB_CHECK(plan->W % 4 == 0, "ingest: W must be a multiple of 4");
  if (int rc = require_device()) return rc;
  IngestArgs a;
  a.rgb = rgb;
  a.pitch = rgb_pitch_bytes;
  a.frame_pitch = rgb_pitch_bytes * src_h;
  a.ch = rgb_channels;
  a.bgr = bgr;
  a.src_h = src_h;
  a.src_w = src_w;
  a.y0 = src_h / 2 - plan->H / 2;  // streaming.py:79-80
  a.x0 = src_w / 2 - plan->W / 2;
  a.depth = depth;
  a.depth_divisor = depth_divisor;
  a.clip = clip;
  a.cube_scale = cube_scale;
  a.near_plane = near_plane;
  a.far_plane = far_plane;
  a.jitter = jitter_or_null;
  a.H = plan->H;
  a.W = plan->W;
  a.B = B;
  a.rgb_integer = plan->bf16 ? 1 : 0;
  return launch_ingest(a, out_frames, frame_elt(plan), static_cast<cudaStream_t>(stream));
}

// the single-layer test entries pack their weights in the KT layout unless a multicast cluster is forced (those CTAs hold
// row slices of ONE tile and keep the row-major tensor map) or PB200_NO_BTILED asks for the row-major path
static bool debug_b_tiled() {
  if (getenv("PB200_NO_BTILED")) return false;
  const char* f = getenv("PB200_FORCE_CLUSTER");
  return !(f != nullptr && f[1] == 0 && (f[0] == '2' || f[0] == '4'));
}

extern "C" int pb200_debug_conv_bf16(const void* in_nhwc, int B, int H, int W, int Cin,
                                     const float* w_oihw_host, const float* bias_host, int Cout, int R,
                                     int stride, int pad, const void* residual_nhwc_or_null, int relu,
                                     void* out_nhwc, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  if (int rc = umma_conv_init()) return rc;
  PB_CHECK(in_nhwc && w_oihw_host && bias_host && out_nhwc, "debug_conv: null argument");
  const size_t K = static_cast<size_t>(R) * R * Cin;
  const bool tiled = debug_b_tiled() && K % 64 == 0;
  std::vector<uint16_t> wp(K * Cout);
  for (int co = 0; co < Cout; ++co)
    for (int ci = 0; ci < Cin; ++ci)
      for (int r = 0; r < R; ++r)
        for (int s = 0; s < R; ++s) {
          const size_t k = (static_cast<size_t>(r) * R + s) * Cin + ci;
          wp[tiled ? umma_pack_kt_offset(0, co, static_cast<int>(k), Cout, static_cast<int>(K)) / 2 : co * K + k] =
              f32_to_bf16_rne(w_oihw_host[((static_cast<size_t>(co) * Cin + ci) * R + r) * R + s]);
        }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void *dw = nullptr, *db = nullptr;
  PB_CUDA(cudaMalloc(&dw, wp.size() * 2));
  PB_CUDA(cudaMalloc(&db, sizeof(float) * Cout));
  int rc = PB200_OK;
  if (cudaMemcpyAsync(dw, wp.data(), wp.size() * 2, cudaMemcpyHostToDevice, st) != cudaSuccess ||
      cudaMemcpyAsync(db, bias_host, sizeof(float) * Cout, cudaMemcpyHostToDevice, st) != cudaSuccess)
    rc = PB200_ERR_CUDA;
  UmmaConvLaunch L;
  if (rc == PB200_OK) {
    int bn = Cout < 256 ? Cout : 256;
    rc = build_umma_conv(&L, in_nhwc, B, H, W, Cin, dw, static_cast<const float*>(db),
                         residual_nhwc_or_null, out_nhwc, Cout, R, stride, pad, relu, bn, 0, 0, nullptr, 0, 0, 0, nullptr,
                         tiled ? 1 : 0);
  }
  if (rc == PB200_OK) rc = launch_umma_conv(L, st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (rc == PB200_OK && e != cudaSuccess) rc = cuda_fail(e, "debug_conv sync", __FILE__, __LINE__);
  cudaFree(dw);
  cudaFree(db);
  return rc;
}

extern "C" int pb200_debug_conv_ds_bf16(const void* in_nhwc, int B, int H, int W, int Cin, const float* w_oihw_host,
                                        const float* bias_host, int Cout, const void* ds_in_nhwc, int ds_Cin,
                                        const float* wds_oi_host, int relu, void* out_nhwc, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  if (int rc = umma_conv_init()) return rc;
  PB_CHECK(in_nhwc && w_oihw_host && bias_host && ds_in_nhwc && wds_oi_host && out_nhwc, "debug_conv_ds: null argument");
  const int R = 3;
  const size_t K = static_cast<size_t>(R) * R * Cin, Krow = K + ds_Cin;  // K-extension: [taps x Cin | ds_Cin]
  std::vector<uint16_t> wp(Krow * Cout);
  const bool tiled = debug_b_tiled();
  auto at = [&](int co, size_t k) -> uint16_t& {
    return wp[tiled ? umma_pack_kt_offset(0, co, static_cast<int>(k), Cout, static_cast<int>(Krow)) / 2 : co * Krow + k];
  };
  for (int co = 0; co < Cout; ++co) {
    for (int ci = 0; ci < Cin; ++ci)
      for (int r = 0; r < R; ++r)
        for (int s = 0; s < R; ++s)
          at(co, (static_cast<size_t>(r) * R + s) * Cin + ci) =
              f32_to_bf16_rne(w_oihw_host[((static_cast<size_t>(co) * Cin + ci) * R + r) * R + s]);
    for (int ci = 0; ci < ds_Cin; ++ci) at(co, K + ci) = f32_to_bf16_rne(wds_oi_host[static_cast<size_t>(co) * ds_Cin + ci]);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void *dw = nullptr, *db = nullptr;
  PB_CUDA(cudaMalloc(&dw, wp.size() * 2));
  PB_CUDA(cudaMalloc(&db, sizeof(float) * Cout));
  int rc = PB200_OK;
  if (cudaMemcpyAsync(dw, wp.data(), wp.size() * 2, cudaMemcpyHostToDevice, st) != cudaSuccess ||
      cudaMemcpyAsync(db, bias_host, sizeof(float) * Cout, cudaMemcpyHostToDevice, st) != cudaSuccess)
    rc = PB200_ERR_CUDA;
  UmmaConvLaunch L;
  if (rc == PB200_OK) {
    int bn = Cout < 256 ? Cout : 256;
    rc = build_umma_conv(&L, in_nhwc, B, H, W, Cin, dw, static_cast<const float*>(db), nullptr, out_nhwc, Cout, R, 1, 1,
                         relu, bn, 0, 0, ds_in_nhwc, 2 * H, 2 * W, ds_Cin, nullptr, tiled ? 1 : 0);
  }
  if (rc == PB200_OK) rc = launch_umma_conv(L, st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (rc == PB200_OK && e != cudaSuccess) rc = cuda_fail(e, "debug_conv_ds sync", __FILE__, __LINE__);
  cudaFree(dw);
  cudaFree(db);
  return rc;
}

extern "C" int pb200_debug_conv_split(const void* in_nhwc2, int B, int H, int W, int Cin,
                                      const float* w_oihw_host, const float* bias_host, int Cout, int R,
                                      int stride, int pad, const void* residual_nhwc2_or_null, int relu,
                                      void* out_nhwc2, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  if (int rc = umma_conv_init()) return rc;
  PB_CHECK(in_nhwc2 && w_oihw_host && bias_host && out_nhwc2, "debug_conv_split: null argument");
  const size_t K = static_cast<size_t>(R) * R * Cin;
  std::vector<uint16_t> wp(2 * K * Cout);
  for (int co = 0; co < Cout; ++co)
    for (int ci = 0; ci < Cin; ++ci)
      for (int r = 0; r < R; ++r)
        for (int s = 0; s < R; ++s) {
          const float wf = w_oihw_host[((static_cast<size_t>(co) * Cin + ci) * R + r) * R + s];
          const uint16_t hi = f32_to_bf16_rne(wf);
          const uint32_t hbits = static_cast<uint32_t>(hi) << 16;
          float hf;
          memcpy(&hf, &hbits, 4);
          const size_t k = (static_cast<size_t>(r) * R + s) * Cin + ci;
          wp[co * 2 * K + k] = hi;
          wp[co * 2 * K + K + k] = f32_to_bf16_rne(wf - hf);
        }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void *dw = nullptr, *db = nullptr;
  PB_CUDA(cudaMalloc(&dw, wp.size() * 2));
  PB_CUDA(cudaMalloc(&db, sizeof(float) * Cout));
  int rc = PB200_OK;
  if (cudaMemcpyAsync(dw, wp.data(), wp.size() * 2, cudaMemcpyHostToDevice, st) != cudaSuccess ||
      cudaMemcpyAsync(db, bias_host, sizeof(float) * Cout, cudaMemcpyHostToDevice, st) != cudaSuccess)
    rc = PB200_ERR_CUDA;
  UmmaConvLaunch L;
  if (rc == PB200_OK) {
    int bn = Cout < 256 ? Cout : 256;
    rc = build_umma_conv(&L, in_nhwc2, B, H, W, Cin, dw, static_cast<const float*>(db),
                         residual_nhwc2_or_null, out_nhwc2, Cout, R, stride, pad, relu, bn, 0, 1);
  }
  if (rc == PB200_OK) rc = launch_umma_conv(L, st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (rc == PB200_OK && e != cudaSuccess) rc = cuda_fail(e, "debug_conv_split sync", __FILE__, __LINE__);
  cudaFree(dw);
  cudaFree(db);
  return rc;
}

extern "C" int pb200_debug_stem_pool(pb200_plan* plan, const void* frames, void* out_p8, int B,
                                     pb200_stream stream) {
  PB_CHECK(plan && frames && out_p8 && plan->bf16 && plan->loaded, "debug_stem_pool: needs a loaded bf16 / fp16x2 plan");
  const ConvSpec& c = plan->convs[0];
  return launch_stem_pool(frames, plan->arena + c.w_off, reinterpret_cast<const float*>(plan->arena + c.b_off),
                          out_p8, B, plan->H, plan->W, static_cast<cudaStream_t>(stream), plan->f16 ? 1 : 0);
}

extern "C" int pb200_debug_conv_halo(const void* in_nhwc, int B, int H, int W, int Cin, const float* w_oihw_host,
                                     const float* bias_host, int Cout, const void* residual_nhwc_or_null, int relu,
                                     void* out_nhwc, const void* ds_in_nhwc_or_null, int ds_Cin,
                                     const float* wds_oi_host_or_null, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(in_nhwc && w_oihw_host && bias_host && out_nhwc, "debug_conv_halo: null argument");
  const bool ds = ds_in_nhwc_or_null != nullptr;
  PB_CHECK(!ds || wds_oi_host_or_null != nullptr, "debug_conv_halo: projection weights missing");
  PB_CHECK(conv_halo_supported(B, H, W, Cin, Cout, 3, 1, 1, ds ? ds_Cin : 0), "debug_conv_halo: unsupported shape");
  const int K = 9 * Cin, Krow = K + (ds ? ds_Cin : 0);
  std::vector<uint16_t> wp(static_cast<size_t>(Krow) * Cout);
  for (int co = 0; co < Cout; ++co) {
    for (int ci = 0; ci < Cin; ++ci)
      for (int r = 0; r < 3; ++r)
        for (int s = 0; s < 3; ++s)
          wp[umma_pack_kt_offset(0, co, (r * 3 + s) * Cin + ci, Cout, Krow) / 2] =
              f32_to_bf16_rne(w_oihw_host[((static_cast<size_t>(co) * Cin + ci) * 3 + r) * 3 + s]);
    for (int ci = 0; ds && ci < ds_Cin; ++ci)
      wp[umma_pack_kt_offset(0, co, K + ci, Cout, Krow) / 2] = f32_to_bf16_rne(wds_oi_host_or_null[static_cast<size_t>(co) * ds_Cin + ci]);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void *dw = nullptr, *db = nullptr;
  PB_CUDA(cudaMalloc(&dw, wp.size() * 2));
  PB_CUDA(cudaMalloc(&db, sizeof(float) * Cout));
  int rc = PB200_OK;
  if (cudaMemcpyAsync(dw, wp.data(), wp.size() * 2, cudaMemcpyHostToDevice, st) != cudaSuccess ||
      cudaMemcpyAsync(db, bias_host, sizeof(float) * Cout, cudaMemcpyHostToDevice, st) != cudaSuccess)
    rc = PB200_ERR_CUDA;
  HaloLaunch L;
  if (rc == PB200_OK)
    rc = build_conv_halo(&L, in_nhwc, B, H, W, Cin, dw, static_cast<const float*>(db), residual_nhwc_or_null, out_nhwc, Cout,
                         relu, ds_in_nhwc_or_null, ds ? ds_Cin : 0);
  if (rc == PB200_OK) rc = launch_conv_halo(L, st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (rc == PB200_OK && e != cudaSuccess) rc = cuda_fail(e, "debug_conv_halo sync", __FILE__, __LINE__);
  cudaFree(dw);
  cudaFree(db);
  return rc;
}

extern "C" int pb200_debug_conv_f16x2(const void* in_nhwc, int B, int H, int W, int Cin, const float* w_oihw_host,
                                      const float* bias_host, int Cout, int R, int stride, int pad,
                                      const void* residual_nhwc_or_null, int relu, void* out_nhwc,
                                      const void* ds_in_nhwc_or_null, int ds_Cin, const float* wds_oi_host_or_null,
                                      pb200_stream stream) {
  if (int rc = require_device()) return rc;
  if (int rc = umma_conv_init()) return rc;
  PB_CHECK(in_nhwc && w_oihw_host && bias_host && out_nhwc, "debug_conv_f16x2: null argument");
  const bool ds = ds_in_nhwc_or_null != nullptr;
  PB_CHECK(!ds || (wds_oi_host_or_null != nullptr && R == 3 && stride == 1 && pad == 1), "debug_conv_f16x2: bad projection");
  const size_t K = static_cast<size_t>(R) * R * Cin, Krow = K + (ds ? ds_Cin : 0);
  std::vector<uint16_t> wp(2 * Krow * Cout);  // [Cout][hi Krow | lo Krow], or the KT layout (hi slabs, then lo slabs)
  const bool tiled = debug_b_tiled();
  auto put = [&](int co, size_t k, float wf) {
    const uint16_t hi = f32_to_f16_rne(wf);
    const uint16_t lo = f32_to_f16_rne(wf - f16_bits_to_f32(hi));
    if (tiled) {
      wp[umma_pack_kt_offset(0, co, static_cast<int>(k), Cout, static_cast<int>(Krow)) / 2] = hi;
      wp[umma_pack_kt_offset(1, co, static_cast<int>(k), Cout, static_cast<int>(Krow)) / 2] = lo;
    } else {
      wp[co * 2 * Krow + k] = hi;
      wp[co * 2 * Krow + Krow + k] = lo;
    }
  };
  for (int co = 0; co < Cout; ++co) {
    for (int ci = 0; ci < Cin; ++ci)
      for (int r = 0; r < R; ++r)
        for (int s = 0; s < R; ++s)
          put(co, (static_cast<size_t>(r) * R + s) * Cin + ci, w_oihw_host[((static_cast<size_t>(co) * Cin + ci) * R + r) * R + s]);
    for (int ci = 0; ds && ci < ds_Cin; ++ci) put(co, K + ci, wds_oi_host_or_null[static_cast<size_t>(co) * ds_Cin + ci]);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void *dw = nullptr, *db = nullptr;
  PB_CUDA(cudaMalloc(&dw, wp.size() * 2));
  PB_CUDA(cudaMalloc(&db, sizeof(float) * Cout));
  int rc = PB200_OK;
  if (cudaMemcpyAsync(dw, wp.data(), wp.size() * 2, cudaMemcpyHostToDevice, st) != cudaSuccess ||
      cudaMemcpyAsync(db, bias_host, sizeof(float) * Cout, cudaMemcpyHostToDevice, st) != cudaSuccess)
    rc = PB200_ERR_CUDA;
  UmmaConvLaunch L;
  if (rc == PB200_OK) {
    int bn = Cout < 256 ? Cout : 256;
    const char* fbn = getenv("PB200_FORCE_BLOCK_N");  // tests: exercise every tile width
    if (fbn != nullptr && atoi(fbn) > 0 && Cout % atoi(fbn) == 0) bn = atoi(fbn);
    rc = build_umma_conv(&L, in_nhwc, B, H, W, Cin, dw, static_cast<const float*>(db), residual_nhwc_or_null, out_nhwc,
                         Cout, R, stride, pad, relu, bn, 0, 2, ds_in_nhwc_or_null, ds ? 2 * H : 0, ds ? 2 * W : 0,
                         ds ? ds_Cin : 0, nullptr, tiled ? 1 : 0);
  }
  if (rc == PB200_OK) rc = launch_umma_conv(L, st);
  cudaError_t e = cudaStreamSynchronize(st);
  if (rc == PB200_OK && e != cudaSuccess) rc = cuda_fail(e, "debug_conv_f16x2 sync", __FILE__, __LINE__);
  cudaFree(dw);
  cudaFree(db);
  return rc;
}

extern "C" int pb200_debug_conv_p8_f16x2(const void* in_p8, int B, int H, int W, const float* w_oihw_host,
                                         const float* bias_host, const void* residual_p8_or_null, int relu,
                                         void* out, int out_nhwc, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  if (int rc = conv_p8_init()) return rc;
  PB_CHECK(in_p8 && w_oihw_host && bias_host && out, "debug_conv_p8_f16x2: null argument");
  const size_t img_bytes = 64 * 576 * 2;
  std::vector<uint8_t> img(2 * img_bytes);  // hi image, then the lo image
  for (int co = 0; co < 64; ++co)
    for (int ci = 0; ci < 64; ++ci)
      for (int r = 0; r < 3; ++r)
        for (int s = 0; s < 3; ++s) {
          const float wf = w_oihw_host[((static_cast<size_t>(co) * 64 + ci) * 3 + r) * 3 + s];
          const uint16_t hi = f32_to_f16_rne(wf);
          const size_t off = conv_p8_wimg_offset(co, (r * 3 + s) * 64 + ci);
          *reinterpret_cast<uint16_t*>(img.data() + off) = hi;
          *reinterpret_cast<uint16_t*>(img.data() + img_bytes + off) = f32_to_f16_rne(wf - f16_bits_to_f32(hi));
        }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void *dw = nullptr, *db = nullptr;
  PB_CUDA(cudaMalloc(&dw, img.size()));
  PB_CUDA(cudaMalloc(&db, sizeof(float) * 64));
  int rc = PB200_OK;
  if (cudaMemcpyAsync(dw, img.data(), img.size(), cudaMemcpyHostToDevice, st) != cudaSuccess ||
      cudaMemcpyAsync(db, bias_host, sizeof(float) * 64, cudaMemcpyHostToDevice, st) != cudaSuccess)
    rc = PB200_ERR_CUDA;
  if (rc == PB200_OK)
    rc = launch_conv_p8(in_p8, dw, static_cast<const float*>(db), residual_p8_or_null, out, out_nhwc, relu, B, H, W, st, 1,
                        getenv("PB200_DEBUG_P8_REVERSE") != nullptr ? 1 : 0);
  cudaError_t e = cudaStreamSynchronize(st);
  if (rc == PB200_OK && e != cudaSuccess) rc = cuda_fail(e, "debug_conv_p8_f16x2 sync", __FILE__, __LINE__);
  cudaFree(dw);
  cudaFree(db);
  return rc;
}

extern "C" int pb200_debug_conv_p8(const void* in_p8, int B, int H, int W, const float* w_oihw_host,
                                   const float* bias_host, const void* residual_p8_or_null, int relu,
                                   void* out, int out_nhwc, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  if (int rc = conv_p8_init()) return rc;
  PB_CHECK(in_p8 && w_oihw_host && bias_host && out, "debug_conv_p8: null argument");
  std::vector<uint8_t> img(64 * 576 * 2);
  for (int co = 0; co < 64; ++co)
    for (int ci = 0; ci < 64; ++ci)
      for (int r = 0; r < 3; ++r)
        for (int s = 0; s < 3; ++s)
          *reinterpret_cast<uint16_t*>(img.data() + conv_p8_wimg_offset(co, (r * 3 + s) * 64 + ci)) =
              f32_to_bf16_rne(w_oihw_host[((static_cast<size_t>(co) * 64 + ci) * 3 + r) * 3 + s]);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  void *dw = nullptr, *db = nullptr;
  PB_CUDA(cudaMalloc(&dw, img.size()));
  PB_CUDA(cudaMalloc(&db, sizeof(float) * 64));
  int rc = PB200_OK;
  if (cudaMemcpyAsync(dw, img.data(), img.size(), cudaMemcpyHostToDevice, st) != cudaSuccess ||
      cudaMemcpyAsync(db, bias_host, sizeof(float) * 64, cudaMemcpyHostToDevice, st) != cudaSuccess)
    rc = PB200_ERR_CUDA;
  if (rc == PB200_OK)
    rc = launch_conv_p8(in_p8, dw, static_cast<const float*>(db), residual_p8_or_null, out, out_nhwc, relu, B, H,
                        W, st,
                        0, getenv("PB200_DEBUG_P8_REVERSE") != nullptr ? 1 : 0);
  cudaError_t e = cudaStreamSynchronize(st);
  if (rc == PB200_OK && e != cudaSuccess) rc = cuda_fail(e, "debug_conv_p8 sync", __FILE__, __LINE__);
  cudaFree(dw);
  cudaFree(db);
  return rc;
}
