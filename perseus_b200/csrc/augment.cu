// Train-time augmentation of a batch, fused: KeypointAugmentation.forward
// (perseus/detector/augmentations.py:434-488) in two launches.
//
//   transplant_count_kernel : RandomTransplantationWithDepth's acceptance statistic (augmentations.py:172-251):
//                             per image, the number of cube pixels the transplanted segmentation would have
//   augment_kernel          : per 32x32 output tile (+2 halo for the 5x5 blur), every stage of the pipeline on
//                             the fly from the ORIGINAL batch:
//                               transplantation select (:203-247)  ->  affine warp, bilinear, zero padding
//                               (kornia RandomAffine, :351-359)    ->  two rectangle erasures (:361-372)
//                               -> Planckian gains (:375-376)      ->  colour jiggle in its sampled order (:378-386)
//                               -> 5x5 Gaussian blur, reflect border (:388-389)  ->  plasma shadow (:391-392)
//                               -> depth bias / Gaussian noise / near + far planes (:17-169, 395-431)
//                             and the key points mapped by the same affine matrix and normalised (:481).
//
// All PER-SAMPLE random draws (donor, matrix, rectangles, gains, order, sigma, shadow) are explicit inputs sampled
// by the host (perseus_b200/detector/augmentations.py); the PER-PIXEL depth noise comes from Philox4x32-10 keyed
// by (seed; pixel, image, stream) so that oracle/augment_oracle.py reproduces it bit for bit.
// Input  : float [B][C_in][H][W] (NCHW, the reference's own batch: RGB in [0,1], depth in data units, seg 0/1)
// Output : float [B][C_out][H][W], C_out <= C_in  (train.py:284 slices the channels the model uses)
// Memory-bound gather: ~40 B/pixel read through L1/L2 (acceptor + donor, 5 planes), 4*C_out B/pixel written.
#include "common.cuh"

namespace pb200 {

constexpr int kAugTile = 32;
constexpr int kAugHalo = 2;
constexpr int kAugSpan = kAugTile + 2 * kAugHalo;  // 36
constexpr int kAugThreads = 256;
constexpr int kAugIP = PB200_AUG_IPARAMS;
constexpr int kAugFP = PB200_AUG_FPARAMS;

struct AugArgs {
  const float* images;
  const int* iparams;
  const float* fparams;
  const int* counts;     // transplant_count_kernel's output, or null
  const float* plasma;   // [B][H][W] or null
  float* out;
  int B, C_in, C_out, H, W;
  unsigned long long seed;
  pb200_aug_config cfg;
  float drop[6];  // (p, keep) of the bias / near / far dropout masks
};

// ---- Philox4x32-10 (Salmon et al. 2011), the same generator family torch.cuda uses -----------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const unsigned hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const unsigned hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ float u01(unsigned w) { return static_cast<float>(w >> 8) * 5.9604644775390625e-8f; }  // [0,1)

// ---- transplantation (augmentations.py:203-240) at one source pixel ----------------------------------------------
struct Px5 {
  float r, g, b, d, s;
};

__device__ __forceinline__ Px5 load5(const float* __restrict__ img, long long plane, int C) {
  Px5 p;
  p.r = __ldg(img);
  p.g = __ldg(img + plane);
  p.b = __ldg(img + 2 * plane);
  p.d = C > 3 ? __ldg(img + 3 * plane) : 0.f;
  p.s = C > 4 ? __ldg(img + 4 * plane) : 0.f;
  return p;
}

// acceptor pixel a, donor pixel d -> the transplanted pixel
__device__ __forceinline__ Px5 transplant(const Px5& a, const Px5& d) {
  const bool acc_cube = a.s == 1.0f;                       // :217
  bool mask = !acc_cube;                                   // :218
  const float da = acc_cube ? a.d : a.d * 0.f;             // :222  depth * ind (0 * x keeps NaN semantics)
  const float dd = acc_cube ? d.d : d.d * 0.f;             // :223
  if (dd < da) mask = true;                                // :224-225
  const bool don_cube = d.s == 1.0f;                       // :228
  if (don_cube) mask = false;                              // :229
  Px5 o;
  o.r = mask ? d.r : a.r;                                  // :232-234
  o.g = mask ? d.g : a.g;
  o.b = mask ? d.b : a.b;
  o.d = mask ? d.d : a.d;                                  // :235-237
  o.s = (don_cube && !acc_cube) ? 0.f : (mask ? 0.f : 1.f);  // :238-243
  return o;
}

__global__ void __launch_bounds__(256)
transplant_count_kernel(const float* __restrict__ images, const int* __restrict__ iparams, int B, int H, int W,
                        int* __restrict__ counts) {
  const int b = blockIdx.y;
  const int donor = iparams[b * kAugIP + 0];
  if (donor < 0) return;
  const long long plane = static_cast<long long>(H) * W;
  const float* a = images + static_cast<long long>(b) * 5 * plane;
  const float* d = images + static_cast<long long>(donor) * 5 * plane;
  int n = 0;
  for (long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x; i < plane;
       i += static_cast<long long>(gridDim.x) * 256) {
    Px5 pa, pd;
    pa.d = __ldg(a + 3 * plane + i);
    pa.s = __ldg(a + 4 * plane + i);
    pd.d = __ldg(d + 3 * plane + i);
    pd.s = __ldg(d + 4 * plane + i);
    pa.r = pa.g = pa.b = pd.r = pd.g = pd.b = 0.f;
    n += transplant(pa, pd).s == 1.0f ? 1 : 0;
  }
  n = __reduce_add_sync(0xffffffffu, n);
  if ((threadIdx.x & 31) == 0 && n != 0) atomicAdd(&counts[b], n);
}

// ---- colour operators (kornia.enhance definitions; kornia>=0.7.1 is the reference's dependency, pyproject.toml:20) ----
constexpr float kTwoPi = 6.283185307179586f;

// kornia.color.rgb_to_hsv: h in [0, 2pi), s = delta / (max + 1e-8), v = max; first maximum wins ties
__device__ __forceinline__ void rgb_to_hsv(float r, float g, float b, float& h, float& s, float& v) {
  const float mx = fmaxf(r, fmaxf(g, b)), mn = fminf(r, fminf(g, b));
  float delta = mx - mn;
  v = mx;
  s = __fdividef(delta, mx + 1e-8f);
  if (delta == 0.f) delta = 1.f;
  const float rc = mx - r, gc = mx - g, bc = mx - b;
  float hh;
  if (r == mx) hh = bc - gc;
  else if (g == mx) hh = (rc - bc) + 2.f * delta;
  else hh = (gc - rc) + 4.f * delta;
  hh = __fdividef(hh, delta) * (1.f / 6.f);
  hh = hh - floorf(hh);  // python-style % 1.0
  h = kTwoPi * hh;
}

// kornia.color.hsv_to_rgb
__device__ __forceinline__ void hsv_to_rgb(float h, float s, float v, float& r, float& g, float& b) {
  const float h6 = (h * (1.f / kTwoPi)) * 6.f;
  float hi = floorf(h6);
  hi = hi - 6.f * floorf(hi * (1.f / 6.f));        // floor(h*6) % 6  (hi is a small integer: exact)
  const float m6 = h6 - 6.f * floorf(h6 * (1.f / 6.f));  // (h*6) % 6
  const float f = m6 - hi;
  const float p = v * (1.f - s), q = v * (1.f - f * s), t = v * (1.f - (1.f - f) * s);
  switch (static_cast<int>(hi)) {
    case 0: r = v; g = t; b = p; break;
    case 1: r = q; g = v; b = p; break;
    case 2: r = p; g = v; b = t; break;
    case 3: r = p; g = q; b = v; break;
    case 4: r = t; g = p; b = v; break;
    default: r = v; g = p; b = q; break;
  }
}

__device__ __forceinline__ float clamp01(float x) { return fminf(fmaxf(x, 0.f), 1.f); }

// RandomPlanckianJitter gains + ColorJiggle's four operators in the sampled order
__device__ __forceinline__ void colour_ops(float& r, float& g, float& b, int flags, int order,
                                           const float* __restrict__ fp) {
  if (flags & PB200_AUG_PLANCKIAN) {  // r * r_w, b * b_w, clamp(max = 1)
    r = fminf(r * fp[12], 1.f);
    g = fminf(g, 1.f);
    b = fminf(b * fp[13], 1.f);
  }
  if (flags & PB200_AUG_JIGGLE) {
#pragma unroll 1
    for (int k = 0; k < 4; ++k) {
      const int op = (order >> (4 * k)) & 15;
      if (op == 0) {  // adjust_brightness(x, factor - 1): x + (factor - 1), clamp
        const float add = fp[14] - 1.f;
        r = clamp01(r + add); g = clamp01(g + add); b = clamp01(b + add);
      } else if (op == 1) {  // adjust_contrast(x, factor): x * factor, clamp
        r = clamp01(r * fp[15]); g = clamp01(g * fp[15]); b = clamp01(b * fp[15]);
      } else if (op == 2) {  // adjust_saturation
        float h, s, v;
        rgb_to_hsv(r, g, b, h, s, v);
        s = clamp01(s * fp[16]);
        hsv_to_rgb(h, s, v, r, g, b);
      } else {  // adjust_hue(x, factor * 2pi): h = fmod(h + shift, 2pi)
        float h, s, v;
        rgb_to_hsv(r, g, b, h, s, v);
        h = fmodf(h + fp[17], kTwoPi);
        hsv_to_rgb(h, s, v, r, g, b);
      }
    }
  }
}

// ---- depth operators (augmentations.py:17-169), fp32 in the reference's own operation order ---------------------
__device__ __forceinline__ float dropout_mask(float u, float p, float keep) {
  // F.dropout(ones, p, training): 0 with probability p, else keep = 1/(1-p); the host passes p = 0, keep = 1 when the
  // dropout is inactive and keep = 0 for p >= 1
  return u >= p ? keep : 0.f;
}
__device__ __forceinline__ float centred(float dev, float mask, float u) {  // dev * mask * 2 * (U - 0.5)
  return __fmul_rn(__fmul_rn(__fmul_rn(dev, mask), 2.f), __fsub_rn(u, 0.5f));
}

struct DropKeep {
  float p_bias, keep_bias, p_near, keep_near, p_far, keep_far;
};

__device__ __forceinline__ float depth_ops(float d, const pb200_aug_config& c, const DropKeep& K, unsigned long long seed,
                                           unsigned pix, unsigned b) {
  const uint2 key = make_uint2(static_cast<unsigned>(seed), static_cast<unsigned>(seed >> 32));
  const uint4 w0 = philox4x32_10(make_uint4(pix, b, 0u, 0u), key);
  const uint4 w1 = philox4x32_10(make_uint4(pix, b, 1u, 0u), key);
  if (c.depth_bias) {  // DepthBiasAugmentation.forward :42-50
    const float s = __fmul_rn(c.cube_scale, d);
    const float bias = centred(c.dev_bias, dropout_mask(u01(w0.x), K.p_bias, K.keep_bias), u01(w0.y));
    d = __fdiv_rn(__fadd_rn(s, bias), c.cube_scale);
  }
  if (c.depth_noise) {  // DepthGaussianNoiseAugmentation.forward :76-78 (its own cube_scale = 0.035)
    const float s = __fmul_rn(c.noise_cube_scale, d);
    const float u1 = 1.f - u01(w0.z);  // (0, 1]
    const float z = sqrtf(-2.f * __logf(u1)) * cospif(2.f * u01(w0.w));
    d = __fdiv_rn(__fadd_rn(s, __fmul_rn(c.std_noise, z)), c.noise_cube_scale);
  }
  if (c.near_plane || c.far_plane) {  // DepthPlaneAugmentation.forward :137-168
    float s = __fmul_rn(c.cube_scale, d);
    if (c.near_plane) {
      const float plane = __fadd_rn(c.near_mean, centred(c.near_dev, dropout_mask(u01(w1.x), K.p_near, K.keep_near), u01(w1.y)));
      if (s < plane) s = c.near_value;
    }
    if (c.far_plane) {
      const float plane = __fadd_rn(c.far_mean, centred(c.far_dev, dropout_mask(u01(w1.z), K.p_far, K.keep_far), u01(w1.w)));
      if (s > plane) s = c.far_value;
    }
    d = __fdiv_rn(s, c.cube_scale);
  }
  return d;
}

// ---- the fused kernel ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int reflect(int i, int n) {  // torch 'reflect' padding (no edge repeat), |overhang| <= 2 < n
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return i;
}

// the (transplanted) source image at integer pixel (y, x); zeros outside (grid_sample padding_mode='zeros')
__device__ __forceinline__ Px5 fetch(const float* __restrict__ acc, const float* __restrict__ don, long long plane, int C,
                                     int H, int W, int y, int x) {
  Px5 z = {0.f, 0.f, 0.f, 0.f, 0.f};
  if (y < 0 || y >= H || x < 0 || x >= W) return z;
  const long long o = static_cast<long long>(y) * W + x;
  const Px5 a = load5(acc + o, plane, C);
  if (don == nullptr) return a;
  return transplant(a, load5(don + o, plane, C));
}

__global__ void __launch_bounds__(kAugThreads) augment_kernel(AugArgs A) {
  __shared__ float s_px[5][kAugSpan][kAugSpan + 1];
  __shared__ float s_fp[kAugFP];
  __shared__ int s_ip[kAugIP];
  const int b = blockIdx.z;
  if (threadIdx.x < kAugFP) s_fp[threadIdx.x] = A.fparams[b * kAugFP + threadIdx.x];
  if (threadIdx.x < kAugIP) s_ip[threadIdx.x] = A.iparams[b * kAugIP + threadIdx.x];
  __syncthreads();
  const int H = A.H, W = A.W, C = A.C_in;
  const long long plane = static_cast<long long>(H) * W;
  const int flags = s_ip[10], order = s_ip[9];
  int donor = C == 5 ? s_ip[0] : -1;  // augmentations.py:201: transplantation needs the segmentation channel
  if (donor >= 0) {                   // :246-248: keep the transplant only if the new cube ratio is within bounds
    const float ratio = static_cast<float>(A.counts[b]) / static_cast<float>(plane);
    if (!(ratio >= A.cfg.lb_seg_ratio && ratio <= A.cfg.ub_seg_ratio)) donor = -1;
  }
  const float* acc = A.images + static_cast<long long>(b) * C * plane;
  const float* don = donor >= 0 ? A.images + static_cast<long long>(donor) * C * plane : nullptr;
  const int ty0 = blockIdx.y * kAugTile, tx0 = blockIdx.x * kAugTile;
  const bool blur = (flags & PB200_AUG_BLUR) != 0;
  const DropKeep dk = {A.drop[0], A.drop[1], A.drop[2], A.drop[3], A.drop[4], A.drop[5]};

  // phase 1: warp + erase + colour for the tile (and its halo when the sample is blurred)
  for (int t = threadIdx.x; t < kAugSpan * kAugSpan; t += kAugThreads) {
    const int ly = t / kAugSpan, lx = t - ly * kAugSpan;
    const bool interior = ly >= kAugHalo && ly < kAugHalo + kAugTile && lx >= kAugHalo && lx < kAugHalo + kAugTile;
    if (!blur && !interior) continue;
    int y = ty0 + ly - kAugHalo, x = tx0 + lx - kAugHalo;
    const int lim = blur ? kAugHalo : 0;  // a ragged edge tile: nothing beyond the image (+ its reflected halo) is read
    if (y >= H + lim || x >= W + lim) continue;
    y = reflect(y, H);
    x = reflect(x, W);
    Px5 p;
    bool erased = false;
#pragma unroll
    for (int e = 0; e < 2; ++e) {  // RandomErasing: rectangle [ey, ey+eh) x [ex, ex+ew) := 0, all channels
      const int ex = s_ip[1 + 4 * e], ey = s_ip[2 + 4 * e], ew = s_ip[3 + 4 * e], eh = s_ip[4 + 4 * e];
      erased |= (ew > 0 && x >= ex && x < ex + ew && y >= ey && y < ey + eh);
    }
    if (erased) {
      p = Px5{0.f, 0.f, 0.f, 0.f, 0.f};
    } else if (flags & PB200_AUG_AFFINE) {
      // sampling position of output pixel (x, y) in the source (kornia warp_affine -> affine_grid + grid_sample,
      // align_corners=False; the host folds kornia's normalisation convention into S)
      const float sx = fmaf(s_fp[6], static_cast<float>(x), fmaf(s_fp[7], static_cast<float>(y), s_fp[8]));
      const float sy = fmaf(s_fp[9], static_cast<float>(x), fmaf(s_fp[10], static_cast<float>(y), s_fp[11]));
      const float fx = floorf(sx), fy = floorf(sy);
      const float wx1 = sx - fx, wy1 = sy - fy, wx0 = 1.f - wx1, wy0 = 1.f - wy1;
      p = Px5{0.f, 0.f, 0.f, 0.f, 0.f};
      if (fx >= -1.f && fx <= static_cast<float>(W) && fy >= -1.f && fy <= static_cast<float>(H)) {
        const int x0 = static_cast<int>(fx), y0 = static_cast<int>(fy);
        const Px5 p00 = fetch(acc, don, plane, C, H, W, y0, x0), p01 = fetch(acc, don, plane, C, H, W, y0, x0 + 1);
        const Px5 p10 = fetch(acc, don, plane, C, H, W, y0 + 1, x0), p11 = fetch(acc, don, plane, C, H, W, y0 + 1, x0 + 1);
        const float w00 = wx0 * wy0, w01 = wx1 * wy0, w10 = wx0 * wy1, w11 = wx1 * wy1;
        p.r = p00.r * w00 + p01.r * w01 + p10.r * w10 + p11.r * w11;
        p.g = p00.g * w00 + p01.g * w01 + p10.g * w10 + p11.g * w11;
        p.b = p00.b * w00 + p01.b * w01 + p10.b * w10 + p11.b * w11;
        p.d = p00.d * w00 + p01.d * w01 + p10.d * w10 + p11.d * w11;
        p.s = p00.s * w00 + p01.s * w01 + p10.s * w10 + p11.s * w11;
      }
    } else {
      p = fetch(acc, don, plane, C, H, W, y, x);
    }
    colour_ops(p.r, p.g, p.b, flags, order, s_fp);
    s_px[0][ly][lx] = p.r;
    s_px[1][ly][lx] = p.g;
    s_px[2][ly][lx] = p.b;
    s_px[3][ly][lx] = p.d;
    s_px[4][ly][lx] = p.s;
  }
  __syncthreads();

  // phase 2: blur + shadow + depth operators, planar stores (a warp writes 32 consecutive floats per plane)
  for (int t = threadIdx.x; t < kAugTile * kAugTile; t += kAugThreads) {
    const int ly = t / kAugTile, lx = t - ly * kAugTile;
    const int y = ty0 + ly, x = tx0 + lx;
    if (y >= H || x >= W) continue;
    float rgb[3];
    if (blur) {  // kornia gaussian_blur2d (5, 5), separable, border 'reflect': horizontal pass, then vertical
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float accv = 0.f;
#pragma unroll
        for (int dy = 0; dy < 5; ++dy) {
          float row = 0.f;
#pragma unroll
          for (int dx = 0; dx < 5; ++dx) row = fmaf(s_fp[18 + dx], s_px[c][ly + dy][lx + dx], row);
          accv = fmaf(s_fp[18 + dy], row, accv);
        }
        rgb[c] = accv;
      }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) rgb[c] = s_px[c][ly + kAugHalo][lx + kAugHalo];
    }
    const long long o = static_cast<long long>(y) * W + x;
    if ((flags & PB200_AUG_SHADOW) && A.plasma != nullptr) {  // RandomPlasmaShadow: (plasma < quantity) * intensity
      const float shade = __ldg(A.plasma + static_cast<long long>(b) * plane + o) < s_fp[24] ? s_fp[23] : 0.f;
#pragma unroll
      for (int c = 0; c < 3; ++c) rgb[c] = clamp01(rgb[c] + shade);
    }
    float* out = A.out + static_cast<long long>(b) * A.C_out * plane + o;
#pragma unroll
    for (int c = 0; c < 3; ++c) out[c * plane] = rgb[c];
    if (A.C_out > 3) {
      const float d = s_px[3][ly + kAugHalo][lx + kAugHalo];
      out[3 * plane] = depth_ops(d, A.cfg, dk, A.seed, static_cast<unsigned>(o), static_cast<unsigned>(b));
    }
    if (A.C_out > 4) out[4 * plane] = s_px[4][ly + kAugHalo][lx + kAugHalo];
  }
}

// key points: RandomAffine maps them by the forward matrix; normalize_pixel_coordinates (augmentations.py:481)
__global__ void aug_keypoints_kernel(const float* __restrict__ kp, const int* __restrict__ iparams,
                                     const float* __restrict__ fparams, int B, int K, int H, int W,
                                     float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * K) return;
  const int b = i / K;
  float u = kp[2 * i], v = kp[2 * i + 1];
  if (iparams[b * kAugIP + 10] & PB200_AUG_AFFINE) {
    const float* M = fparams + b * kAugFP;
    const float nu = fmaf(M[0], u, fmaf(M[1], v, M[2]));
    const float nv = fmaf(M[3], u, fmaf(M[4], v, M[5]));
    u = nu;
    v = nv;
  }
  // kornia: factor = 2 / (size - 1).clamp(eps); out = factor * x - 1
  const float fw = 2.f / fmaxf(static_cast<float>(W - 1), 1e-8f), fh = 2.f / fmaxf(static_cast<float>(H - 1), 1e-8f);
  out[2 * i] = __fsub_rn(__fmul_rn(fw, u), 1.f);
  out[2 * i + 1] = __fsub_rn(__fmul_rn(fh, v), 1.f);
}

}  // namespace pb200

using namespace pb200;

extern "C" int pb200_augment(const float* images, const int32_t* iparams, const float* fparams,
                             const float* plasma_or_null, const pb200_aug_config* cfg, uint64_t seed, int32_t* counts,
                             const float* keypoints_or_null, float* out_images, float* out_keypoints_or_null, int B,
                             int C_in, int C_out, int H, int W, int K, pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(images && iparams && fparams && cfg && counts && out_images, "augment: null argument");
  PB_CHECK(B >= 0 && C_in >= 3 && C_in <= 5 && C_out >= 3 && C_out <= C_in, "augment: channels in %d out %d", C_in, C_out);
  PB_CHECK(H >= 3 && W >= 3 && static_cast<long long>(H) * W < (1LL << 31), "augment: H x W = %d x %d unsupported", H, W);
  PB_CHECK(images != out_images, "augment: in-place operation is not supported (gather)");
  if (B == 0) return PB200_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PB_CUDA(cudaMemsetAsync(counts, 0, sizeof(int32_t) * B, st));
  if (C_in == 5) {
    dim3 g(static_cast<unsigned>((static_cast<long long>(H) * W + 256 * 8 - 1) / (256 * 8)), B);
    transplant_count_kernel<<<g, 256, 0, st>>>(images, iparams, B, H, W, counts);
    PB_CUDA(cudaGetLastError());
  }
  AugArgs A;
  A.images = images; A.iparams = iparams; A.fparams = fparams; A.counts = counts; A.plasma = plasma_or_null;
  const float ps[3] = {cfg->p_bias, cfg->p_near, cfg->p_far};
  for (int i = 0; i < 3; ++i) {  // F.dropout(ones, p, training): keep = 1/(1-p) in fp32
    const bool active = cfg->training && ps[i] > 0.f;
    A.drop[2 * i] = active ? ps[i] : 0.f;
    A.drop[2 * i + 1] = !active ? 1.f : (ps[i] >= 1.f ? 0.f : 1.f / (1.f - ps[i]));
    if (active && ps[i] >= 1.f) A.drop[2 * i] = 2.f;  // u01 < 1 always: every element dropped
  }
  A.out = out_images; A.B = B; A.C_in = C_in; A.C_out = C_out; A.H = H; A.W = W; A.seed = seed; A.cfg = *cfg;
  dim3 grid((W + kAugTile - 1) / kAugTile, (H + kAugTile - 1) / kAugTile, B);
  augment_kernel<<<grid, kAugThreads, 0, st>>>(A);
  PB_CUDA(cudaGetLastError());
  if (keypoints_or_null != nullptr && out_keypoints_or_null != nullptr && K > 0) {
    aug_keypoints_kernel<<<(B * K + 255) / 256, 256, 0, st>>>(keypoints_or_null, iparams, fparams, B, K, H, W,
                                                             out_keypoints_or_null);
    PB_CUDA(cudaGetLastError());
  }
  return PB200_OK;
}

// ---- diamond-square plasma fractal (RandomPlasmaShadow's shade map) -------------------------------------------
// One CTA per image, every level in one launch: the S x S grid (S = 2^k + 1) lives in global scratch (264 KB at
// 257^2: L2-resident), __syncthreads() orders the levels.  Noise of grid point (y, x): Philox(seed; y*S+x, image, 2).
namespace pb200 {

__device__ __forceinline__ float plasma_noise(unsigned long long seed, unsigned pt, unsigned b) {
  const uint4 w = philox4x32_10(make_uint4(pt, b, 2u, 0u), make_uint2(static_cast<unsigned>(seed), static_cast<unsigned>(seed >> 32)));
  return __fsub_rn(__fmul_rn(2.f, u01(w.x)), 1.f);  // [-1, 1)
}

__global__ void __launch_bounds__(1024)
plasma_kernel(const float* __restrict__ roughness, unsigned long long seed, float* scratch, float* __restrict__ out,
              int S, int H, int W) {
  __shared__ float s_red[2][32];
  const unsigned b = blockIdx.x;
  float* g = scratch + static_cast<long long>(b) * S * S;
  const float rough = roughness[b];
  const int st0 = (S - 1) / 2;
  if (threadIdx.x < 9) {
    const int y = (threadIdx.x / 3) * st0, x = (threadIdx.x % 3) * st0;
    g[y * S + x] = plasma_noise(seed, y * S + x, b);
  }
  __syncthreads();
  float amp = 1.f;
  for (int st = st0; st >= 2; st >>= 1) {
    const int half = st >> 1;
    amp = __fmul_rn(amp, rough);
    const int n = (S - 1) / st;  // cells per side
    // diamond: cell centres
    for (int t = threadIdx.x; t < n * n; t += blockDim.x) {
      const int y = (t / n) * st + half, x = (t % n) * st + half;
      const float m = __fmul_rn(0.25f, __fadd_rn(__fadd_rn(g[(y - half) * S + x - half], g[(y - half) * S + x + half]),
                                                 __fadd_rn(g[(y + half) * S + x - half], g[(y + half) * S + x + half])));
      g[y * S + x] = __fadd_rn(m, __fmul_rn(amp, plasma_noise(seed, y * S + x, b)));
    }
    __syncthreads();
    // square: edge midpoints — (row multiple of st, column odd multiple of half) and the transpose
    const int per = n * (n + 1);
    for (int t = threadIdx.x; t < 2 * per; t += blockDim.x) {
      int y, x;
      if (t < per) {
        y = (t / n) * st;
        x = (t % n) * st + half;
      } else {
        const int u = t - per;
        x = (u / n) * st;
        y = (u % n) * st + half;
      }
      float sum = 0.f;
      int cnt = 0;
      if (y - half >= 0) { sum = __fadd_rn(sum, g[(y - half) * S + x]); ++cnt; }
      if (x - half >= 0) { sum = __fadd_rn(sum, g[y * S + x - half]); ++cnt; }
      if (x + half < S) { sum = __fadd_rn(sum, g[y * S + x + half]); ++cnt; }
      if (y + half < S) { sum = __fadd_rn(sum, g[(y + half) * S + x]); ++cnt; }
      g[y * S + x] = __fadd_rn(__fdiv_rn(sum, static_cast<float>(cnt)), __fmul_rn(amp, plasma_noise(seed, y * S + x, b)));
    }
    __syncthreads();
  }
  // min-max normalisation over the H x W crop
  float mn = INFINITY, mx = -INFINITY;
  for (int t = threadIdx.x; t < H * W; t += blockDim.x) {
    const float v = g[(t / W) * S + (t % W)];
    mn = fminf(mn, v);
    mx = fmaxf(mx, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if ((threadIdx.x & 31) == 0) {
    s_red[0][threadIdx.x >> 5] = mn;
    s_red[1][threadIdx.x >> 5] = mx;
  }
  __syncthreads();
  mn = s_red[0][0];
  mx = s_red[1][0];
  for (int i = 1; i < static_cast<int>(blockDim.x >> 5); ++i) {
    mn = fminf(mn, s_red[0][i]);
    mx = fmaxf(mx, s_red[1][i]);
  }
  const float range = __fsub_rn(mx, mn);
  float* o = out + static_cast<long long>(b) * H * W;
  for (int t = threadIdx.x; t < H * W; t += blockDim.x) {
    const float v = g[(t / W) * S + (t % W)];
    o[t] = range > 0.f ? __fdiv_rn(__fsub_rn(v, mn), range) : 0.f;
  }
}

}  // namespace pb200

extern "C" int pb200_plasma_side(int H, int W) {
  int m = H > W ? H : W;
  int s = 2;
  while (s + 1 < m) s <<= 1;
  return s + 1;
}

extern "C" int pb200_plasma(const float* roughness, uint64_t seed, float* scratch, float* out, int B, int H, int W,
                            pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(roughness && scratch && out && B >= 0 && H >= 1 && W >= 1, "plasma: bad argument");
  if (B == 0) return PB200_OK;
  const int S = pb200_plasma_side(H, W);
  PB_CHECK(static_cast<long long>(S) * S < (1LL << 31), "plasma: %d x %d too large", H, W);
  plasma_kernel<<<B, 1024, 0, static_cast<cudaStream_t>(stream)>>>(roughness, seed, scratch, out, S, H, W);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}

// ---- the same fractal with NO caller scratch and NO allocation (no counterpart in the reference: added
// for callers that cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops) ------
// The S x S working grids live in a file-scope device array of PB200_PLASMA_STATIC_MAX_FLOATS floats (4.03 MB = sixteen
// 257 x 257 grids) that every GPU that loads the library holds until exit.  In host code `g_plasma_static` is only a
// stand-in for the symbol, NOT a device address: the address is fetched once per device, with that device current, by
// cudaGetSymbolAddress and passed to the kernel.  Every call shares the array: calls that overlap on the GPU (two streams,
// two host threads) race on it; calls on one stream are ordered.  Nothing has to start at zero -- the kernel writes every
// grid point (corners / centres / edge midpoints of every level) before any read of it -- so the previous call's contents
// are simply overwritten.
namespace pb200 {
static __device__ __align__(256) float g_plasma_static[PB200_PLASMA_STATIC_MAX_FLOATS];
}  // namespace pb200

extern "C" int pb200_plasma_static(const float* roughness, uint64_t seed, float* out, int B, int H, int W,
                                   pb200_stream stream) {
  if (int rc = require_device()) return rc;
  PB_CHECK(roughness && out && B >= 0 && H >= 1 && W >= 1, "plasma_static: bad argument");
  const int S = pb200_plasma_side(H, W);
  const long long need = static_cast<long long>(B) * S * S;
  PB_CHECK(static_cast<long long>(S) * S < (1LL << 31) && need <= static_cast<long long>(PB200_PLASMA_STATIC_MAX_FLOATS),
           "plasma_static: %d images of %d x %d need %lld scratch floats, the built-in scratch holds %lld "
           "(PB200_PLASMA_STATIC_MAX_FLOATS); use pb200_plasma with caller scratch for larger problems",
           B, H, W, need, static_cast<long long>(PB200_PLASMA_STATIC_MAX_FLOATS));
  if (B == 0) return PB200_OK;
  static float* dev_ptr[kMaxDevices] = {};  // per device: the symbol's address in THAT device's copy of the module
  const int dev = current_device();
  if (dev_ptr[dev] == nullptr) {
    float* ptr = nullptr;
    PB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_plasma_static));
    PB_CHECK(ptr != nullptr, "plasma_static: cudaGetSymbolAddress returned a null address on device %d", dev);
    dev_ptr[dev] = ptr;
  }
  plasma_kernel<<<B, 1024, 0, static_cast<cudaStream_t>(stream)>>>(roughness, seed, dev_ptr[dev], out, S, H, W);
  PB_CUDA(cudaGetLastError());
  return PB200_OK;
}
