// Buffer layouts and internal launch interfaces shared by the kernels and the plan.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace pb200 {

// frames layout: [B][H + 6][W + 8][4], zero margins.  The left margin is 4 (not the conv's 3) so
// that every 4-pixel store group is either all margin or all image and the stem's TMA rows stay
// 16-byte aligned; the stem kernel compensates with a zero weight for its first (8th) tap.
constexpr int kFrameMarginTop = 3;
constexpr int kFrameMarginBottom = 3;
constexpr int kFrameMarginLeft = 4;
constexpr int kFrameMarginRight = 4;
constexpr int kFrameC = 4;
// The stem is evaluated as a 7 (rows) x 8 (taps) x 4 (channels) filter over the frames buffer:
// K index = r*32 + s8*4 + c with s8 = s + 1 (tap 0 has zero weight).
constexpr int kStemTaps = 8;
constexpr int kStemK = 7 * kStemTaps * kFrameC;  // 224

struct IngestArgs {
  const uint8_t* rgb;
  long long pitch;        // bytes per source row
  long long frame_pitch;  // bytes per source frame
  int ch;                 // 3 or 4
  int bgr;
  int src_h, src_w;
  int y0, x0;  // crop origin
  const float* depth;
  float depth_divisor;
  int clip;
  float cube_scale, near_plane, far_plane;
  const float* jitter;  // [B][2][H][W] or null
  int H, W, B;
  int rgb_integer;  // 1: emit 0..255 (16-bit modes), 0: emit k/255 (fp32 mode)
};

// element type of a frames buffer
enum FrameElt { FRAME_F32 = 0, FRAME_BF16 = 1, FRAME_F16 = 2 };
// fp16 frames ("fp16x2" mode) carry RGB as k * 2^-12 and depth as d * 2^-4 (exact power-of-two scalings that keep the
// stem's folded weights -- w * 4096/255 and w * 16 -- and their lo parts in fp16's normal range)
constexpr float kF16RgbScale = 1.0f / 4096.0f;
constexpr float kF16DepthScale = 1.0f / 16.0f;

int launch_ingest(const IngestArgs& a, void* out, int elt /* FrameElt */, cudaStream_t st);
int launch_repack_nchw(const float* x, int C, int H, int W, int B, void* out, int elt /* FrameElt */,
                       cudaStream_t st);


// ---- "P8" activation layout: [n][C/8][H+2][W+2][8] bf16, zero border --------------
// One 16-byte element = 8 consecutive channels of one pixel; a channel-chunk plane is a dense
// (H+2) x (W+2) image.  A 3x3/1 conv over the FLATTENED padded plane is out[q] = sum_{r,s} in[q + r*(W+2) + s] . w[r][s]
// with q = h*(W+2) + w, so every filter tap is the same shared-memory patch read at a 16-byte-granular
// offset through a no-swizzle K-major UMMA descriptor (row pitch 16 B, LBO = plane pitch).
// Producers write the whole tensor (interior + zero border) every time.
inline size_t p8_plane_elems(int H, int W) { return static_cast<size_t>(H + 2) * (W + 2); }
inline size_t p8_image_bytes(int C, int H, int W) { return static_cast<size_t>(C / 8) * p8_plane_elems(H, W) * 16; }
constexpr size_t kP8SlackBytes = 4096;  // the last tile of the last image reads 2 elements past the end

// Fused stem: conv1 7x7/2 + folded BN + ReLU + maxpool 3x3/2 on the frames buffer -> P8 [B][8][66][66][8].
// wimg: 28672-byte shared-memory image of the folded stem weights (no-swizzle K-major core-matrix
// order, K = r*32 + s8*4 + c): element (co, k) at (co/8)*3584 + (k/8)*128 + (co%8)*16 + (k%8)*2.
constexpr int kStemWimgBytes = 64 * kStemK * 2;
inline size_t stem_wimg_offset(int co, int k) {
  return static_cast<size_t>(co / 8) * (kStemK / 8 * 128) + static_cast<size_t>(k / 8) * 128 + (co % 8) * 16 + (k % 8) * 2;
}
int stem_pool_init();

// conv_p8.cu: 3x3/1 conv 64 -> 64 on P8 with the weights resident in shared memory.
// wimg: 73728-byte image, K = (r*3 + s)*64 + ci: element (co, k) at (co/8)*9216 + (k/8)*128 + (co%8)*16 + (k%8)*2.
inline size_t conv_p8_wimg_offset(int co, int k) {
  return static_cast<size_t>(co / 8) * 9216 + static_cast<size_t>(k / 8) * 128 + (co % 8) * 16 + (k % 8) * 2;
}
int conv_p8_init();
// x2 = 1: the "fp16x2" mode -- fp16 P8 activations, wimg = the hi image followed by the lo image (2 x 73728 bytes)
int launch_conv_p8(const void* in_p8, const void* wimg, const float* bias, const void* residual_p8, void* out,
                   int out_nhwc, int relu, int B, int H, int W, cudaStream_t st, int x2 = 0, int reverse = 0);
// x2 = 1: the "fp16x2" mode -- fp16 frames / output, wimg = the hi image followed by the lo image (2 x 28672 bytes)
// rgb8 != null: the ingest step fused into the stem -- it reads the tight, 16-byte aligned uint8 [B][H][256][3] batch
// itself (bgr: source pixels are B, G, R) and `frames` is not touched
int launch_stem_pool(const void* frames, const void* wimg, const float* bias, void* out_p8, int B, int H,
                     int W, cudaStream_t st, int x2 = 0, const void* rgb8 = nullptr, int bgr = 0, int reverse = 0);

// 3x3 stride-2 pad-1 max pooling, NHWC (C % 8 == 0 for bf16, % 4 for fp32).
int launch_maxpool(const void* in, void* out, int B, int H, int W, int C, bool bf16, cudaStream_t st);
// global average pool + Linear(C -> n_out) + pixel decode; in is NHWC [B][HW][C].
// mode: 0 = fp32 input, 1 = bf16, 2 = split bf16 rows [hi(C) | lo(C)], 3 = fp16
int launch_head(const void* in, const float* fc_w, const float* fc_b, float* out_norm, float* out_px,
                int B, int HW, int C, int n_out, int imgH, int imgW, int mode, cudaStream_t st);
// split-bf16 mode: fp32 NHWC in -> [hi(C) | lo(C)] bf16 NHWC out
int launch_maxpool_split(const void* in, void* out, int B, int H, int W, int C, cudaStream_t st);
int launch_decode(const float* xy_norm, float* xy_px, long long n_points, int H, int W,
                  cudaStream_t st);

// fp32 SIMT implicit-GEMM convolution, NHWC, weights [K][Cout] with K = (r, s, c).
struct SimtConv {
  const float* in;   // [B][Hin][Win][Cin]
  const float* w;    // [R*S*Cin][Cout]
  const float* bias; // [Cout]
  const float* residual;  // [B][Ho][Wo][Cout] or null
  float* out;
  int B, Hin, Win, Cin, Ho, Wo, Cout, R, S, stride, pad, relu;
};
int launch_conv_simt(const SimtConv& c, cudaStream_t st);
// dedicated fp32 stem over the frames buffer (fp32 and split-bf16 modes); real_channels = the model's 3 or 4
bool stem_f32_supported(const SimtConv& c);
int launch_stem_f32(const SimtConv& c, int real_channels, cudaStream_t st);

}  // namespace pb200
