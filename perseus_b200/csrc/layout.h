// Buffer layouts and internal launch interfaces shared by the kernels and the plan.
#pragma once
#include <cuda_runtime.h>
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
  int rgb_integer;  // 1: emit 0..255 (bf16 modes), 0: emit k/255 (fp32 mode)
};

int launch_ingest(const IngestArgs& a, void* out, bool bf16, cudaStream_t st);
int launch_repack_nchw(const float* x, int C, int H, int W, int B, void* out, bool bf16,
                       cudaStream_t st);

// 3x3 stride-2 pad-1 max pooling, NHWC (C % 8 == 0 for bf16, % 4 for fp32).
int launch_maxpool(const void* in, void* out, int B, int H, int W, int C, bool bf16, cudaStream_t st);
// global average pool + Linear(C -> n_out) + pixel decode; in is NHWC [B][HW][C].
int launch_head(const void* in, const float* fc_w, const float* fc_b, float* out_norm, float* out_px,
                int B, int HW, int C, int n_out, int imgH, int imgW, bool bf16, cudaStream_t st);
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

}  // namespace pb200
