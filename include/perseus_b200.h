/*
 * perseus_b200.h — C ABI of the B200-native perseus hot path (libperseus_b200.so).
 *
 * This is the drop-in boundary: everything the Python host side (perseus_b200/) does on the
 * GPU goes through these entry points.  There are no torch types in the signatures — plain
 * pointers, sizes and a cudaStream_t passed as void*.
 *
 * Conventions
 *   - every function returns int: 0 = OK, non-zero = error; pb200_last_error() returns a
 *     thread-local message for the last failing call on this thread.
 *   - unless stated otherwise pointers are DEVICE pointers owned by the caller; calls are
 *     asynchronous on `stream` and do no allocation after pb200_plan_create().
 *   - there is no CPU fallback anywhere behind this ABI: with no usable sm_100 device the
 *     compute entry points fail with PB200_ERR_NO_DEVICE.
 *
 * Reference interfaces replaced (paths relative to the reference checkout):
 *   pb200_plan_create / _load_weights / _destroy
 *        <- KeypointCNN.__init__               perseus/detector/models.py:9-32
 *        <- checkpoint load                    perseus/detector/validate.py:92-98,
 *                                              scripts/streaming.py:112-119
 *   pb200_ingest
 *        <- ZEDCamera.get_frame math           scripts/streaming.py:68-80
 *        <- dataset u8 -> f32 conversion       perseus/detector/data.py:78,84-85
 *        <- RGBD concat                        perseus/detector/validate.py:122-124,
 *                                              perseus/detector/train.py:266-268
 *        <- val-mode depth near/far clip       perseus/detector/augmentations.py:137-168
 *   pb200_forward / pb200_forward_nchw_f32
 *        <- KeypointCNN.forward                perseus/detector/models.py:34-40
 *           (torchvision ResNet._forward_impl, BasicBlock.forward)
 *        <- kornia denormalize_pixel_coordinates call sites
 *                                              perseus/detector/validate.py:144-153,
 *                                              scripts/streaming.py:129-131
 *   pb200_factors_projection  <- KeypointProjectionFactor.error_func  perseus/smoother/factors.py:216-275
 *   pb200_factors_dynamics    <- PoseDynamicsFactor.error_func        perseus/smoother/factors.py:54-142
 *   pb200_factors_constvel    <- ConstantVelocityFactor.error_func    perseus/smoother/factors.py:160-171
 */
#ifndef PERSEUS_B200_H_
#define PERSEUS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PB200_VERSION 100 /* 0.1.0 */

/* error codes */
#define PB200_OK 0
#define PB200_ERR_INVALID 1     /* bad argument / unsupported shape */
#define PB200_ERR_CUDA 2        /* a CUDA runtime / driver call failed */
#define PB200_ERR_NO_DEVICE 3   /* no sm_100 device: there is no CPU fallback */
#define PB200_ERR_STATE 4       /* e.g. forward before load_weights */

/* arithmetic of the conv backbone (the "precision ladder") */
#define PB200_PREC_FP32 0   /* SIMT fp32 FFMA implicit GEMM: the <=1e-3 px parity mode          */
#define PB200_PREC_BF16 1   /* tcgen05 kind::f16 bf16 x bf16 -> fp32 TMEM: the throughput mode   */
#define PB200_PREC_BF16X3 2 /* split-bf16 (hi*hi + lo*hi + hi*lo) on tcgen05: tensor parity mode */
#define PB200_PREC_FP16X2 3 /* fp16 activations x (fp16 hi + fp16 lo) weights on tcgen05, two MMAs per K step
                             * on one activation tile: the tensor mode that meets 0.05 px (RGB) at speed */

typedef struct pb200_plan pb200_plan;
typedef void* pb200_stream; /* cudaStream_t */

int pb200_version(void);
const char* pb200_last_error(void);

/* ---- detector plan -------------------------------------------------------------------- */

/* Creates the plan for KeypointCNN(n_keypoints, in_channels, H, W) on `device`:
 * allocates packed-weight storage and fixes every buffer offset for up to max_batch frames. */
int pb200_plan_create(pb200_plan** out, int device, int max_batch, int in_channels, int H, int W,
                      int n_keypoints, int precision);
void pb200_plan_destroy(pb200_plan* plan);

/* Bytes of caller-owned device scratch pb200_forward* needs (for max_batch frames). */
int pb200_plan_workspace_bytes(const pb200_plan* plan, size_t* bytes);
/* Bytes of one ingested-frames buffer holding B frames in the plan's internal layout
 * (zero-margined NHWC4: bf16 for the tcgen05 modes, fp32 for PB200_PREC_FP32). */
int pb200_plan_frames_bytes(const pb200_plan* plan, int B, size_t* bytes);

/* Consumes the reference state_dict: n HOST fp32 tensors in the reference layout (OIHW conv
 * weights, BN weight/bias/running_mean/running_var, fc weight/bias), names without the
 * "module." prefix ("resnet.conv1.weight", ...).  Folds eval-mode BatchNorm (eps 1e-5) into
 * the conv weights and a per-channel bias, packs to the kernel layouts and uploads on `stream`
 * (synchronises that stream before returning).  Unknown names (num_batches_tracked) are ignored;
 * a missing tensor is PB200_ERR_INVALID. */
int pb200_plan_load_weights(pb200_plan* plan, const float* const* tensors,
                            const int64_t* const* shapes, const int* ndims,
                            const char* const* names, int n, pb200_stream stream);

/* Fused frame ingest -> plan-format frames buffer.
 *   rgb         : uint8 [B][src_h][src_w][rgb_channels] (3 = RGB/BGR, 4 = RGBA/BGRA), row pitch
 *                 rgb_pitch_bytes (>= src_w*rgb_channels), frame pitch rgb_pitch_bytes*src_h
 *   bgr         : non-zero = channel order is B,G,R(,A)  (ZED camera; streaming.py:68-69)
 *   depth       : float [B][src_h][src_w] or NULL (NULL -> depth channel written as 0)
 *   depth_divisor : after NaN/+-Inf -> 0 the depth is DIVIDED by this in fp32, exactly as
 *                 streaming.py:74-76 does (0.035 for a camera in metres, 1 for dataset units)
 *   clip        : non-zero = val-mode DepthPlaneAugmentation (augmentations.py:137-168):
 *                 s = cube_scale*d; s = 0 where s < near_plane + jn; s = 0 where s > far_plane + jf;
 *                 d = s / cube_scale.  jitter_or_null is float [B][2][H][W] holding (jn, jf) —
 *                 the reference draws them at random even in val mode;
 *                 NULL = the deterministic planes.
 *   The centred H x W window of the source is taken (streaming.py:79-80):
 *   rows src_h/2 - H/2 .., cols src_w/2 - W/2 .. */
int pb200_ingest(const pb200_plan* plan, const uint8_t* rgb, int rgb_channels,
                 int64_t rgb_pitch_bytes, int src_h, int src_w, int bgr, const float* depth,
                 float depth_divisor, int clip, float cube_scale, float near_plane, float far_plane,
                 const float* jitter_or_null, void* out_frames, int B, pb200_stream stream);

/* Backbone + head + decode on ingested frames.
 *   out_norm : float [B][2*n_keypoints] normalised coordinates (the reference forward() output)
 *   out_px   : float [B][n_keypoints][2] pixel coordinates, or NULL */
int pb200_forward(pb200_plan* plan, const void* frames, void* workspace, float* out_norm,
                  float* out_px_or_null, int B, pb200_stream stream);

/* pb200_ingest + pb200_forward in one call for the batched RGB workload (perseus/detector/validate.py:109-153 with
 * data.py:78,84-85: uint8 image batch -> model -> pixel keypoints): rgb is a TIGHT, 16-byte aligned uint8
 * [B][H][W][3] device batch (bgr = 1: pixels are B, G, R, scripts/streaming.py:69), plan is a bf16 / fp16x2 plan with
 * 3 input channels.  The stem kernel converts the rows itself, so the frames tensor is neither written nor read and
 * the ingest launch disappears; results are bit-identical to pb200_ingest (no depth, no crop) + pb200_forward. */
int pb200_forward_rgb8(pb200_plan* plan, const uint8_t* rgb, int bgr, void* workspace, float* out_norm,
                       float* out_px_or_null, int B, pb200_stream stream);

/* Compatibility entry used by KeypointCNN.forward(x): x is float [B][C][H][W] (RGB in [0,1],
 * depth in data units); fuses the NCHW->frames repack in front of pb200_forward. */
int pb200_forward_nchw_f32(pb200_plan* plan, const float* x_nchw, void* workspace, float* out_norm,
                           float* out_px_or_null, int B, pb200_stream stream);

/* Per-kernel timing of one forward (CUDA events around every launch; synchronises).
 * ms has room for pb200_plan_num_ops() floats. */
int pb200_plan_num_ops(const pb200_plan* plan);
const char* pb200_plan_op_name(const pb200_plan* plan, int i);
/* name of the kernel op i launches (depends on the batch size of the last forward for the tcgen05 convs) */
const char* pb200_plan_op_kernel(const pb200_plan* plan, int i);
/* algorithmic FLOPs (2*MAC, unpadded) and algorithmic HBM bytes of op i for B frames */
int pb200_plan_op_work(const pb200_plan* plan, int i, int B, double* flops, double* bytes);
int pb200_forward_profile(pb200_plan* plan, const void* frames, void* workspace, float* out_norm,
                          float* out_px_or_null, int B, pb200_stream stream, float* ms);
/* number of kernel launches one pb200_forward issues (batch sizes up to the L2 chunk of the front layers) */
int pb200_plan_launches_per_forward(const pb200_plan* plan);
/* the same at batch size B: the fused stem + layer1 kernels run once per L2 chunk of frames */
int pb200_plan_launches_at(const pb200_plan* plan, int B);
/* ops launched together starting at op i: 1 = its own launch; k >= 2 = head of ONE chained persistent launch covering ops
 * i..i+k-1 (layers 3-4: torchvision resnet.py layer3 / layer4 as called from perseus/detector/models.py:34-40);
 * 0 = covered by a chain that started earlier.  pb200_forward_profile reports a chain's time at its head op. */
int pb200_plan_op_chain_len(const pb200_plan* plan, int i);

/* ---- stand-alone pieces (also used by the parity tests) ------------------------------- */

/* (x+1)*(W-1)/2, (y+1)*(H-1)/2 on n (x,y) pairs — kornia denormalize_pixel_coordinates. */
int pb200_decode_pixels(const float* xy_norm, float* xy_px, int64_t n_points, int H, int W,
                        pb200_stream stream);

/* One conv layer through the tcgen05 path, for unit tests: NHWC bf16 in/out, OIHW fp32 HOST
 * weights and bias (packed internally; allocates and synchronises — not a product entry). */
int pb200_debug_conv_bf16(const void* in_nhwc, int B, int H, int W, int Cin, const float* w_oihw_host,
                          const float* bias_host, int Cout, int R, int stride, int pad,
                          const void* residual_nhwc_or_null, int relu, void* out_nhwc,
                          pb200_stream stream);

/* Same for the split-bf16 ("bf16x3") mode: activations, residual and output are NHWC bf16 with 2*C
 * channels per pixel, [hi(C) | lo(C)], value = hi + lo. */
/* Unit-test entry of the fused projection (K-extension): out = [relu](conv3x3/1(in) + bias +
 * conv1x1/2(ds_in)) with ds_in bf16 [B][2H][2W][ds_Cin]; wds_oi_host fp32 [Cout][ds_Cin]; HOST weights. */
int pb200_debug_conv_ds_bf16(const void* in_nhwc, int B, int H, int W, int Cin, const float* w_oihw_host,
                             const float* bias_host, int Cout, const void* ds_in_nhwc, int ds_Cin,
                             const float* wds_oi_host, int relu, void* out_nhwc, pb200_stream stream);
int pb200_debug_conv_split(const void* in_nhwc2, int B, int H, int W, int Cin, const float* w_oihw_host,
                           const float* bias_host, int Cout, int R, int stride, int pad,
                           const void* residual_nhwc2_or_null, int relu, void* out_nhwc2,
                           pb200_stream stream);

/* Unit-test entries of the P8-layout kernels (allocate / synchronise — not product entries).
 * P8 = bf16 [n][C/8][H+2][W+2][8] with a zero border.
 *   pb200_debug_stem_pool : fused conv1+bn1+relu+maxpool of a loaded bf16 plan: frames -> P8 [B][8][H/4+2][W/4+2][8]
 *   pb200_debug_conv_p8   : one 3x3/1 64->64 conv (+bias, +residual, ReLU) P8 -> P8 or NHWC; OIHW fp32 HOST weights */
int pb200_debug_stem_pool(pb200_plan* plan, const void* frames, void* out_p8, int B, pb200_stream stream);
int pb200_debug_conv_p8(const void* in_p8, int B, int H, int W, const float* w_oihw_host,
                        const float* bias_host, const void* residual_p8_or_null, int relu, void* out,
                        int out_nhwc, pb200_stream stream);
/* The same two layer entries in the "fp16x2" mode (PB200_PREC_FP16X2): activations, residual and output are fp16
 * (NHWC / P8), the fp32 HOST weights are split into fp16 hi + fp16 lo internally.  pb200_debug_conv_f16x2 optionally
 * accumulates the 1x1/2 projection of ds_in [B][2H][2W][ds_Cin] (weights wds [Cout][ds_Cin]) into the same tile
 * (R = 3, stride 1 only); pass NULL / 0 / NULL for a plain conv. */
int pb200_debug_conv_f16x2(const void* in_nhwc, int B, int H, int W, int Cin, const float* w_oihw_host,
                           const float* bias_host, int Cout, int R, int stride, int pad,
                           const void* residual_nhwc_or_null, int relu, void* out_nhwc,
                           const void* ds_in_nhwc_or_null, int ds_Cin, const float* wds_oi_host_or_null,
                           pb200_stream stream);
/* One 3x3/1 conv with Cout = 128 through the halo-patch / resident-weights pair kernel (csrc/conv_halo.cu, layer2), NHWC bf16,
 * optionally with the fused 1x1/2 projection of ds_in [B][2H][2W][ds_Cin] (weights wds [Cout][ds_Cin]). */
int pb200_debug_conv_halo(const void* in_nhwc, int B, int H, int W, int Cin, const float* w_oihw_host,
                          const float* bias_host, int Cout, const void* residual_nhwc_or_null, int relu, void* out_nhwc,
                          const void* ds_in_nhwc_or_null, int ds_Cin, const float* wds_oi_host_or_null, pb200_stream stream);
int pb200_debug_conv_p8_f16x2(const void* in_p8, int B, int H, int W, const float* w_oihw_host,
                              const float* bias_host, const void* residual_p8_or_null, int relu, void* out,
                              int out_nhwc, pb200_stream stream);

/* ---- smoother factors (float64, batched) ---------------------------------------------- */

/* KeypointProjectionFactor over N poses x P body points.
 *   poses      : double [N][12]  row-major R (9) then t (3)    (gtsam.Pose3)
 *   meas       : double [N][P][2]
 *   pts_body   : double [P][3]
 *   K5         : double [5] fx, fy, s, u0, v0                  (gtsam.Cal3_S2)
 *   cam_pose12 : double [12] camera pose in the world, or NULL = identity
 *   resid      : double [N][P][2]
 *   jac        : double [N][P][12] 2x6 column-major (Fortran, as GTSAM expects), tangent order
 *                [rot; trans], right perturbation; or NULL
 *   cheirality : uint8 [N][P] 1 where the point is behind the camera (GTSAM would throw), or NULL */
int pb200_factors_projection(const double* poses, const double* meas, const double* pts_body,
                             const double* K5, const double* cam_pose12_or_null, double* resid,
                             double* jac_or_null, uint8_t* cheirality_or_null, int64_t N, int P,
                             pb200_stream stream);

/* PoseDynamicsFactor over N (pose1, ang_vel1, vel1, pose2) tuples.
 *   resid : double [N][6] [rot; trans];  H0,H3 : double [N][36] 6x6 column-major;
 *   H1,H2 : double [N][18] 6x3 column-major; all four NULL = residual only. */
int pb200_factors_dynamics(const double* pose1, const double* ang_vel1, const double* vel1,
                           const double* pose2, double dt, int world_frame, double* resid,
                           double* H0, double* H1, double* H2, double* H3, int64_t N,
                           pb200_stream stream);

/* ConstantVelocityFactor: resid = vel2 - vel1 [N][3]; H0 = -I, H1 = +I ([N][9] each) or NULL. */
int pb200_factors_constvel(const double* vel1, const double* vel2, double* resid, double* H0,
                           double* H1, int64_t N, pb200_stream stream);

/* One launch for a whole chain problem (what the smoother evaluates per linearisation): W windows of T states,
 * poses [W][T][12], ang_vel / vel [W][T][3], meas [W][T][P][2].  Evaluates every KeypointProjectionFactor
 * (factors.py:216-275), every PoseDynamicsFactor between consecutive states of a window (:54-142, pose1 / pose2 read from
 * the one pose array) and both ConstantVelocityFactor residuals (:160-171; their Jacobians are the constants -I, I).
 * Outputs have the layouts of the three calls above: proj_resid [W*T][P][2], proj_jac [W*T][P] 2x6 column-major,
 * cheirality [W*T][P]; dyn_resid [W*(T-1)][6], H0..H3 column-major, cv_*_resid [W*(T-1)][3] (unused when T == 1). */
int pb200_factors_chain(const double* poses, const double* ang_vel, const double* vel, const double* meas,
                        const double* pts_body, const double* K5, const double* cam_pose12_or_null, double dt,
                        int world_frame, double* proj_resid, double* proj_jac, uint8_t* cheirality,
                        double* dyn_resid, double* H0, double* H1, double* H2, double* H3, double* cv_angvel_resid,
                        double* cv_vel_resid, int W, int T, int P, pb200_stream stream);

/* ---- training-step operators (train.py:259-304) -----------------------------------------------
 * Hand-written kernels for the data-parallel training step: tcgen05 forward / data-gradient / weight-gradient
 * convolutions on NHWC bf16, batch-statistics BatchNorm forward + backward, max-pool backward, the regression
 * head, SmoothL1 loss, and the fused gradient-norm + clip + AdamW update.  Every entry is asynchronous on
 * `stream`, allocates nothing, and takes DEVICE pointers.  Replaces loss.backward() through the torchvision
 * module tree + torch.optim.AdamW (train.py:287-303). */

/* fp32 OIHW parameter -> bf16 GEMM operand.  mode 0: forward [Cout][(r,s,ci)];
 * mode 1: data-gradient [Cin][(R-1-r,R-1-s,co)] (flipped + transposed filter). */
int pb200_train_pack_weights(const float* w_oihw, int Cout, int Cin, int R, int mode, void* out_bf16,
                             pb200_stream stream);
/* every weight pack of a step in one launch.  table_dev: int64 [n][8] on the device, one row per conv:
 * {w fp32 OIHW ptr, forward operand ptr, data-gradient operand ptr or 0, Cout, Cin, R, first TILE index,
 *  kind (0 conv, 1 conv1 -> [64][256] im2col operand)}; a conv has (Cout/32)*(Cin/32) tiles (both multiples of 32),
 *  conv1 has 64; total = sum of the tile counts. */
int pb200_train_pack_all(const int64_t* table_dev, int n, int64_t total, pb200_stream stream);
/* convolution on NHWC bf16 with device-packed weights; out = [relu](conv + bias [+ residual]).
 * stats_or_null: double [2][Cout], ZEROED BY THE CALLER: the epilogue adds the per-channel sum and sum of squares of the
 * bf16 outputs (the batch statistics pb200_train_bn_apply consumes) — no separate statistics pass over the tensor. */
int pb200_train_conv(const void* in_nhwc, int B, int H, int W, int Cin, const void* wpacked, const float* bias,
                     int Cout, int R, int stride, int pad, const void* residual_or_null, int relu,
                     void* out_nhwc, double* stats_or_null, pb200_stream stream);
/* data gradient of a STRIDE-2 convolution (3x3 pad 1 or 1x1 pad 0) without zero-stuffing: one small stride-1 convolution
 * of dz per output parity (2a+ph, 2b+pw) with the 1 / 2 / 2 / 4 filter taps that reach it (1 / 0 / 0 / 0 for 1x1), written
 * straight to the strided pixels.  dz bf16 [B][Ho][Wo][Cout]; w_dgrad_packed = pack mode 1 of the layer; dx bf16
 * [B][2Ho][2Wo][Cin]; residual00 (same shape as dx) is added at the parity-(0,0) pixels only — the block's 1x1/2 projection
 * gradient, which exists only there.  A 1x1 layer writes ONLY the parity-(0,0) pixels of dx. */
int pb200_train_dgrad_s2(const void* dz_nhwc, int B, int Ho, int Wo, int Cout, const void* w_dgrad_packed,
                         const float* zero_bias, int Cin, int R, const void* residual00_or_null, void* dx_nhwc,
                         pb200_stream stream);
/* weight gradient: scratch[R*R][Cin][Cout] (fp32, zeroed by the caller) += x^T (*) dz over all pixels. */
int pb200_train_wgrad(const void* x_nhwc, const void* dz_nhwc, int B, int H, int W, int Cin, int Cout, int R,
                      int stride, int pad, float* scratch, pb200_stream stream);
/* scratch [R*R][Cin][Cout] -> OIHW fp32 [Cout][cin_real][R][R] (cin_real <= Cin drops padded channels). */
int pb200_train_unpack_wgrad(const float* scratch, int Cout, int Cin, int R, int cin_real, float* dw_oihw,
                             pb200_stream stream);

/* Batch-statistics BatchNorm on NHWC bf16 [M][C] (nn.BatchNorm2d in train mode; C % 8 == 0, C <= 512).
 * bn_stats: sums[2*C] (fp64; zeroed here) = per-channel sum and sum of squares of z.
 * bn_apply: y = [relu](gamma * (z - mean) * invstd + beta [+ residual]); stores mean / invstd for the backward
 *           and updates running_mean / running_var (momentum, unbiased variance) when they are non-null. */
int pb200_train_bn_stats(const void* z, int64_t M, int C, double* sums, pb200_stream stream);
int pb200_train_bn_apply(const void* z, const double* sums, const float* gamma, const float* beta,
                         float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                         float* save_invstd, const void* residual_or_null, int relu, void* y, int64_t M, int C,
                         pb200_stream stream);
/* BatchNorm backward fused with the ReLU mask of the layer output: g = dy * [y_mask > 0]; y_mask null and beta
 * given: the mask is recomputed from z with the forward's affine map (layers without a residual input) — one
 * tensor less to read per pass; both null: g = dy;
 * dz = gamma * invstd * (g - mean(g) - xhat * mean(g * xhat)); dgamma = sum g * xhat; dbeta = sum g.
 * Optionally also writes g (the gradient the identity branch receives) and a zero-stuffed copy of dz,
 * [B][2*Ho][2*Wo][C] with dz at the even positions, which a stride-2 data gradient convolves at stride 1. */
int pb200_train_bn_backward(const void* dy, const void* y_mask_or_null, const void* z, const float* save_mean,
                            const float* save_invstd, const float* gamma, const float* beta_or_null, double* sums,
                            int64_t M, int C, void* dz,
                            void* g_out_or_null, void* stuffed_out_or_null, int Ho, int Wo, float* dgamma,
                            float* dbeta, pb200_stream stream);
/* conv1 (7x7/2) as a GEMM: im2col of the NCHW fp32 batch into bf16 [B*H/2*W/2][256] (k = (r*7+s)*C + c, zero
 * tail), the matching weight pack [64][256] and the gradient unpack from the wgrad scratch [256][64]. */
int pb200_train_stem_im2col(const float* x_nchw, int B, int C, int H, int W, void* A_bf16, pb200_stream stream);
int pb200_train_stem_pack(const float* w_oihw, int C, void* out_bf16, pb200_stream stream);
int pb200_train_stem_unpack(const float* scratch, int C, float* dw_oihw, pb200_stream stream);
/* 3x3/2 pad-1 max pool on NHWC bf16 with the argmax (one byte per element) kept for the backward. */
int pb200_train_maxpool(const void* in, void* out, void* idx, int B, int H, int W, int C, pb200_stream stream);
int pb200_train_maxpool_backward(const void* dout, const void* idx, void* din, int B, int H, int W, int C,
                                 pb200_stream stream);
/* avgpool + fc forward keeping the pooled features; SmoothL1Loss(beta, mean) with d loss / d pred; the head's
 * backward (fc gradients + the gradient of the last activation, broadcast over its HW pixels). */
int pb200_train_head_forward(const void* y_nhwc, const float* fc_w, const float* fc_b, float* feat, float* pred,
                             int B, int HW, int C, int n_out, pb200_stream stream);
int pb200_train_smooth_l1(const float* pred, const float* target, int n, float beta, float* loss, float* dpred,
                          pb200_stream stream);
int pb200_train_head_backward(const float* dpred, const float* feat, const float* fc_w, int B, int HW, int C,
                              int n_out, float* d_fc_w, float* d_fc_b, void* dy_nhwc, pb200_stream stream);
/* clip_grad_norm_(max_norm) + AdamW on flat fp32 buffers: grads are first scaled by grad_scale (1/world after a
 * sum all-reduce), the global norm of the scaled gradient is reduced into sqnorm_scratch, then one fused update. */
int pb200_train_adamw_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                           float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                           float grad_scale, float max_norm, double* sqnorm_scratch, pb200_stream stream);

/* ---- train-time augmentation -----------------------------------------------------------
 *   pb200_augment <- KeypointAugmentation.forward      perseus/detector/augmentations.py:434-488
 *                    RandomTransplantationWithDepth    :172-251
 *                    kornia RandomAffine / RandomErasing x2 / RandomPlanckianJitter / ColorJiggle /
 *                    RandomGaussianBlur / RandomPlasmaShadow as configured at :351-392
 *                    DepthBias / DepthGaussianNoise / DepthPlane augmentations :17-169, 395-431
 *                    normalize_pixel_coordinates       :481
 * Per-sample draws are inputs (the host samples them); per-pixel depth noise is Philox4x32-10(seed; pixel, image).
 *   iparams [B][PB200_AUG_IPARAMS] int32:  0 donor image (-1 none); 1-4 / 5-8 erase rectangles x, y, w, h (w = 0: none);
 *            9 colour-jiggle order, operator k in nibble k (0 brightness 1 contrast 2 saturation 3 hue);
 *            10 PB200_AUG_* flags; 11 reserved
 *   fparams [B][PB200_AUG_FPARAMS] float:  0-5 forward affine matrix M (2x3, pixels; key points);  6-11 sampling map S
 *            (2x3: source position of output pixel (x, y), kornia warp_affine conventions folded in);
 *            12,13 Planckian red / blue gains; 14 brightness, 15 contrast, 16 saturation factors, 17 hue shift (rad);
 *            18-22 the 5 Gaussian blur taps; 23 shadow intensity, 24 shadow quantity; rest reserved
 *   plasma  [B][H][W] float fractal map in [0,1] for the shadow, or NULL
 *   counts  [B] int32 scratch (transplanted cube-pixel counts; written by the call)
 *   images  float [B][C_in][H][W], C_in 3..5 (R,G,B[,depth[,segmentation]]); out_images float [B][C_out][H][W]
 *   keypoints float [B][K][2] pixels -> out_keypoints normalised to [-1, 1]; both may be NULL */
#define PB200_AUG_IPARAMS 12
#define PB200_AUG_FPARAMS 32
#define PB200_AUG_AFFINE 1
#define PB200_AUG_PLANCKIAN 2
#define PB200_AUG_JIGGLE 4
#define PB200_AUG_BLUR 8
#define PB200_AUG_SHADOW 16
typedef struct pb200_aug_config {
  float cube_scale;        /* AugmentationConfig.cube_scale (0.035) */
  float noise_cube_scale;  /* DepthGaussianNoiseAugmentation keeps its own default (0.035), augmentations.py:396 */
  float lb_seg_ratio, ub_seg_ratio; /* transplantation acceptance bounds (0.02, 0.7) */
  int training;            /* F.dropout active (nn.Module.training) */
  int depth_bias;   float dev_bias, p_bias;
  int depth_noise;  float std_noise;
  int near_plane;   float near_mean, near_dev, p_near, near_value;
  int far_plane;    float far_mean, far_dev, p_far, far_value;
} pb200_aug_config;
int pb200_augment(const float* images, const int32_t* iparams, const float* fparams, const float* plasma_or_null,
                  const pb200_aug_config* cfg /* host */, uint64_t seed, int32_t* counts,
                  const float* keypoints_or_null, float* out_images, float* out_keypoints_or_null, int B, int C_in,
                  int C_out, int H, int W, int K, pb200_stream stream);
/* Diamond-square plasma fractal (kornia.contrib.diamond_square as RandomPlasmaShadow uses it): float [B][H][W]
 * in [0,1] (per-image min-max normalised); roughness float [B]; scratch float [B][S][S], S = 2^k + 1 >= max(H, W)
 * (pb200_plasma_side). */
int pb200_plasma_side(int H, int W);
int pb200_plasma(const float* roughness, uint64_t seed, float* scratch, float* out, int B, int H, int W,
                 pb200_stream stream);
/* pb200_plasma without caller scratch and without ANY allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The B x S x S working grids live in a static device array of PB200_PLASMA_STATIC_MAX_FLOATS floats
 * (4.03 MB = sixteen 257 x 257 grids, held by every GPU that loads the library until exit); B * S * S beyond that is
 * rejected with PB200_ERR_INVALID and a message naming the limit.  Same arguments otherwise, bit-identical output.  All calls
 * share the array: calls that overlap on the GPU (different streams / threads) race; calls on one stream are ordered.  The
 * first call on each device resolves the array's device address (cudaGetSymbolAddress): make it outside graph capture. */
#define PB200_PLASMA_STATIC_MAX_FLOATS (16 * 257 * 257)
int pb200_plasma_static(const float* roughness, uint64_t seed, float* out, int B, int H, int W, pb200_stream stream);

/* ---- fixed-lag smoother linearisation -----------------------------------------------------
 * The consumer of the factor entries above: GTSAM's whitening (noiseModel.Diagonal sigmas, factors.py:42,49) +
 * Gauss-Newton system + solve, specialised to the tracking chain.  W independent windows of T states; state k =
 * [pose tangent (rot, trans); ang_vel; vel] (12).  Inputs are the factor entries' outputs in place:
 *   proj_resid [W*T][P][2], proj_jac [W*T][P][12]      (pb200_factors_projection on the W*T poses)
 *   dyn_resid [W*(T-1)][6], H0..H3                     (pb200_factors_dynamics on (k, k+1) pairs, window-major)
 *   cv_*_resid [W*(T-1)][3]                            (pb200_factors_constvel on consecutive ang_vel / vel)
 *   sigmas13 (HOST): pixel, dynamics[6], const ang_vel[3], const vel[3]
 *   D [W][T][12][12] row-major diagonal blocks of H = sum J^T W J (+ lambda I, + prior I on state 0),
 *   O [W][T-1][12][12] blocks H[k+1][k],  g [W][T][12] = sum J^T W r
 * pb200_smoother_solve: block-tridiagonal Cholesky, H delta = -g; work double [W][T][2][144]; delta [W][T][12];
 * status int32 [W] (1 = a non-positive pivot was met) or NULL. */
int pb200_smoother_normal_equations(const double* proj_resid, const double* proj_jac, const double* dyn_resid,
                                    const double* H0, const double* H1, const double* H2, const double* H3,
                                    const double* cv_angvel_resid, const double* cv_vel_resid,
                                    const double* sigmas13 /* host */, double lambda, double prior, double* D,
                                    double* O, double* g, int W, int T, int P, pb200_stream stream);
int pb200_smoother_solve(const double* D, const double* O, const double* g, double* work, double* delta,
                         int32_t* status_or_null, int W, int T, pb200_stream stream);

/* x <- x (+) delta for n states (gtsam Values::retract): pose * Pose3::Expmap(delta[0:6]) (tangent [rot; trans]),
 * ang_vel + delta[6:9], vel + delta[9:12].  Outputs may alias the inputs. */
int pb200_smoother_retract(const double* poses, const double* ang_vel, const double* vel, const double* delta,
                           double* poses_out, double* ang_vel_out, double* vel_out, int64_t n, pb200_stream stream);

/* Fixed-lag window advance, in place, W windows of T states: drops state 0, appends the dynamics factor's own prediction
 * of the newest state (factors.py:100-105: pose * Exp(dt [ang_vel; v_body]); velocities carried over) and writes the
 * detector's pixel keypoints (float32 [W][P][2], the layout pb200_forward's out_px has) as its measurement row. */
int pb200_smoother_push(double* poses, double* ang_vel, double* vel, double* meas, const float* keypoints_px, double dt,
                        int world_frame, int W, int T, int P, pb200_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* PERSEUS_B200_H_ */
