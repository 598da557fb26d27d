/*
 * scflow_b200 -- C ABI of the B200-native SCFlow refinement hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types.  Every entry point
 * replaces the native work that one reference function reaches through PyTorch (cited as
 * file:line relative to GiaKhangLuu/SCFlow).  The reference binds nothing natively itself (it
 * is pure Python); the binding a maintainer adds is the ctypes stub shown in INTEGRATION.md.
 *
 * Conventions
 *  - All tensor arguments are DEVICE pointers to contiguous fp32 data unless stated otherwise.
 *  - Launchers are asynchronous: they only enqueue on `stream` (a cudaStream_t passed as
 *    void*), never synchronise, never allocate or free device memory and keep no references.
 *  - Return 0 on success, a negative SCFLOW_E_* otherwise (scflow_strerror() names it).
 *    Unsupported configurations are rejected, never routed to another implementation.
 *  - Re-entrant: no global mutable state besides one-time function-attribute setup.
 */
#ifndef SCFLOW_B200_H_
#define SCFLOW_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SCFLOW_ABI_VERSION 7

enum {
  SCFLOW_OK = 0,
  SCFLOW_E_BAD_SHAPE = -1,      /* dimension not supported by the kernel                  */
  SCFLOW_E_UNSUPPORTED = -2,    /* flag / mode the kernels do not implement               */
  SCFLOW_E_MISALIGNED = -3,     /* pointer not 16-byte aligned                            */
  SCFLOW_E_NULL = -4,           /* required pointer is NULL                               */
  SCFLOW_E_LAUNCH = -5,         /* cudaGetLastError() after the launch was not success    */
  SCFLOW_E_WORKSPACE = -6,      /* workspace too small                                    */
  SCFLOW_E_DRIVER = -7,         /* driver entry point (cuTensorMapEncodeTiled) missing    */
  SCFLOW_E_UNIMPLEMENTED = -8,
  SCFLOW_E_STATIC_SCRATCH = -9  /* problem needs more scratch than the library's static device array */
};

/* scflow_conv_tc_desc.epi */
#define SCFLOW_EPI_PLAIN 0
#define SCFLOW_EPI_GRU_ZR 1
#define SCFLOW_EPI_GRU_GATE 2

/* scflow_conv_tc_desc.algo */
#define SCFLOW_CONV_AUTO 0
#define SCFLOW_CONV_TAP_BOXES 1
#define SCFLOW_CONV_PATCH 2
#define SCFLOW_CONV_PATCH_PAIR 3

/* correlation arithmetic */
#define SCFLOW_CONV_DOT_PITCH 20  /* floats per channel row of scflow_conv_tc_desc.dot_w: 3 x 3 taps x 2 outputs, padded */
#define SCFLOW_STATS_REPLICAS 8 /* accumulators per (image, channel) of scflow_conv_tc_desc.stats */
#define SCFLOW_PREC_FP32 0      /* CUDA-core fp32 FMA (exact-mode parity anchor)          */
#define SCFLOW_PREC_TF32 1      /* tcgen05 kind::tf32, fp32 accumulate in TMEM (default)  */

/* how `x * 2. / max(W-1, 1)` (models/utils/corr_lookup.py:64-65) is evaluated */
#define SCFLOW_DIV_IEEE 0       /* true fp32 division: what ATen does on the CPU          */
#define SCFLOW_DIV_RCP 1        /* x*2 * (1/(W-1)): ATen's CUDA div-by-scalar fast path   */

int scflow_abi_version(void);
const char* scflow_strerror(int rc);
/* last cudaError_t seen by a failing launcher on this thread (0 if none) */
int scflow_last_cuda_error(void);

/* ---------------------------------------------------------------------------------------------
 * f4  Around the hot path: the renderer call and the pose-error evaluator (csrc/render.cu).
 *
 * scflow_render_fwd: what Renderer.forward (models/utils/rendering.py:185-248) asks of pytorch3d with the shipped
 * configuration (configs/refine_models/scflow_ycbv_real.py:261-274: HardPhongShader, one face per pixel, blur radius
 * 0, seperate_lights) and what the refiner reads back (models/refiner/base_refiner.py:106-112).  Meshes are packed:
 * verts / normals / colors (sum V, 3), faces (sum F, 3) with vertex indices LOCAL to their mesh, voff / foff
 * [classes + 1] offsets, vmax / fmax the largest mesh; per image: labels (int64), R (n,3,3), t (n,3), K (n,3,3)
 * (OpenCV).  Outputs: image (n, 3, h, w) planar, zbuf (n, h, w) view-space z (-1 = background), mask (n, h, w) or
 * NULL.  workspace: scflow_render_workspace_bytes(n, vmax, h, w) bytes.  No host synchronisation.
 *
 * scflow_add_error_fwd: ADD / ADD-S (symmetric[class] != 0: nearest predicted vertex) mean 3-D error and mean 2-D
 * projection error per object, ADD.eval_pose_error (metrics/add.py:354-394) + project_3d_point
 * (datasets/pose.py:18-70).  err3d / err2d: (n) doubles.
 * ------------------------------------------------------------------------------------------- */
size_t scflow_render_workspace_bytes(int n, int vmax, int h, int w);
int scflow_render_fwd(const float* verts, const float* normals, const float* colors, const int* faces, const int* voff,
                      const int* foff, int vmax, int fmax, const long long* labels, const float* R, const float* t,
                      const float* K, int n, int h, int w, float bg_r, float bg_g, float bg_b, float* image,
                      float* zbuf, float* mask, void* workspace, void* stream);
int scflow_add_error_fwd(const float* verts, const int* voff, int vmax, const long long* labels, const int* symmetric,
                         const float* gt_r, const float* gt_t, const float* pred_r, const float* pred_t,
                         const float* K, double* err3d, double* err2d, int n, void* stream);

/* ---------------------------------------------------------------------------------------------
 * a1  CorrelationPyramid.forward   models/decoder/raft_decoder.py:35-58
 *     replaces torch.matmul (:48-50), the /sqrt(C) pass (:51-52) and 3x nn.AvgPool2d (:54-56).
 * feat1, feat2: (n, c, h, w).  levels[l]: (n*h*w, 1, h>>l, w>>l), l < num_levels (host array of
 * device pointers).  corr[n*P+q][t] = sum_c feat1[n,c,q] feat2[n,c,t] / sqrt(c).
 * The tcgen05 TF32 kernel needs c % 32 == 0, h % 8 == 0, w % 32 == 0, (h*w) % 128 == 0; a SCFLOW_PREC_TF32 request on
 * any other shape runs the fp32 FMA kernel (a stricter result, not a refusal).  num_levels <= 4.
 * ------------------------------------------------------------------------------------------- */
size_t scflow_corr_pyramid_workspace_bytes(int n, int c, int h, int w, int precision);
int scflow_corr_pyramid_fwd(const float* feat1, const float* feat2, float* const* levels,
                            int num_levels, int n, int c, int h, int w, int precision,
                            void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * a2  CorrLookup.forward           models/utils/corr_lookup.py:102-136 (+ bilinear_sample :31-67)
 *     replaces 4x F.grid_sample and ~25 glue launches per iteration.
 * levels as produced above; flow: (n, 2, h, w) in 1/8-resolution pixels;
 * out: (n, num_levels*(2r+1)^2, h, w), channel = (2r+1)^2*l + (2r+1)*a + b, a = x offset.
 * windows (optional, may be NULL): int16 (n*h*w, num_levels, 2, 2r+1) = floor of the
 * un-normalised sample coordinate per x tap / per y tap -- the window selection, exported so
 * tests can check it bit-exactly.  Only radius == 4, bilinear, zeros, align_corners=True.
 * out_layout: SCFLOW_LAYOUT_NCHW writes (n, 324, h, w) as the reference does; SCFLOW_LAYOUT_NHWC
 * writes (n, h, w, 324) -- same values, channels-last memory, the layout the downstream 1x1
 * convolution and the ConvGRU kernel consume (one warp per query, TMA-staged windows).
 * ------------------------------------------------------------------------------------------- */
#define SCFLOW_LAYOUT_NCHW 0
#define SCFLOW_LAYOUT_NHWC 1
#define SCFLOW_LAYOUT_NHWC_GEN1 2   /* NHWC through the first-generation kernel: the cross-check of the tests */
#define SCFLOW_LAYOUT_NHWC_SPLIT 3  /* `out` = split-bf16 planes (n, h, w, [328 hi | 328 lo]) bf16, value = hi + lo,
                                       channels 324..327 zero: the operand of the 1x1 convolution that consumes the
                                       lookup (scflow_conv_tc_fwd); w % 32 == 0, else SCFLOW_E_UNSUPPORTED.
                                       Plane channel order is LANE-MAJOR: plane channel 12 (9 g + a) + 3 level + j
                                       holds the reference's channel 81 level + 9 a + 3 g + j (a, g*3 + j = x, y tap);
                                       pack the consumer's filter with its input channels in that order */
int scflow_corr_lookup_fwd(const float* const* levels, int num_levels, const float* flow,
                           float* out, short* windows, int n, int h, int w, int radius,
                           int div_mode, int out_layout, void* stream);

/* ---------------------------------------------------------------------------------------------
 * a3  ConvGRU.forward (SeqConv)    models/decoder/raft_decoder.py:235-253
 *     replaces 6 cuDNN convs + 6 activation kernels + 4 cat + 8 elementwise per iteration.
 * Since ABI v5 the gates run on the split-bf16 convolution family below (scflow_conv_tc_fwd with
 * epi = SCFLOW_EPI_GRU_ZR / SCFLOW_EPI_GRU_GATE: three launches per pass, fp32-grade products) instead of
 * the TF32 kernels of ABI v4 (scflow_convgru_fwd / _pass_fwd: one TF32 MMA per product cost 5.7e-5 -
 * 8.9e-5 rad of the 1e-4 rad pose budget in this module alone).
 * ------------------------------------------------------------------------------------------- */

/* Channels-last packing of the GRU input x = cat[cxt, motion, flow] (raft_decoder.py:166 and
 * scflow_decoder.py:208) with round-to-nearest TF32, replacing two torch.cat and a layout pass.
 * dst: (npix, cdst) channels-last; channels [dst_off, dst_off + ca + cb + cf) are written from
 *   a: (npix, ca) channels-last (may be NULL with ca = 0), b: (npix, cb) channels-last (ditto),
 *   f: (n, cf, P) planar, P = pixels per image (may be NULL with cf = 0).
 * round_tf32 != 0 rounds every value to the nearest TF32 (what tcgen05 kind::tf32 would otherwise
 * truncate). */
int scflow_pack_nhwc(const float* a, int ca, const float* b, int cb, const float* f, int cf,
                     float* dst, int cdst, int dst_off, long long npix, int P, int round_tf32,
                     void* stream);

/* ---------------------------------------------------------------------------------------------
 * a4  cal_3d_2d_corr / lift_2d_to_3d    models/utils/pose.py:26-64
 *     replaces per crop: nonzero (host sync), 2x inverse, 2x mm, stack.
 * depth (n,H,W); K,R0 (n,3,3); t0 (n,3) -> points3d (n,3,H,W) object-frame (0 on background),
 * mask (n,H,W) uint8 (may be NULL).  Dense form of the ragged lists; no host sync.
 * ------------------------------------------------------------------------------------------- */
int scflow_lift_points(const float* depth, const float* K, const float* R0, const float* t0,
                       float* points3d, unsigned char* mask, int n, int H, int W, void* stream);

/* ---------------------------------------------------------------------------------------------
 * a6 + a5 + a7(i)  one launch per refinement iteration:
 *   get_pose_from_delta_pose             models/utils/pose.py:124-169  (ortho6d, 'exp' depth)
 *   get_flow_from_delta_pose_and_points  models/utils/pose.py:66-88
 *   1/scale * F.interpolate(flow, 1/scale, bilinear, align_corners=True)
 *                                        models/decoder/scflow_decoder.py:197-198
 * dR6 (n,6), dt (n,3) may both be NULL: then (R_prev,t_prev) is used as the pose directly.
 * Object points come from `points3d`+`mask` if points3d != NULL, else are recomputed from
 * depth/K/R0/t0 (4 B/pixel instead of 13).  Outputs (each may be NULL to skip):
 *   R_out (n,3,3), t_out (n,3), flow_full (n,2,H,W), flow_lr (n,2,H/scale,W/scale)
 * flow_lr is the NEXT iteration's low-resolution flow, already multiplied by 1/scale.
 * ------------------------------------------------------------------------------------------- */
int scflow_pose_flow_fwd(const float* dR6, const float* dt, const float* R_prev,
                         const float* t_prev, const float* K, const float* R0, const float* t0,
                         const float* depth, const float* points3d, const unsigned char* mask,
                         float* R_out, float* t_out, float* flow_full, float* flow_lr,
                         int n, int H, int W, int scale, float invalid_num, float weight,
                         void* stream);

/* ---------------------------------------------------------------------------------------------
 * a7(ii,iii)  scale * F.interpolate(x [+ y], scale, bilinear, align_corners=True)
 *             models/decoder/scflow_decoder.py:223-228
 * in, add (may be NULL): (n, c, h, w) -> out (n, c, h*scale, w*scale), out = mul * up(in + add)
 * ------------------------------------------------------------------------------------------- */
int scflow_upsample_fwd(const float* in, const float* add, float* out, int n, int c, int h,
                        int w, int scale, float mul, void* stream);

/* align_corners=True bilinear down-sampling by an integer factor, out = mul * down(in):
 * the stand-alone form of a7(i) for callers that hold a full-resolution flow already. */
int scflow_downsample_fwd(const float* in, float* out, int n, int c, int H, int W, int scale,
                          float mul, void* stream);

/* ---------------------------------------------------------------------------------------------
 * f3 (glue)  The InstanceNorm2d / eval BatchNorm2d + ReLU + residual passes of the RAFT encoders
 *            (models/encoder/raft_encoder.py:286-314, BasicBlock models/backbone/resnet.py:12-92) are part
 *            of the convolution family below: InstanceNorm statistics are accumulated by the convolution
 *            epilogue (scflow_conv_tc_desc.stats) and applied by scflow_norm_apply_split; eval BatchNorm
 *            is folded into the packed filters.  What remains here is layout glue.
 * ------------------------------------------------------------------------------------------- */
/* (n, P, C) channels-last -> (n, C, P) planar, with optional round-to-nearest TF32: the hand-over
 * of channels-last encoder features to scflow_corr_pyramid_fwd (planar operands), replacing a
 * torch .contiguous() pass.  in != out. */
int scflow_nhwc_to_nchw(const float* in, float* out, int n, int P, int C, int round_tf32, void* stream);

/* GroupNorm with 4 channels per group (+ ReLU) on channels-last (n, P, C), one launch: the conv -> GN ->
 * ReLU layers of MultiClassPoseHead (models/head/pose_head.py:128-137; norm_cfg GN num_groups=32 on 128
 * channels).  gamma/beta: [C].  x == out allowed.  C/4 must divide 512.  Outputs (either or both): out fp32
 * (n, P, C); out_split = the result as split planes (out_pitch bf16 per pixel, lo plane out_lo further), the
 * operand of the next scflow_conv_tc_fwd. */
int scflow_group_norm4_act_nhwc(const float* x, const float* gamma, const float* beta, float* out, void* out_split,
                                long long out_pitch, int out_lo, int n, int P, int C, float eps, int relu,
                                void* stream);

/* Fully connected tail of MultiClassPoseHead (models/head/pose_head.py:139-147, 201-211): fc0 + ReLU, fc1 +
 * ReLU, rotation_pred / translation_pred restricted to the class of label[0] (index_select(...)[:, 0, :],
 * pose_head.py:209-210).  x: (M, K) row-major; w0: (n0, K), w1: (n1, n0), wr: (6*classes, n1), wt:
 * (3*classes, n1) row-major (torch Linear layout); label: int64 [M] on the device; workspace:
 * scflow_pose_fc_workspace_floats(M, n0, n1, splitk) floats; dr: (M, 6), dt: (M, 3).  Exact fp32, deterministic. */
size_t scflow_pose_fc_workspace_floats(int M, int n0, int n1, int splitk);
int scflow_pose_fc_fwd(const float* x, const float* w0, const float* b0, const float* w1, const float* b1,
                       const float* wr, const float* br, const float* wt, const float* bt,
                       const long long* label, float* workspace, float* dr, float* dt, int M, int K, int n0,
                       int n1, int num_class, int splitk, void* stream);
/* The last stage alone (pose_head.py:144-147, 201-211): a1 = ReLU(b1 + sum over `splitk` blocks of part1
 * (splitk, M, n1)), then rotation_pred / translation_pred restricted to the class of label[0].  With splitk = 1,
 * part1 is fc1's pre-activation without bias: the hand-over from the tensor-core form of fc0 / fc1
 * (scflow_conv_tc_fwd: fc0 is the hs x ws convolution that flatten + Linear amounts to, fc1 a 1x1 one). */
int scflow_pose_tail_fwd(const float* part1, const float* b1, const float* wr, const float* br, const float* wt,
                         const float* bt, const long long* label, float* dr, float* dt, int M, int splitk, int n1,
                         int num_class, void* stream);


/* scflow_pose_fc_fwd without a workspace argument and without any allocation: the split-K partial sums live
 * in a file-scope `static __device__` array of the library (SCFLOW_POSE_FC_STATIC_MAX_FLOATS floats = 4 MiB,
 * held by every GPU that loads the library until exit).  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, results and launches as scflow_pose_fc_fwd otherwise.
 * Limit: scflow_pose_fc_workspace_floats(M, n0, n1, splitk) <= SCFLOW_POSE_FC_STATIC_MAX_FLOATS, else
 * SCFLOW_E_STATIC_SCRATCH (the shipped head, n0 = 1024, n1 = 256, splitk = 4, fits up to M = 204 crops).
 * Every call on a device shares the array: nothing in it needs clearing (it is fully overwritten before it
 * is read), but calls that overlap on the GPU (different streams or threads) race -- serialise them. */
#define SCFLOW_POSE_FC_STATIC_MAX_FLOATS (1 << 20)
int scflow_pose_fc_fwd_static(const float* x, const float* w0, const float* b0, const float* w1, const float* b1,
                              const float* wr, const float* br, const float* wt, const float* bt,
                              const long long* label, float* dr, float* dt, int M, int K, int n0, int n1,
                              int num_class, int splitk, void* stream);

/* ------------------------------------------------------------------ degenerate convolutions (8f-f2)
 * Convolutions of the decoder with 1-2 input or 1-2 output channels, which cuDNN serves with generic
 * engines: MotionEncoder.flow_net[0] / delta_flow_encoder[0] (7x7, 2 -> 128), mask_encoder[0] (3x3,
 * 1 -> 64) -- raft_decoder.py:152-160, scflow_decoder.py:96-106 -- and XHead.predict_layer (3x3
 * 256 -> 2, 1x1 256 -> 1; raft_decoder.py:281-287).  Stride 1, zero padding k/2, exact fp32. */

/* out (n, h, w, cout) channels-last = act(conv(in) + bias).  in: element (n, ci, y, x) at
 * n*stride_n + ci*stride_c + (y*w + x)*stride_px floats (planar or channels-last callers);
 * wpack: [cin][k][k][cout] (Conv2d weight (cout,cin,k,k) permuted 1,2,3,0); cin in {1,2},
 * ksize in {3,7}, cout % 32 == 0, w % 8 == 0.  Outputs (either or both): out fp32, out_split = the same
 * values as split-bf16 planes (out_pitch bf16 per pixel, lo plane out_lo further): the operand of the
 * tensor-core convolution that follows in flow_net / delta_flow_encoder / mask_encoder. */
int scflow_conv_small_cin_fwd(const float* in, long long stride_n, long long stride_c, long long stride_px,
                              const float* wpack, const float* bias, float* out, void* out_split,
                              long long out_pitch, int out_lo, int n, int cin, int h, int w, int cout, int ksize,
                              int relu, void* stream);

/* Stem input of the RAFT encoders (raft_encoder.py:286-288: conv1 7x7 stride 2 on 3 channels) as the 2x2
 * space-to-depth image of the equivalent 4x4 stride-1 convolution: x planar (n, c <= 4, h, w), h and w even
 * -> (n, h/2+3, w/2+3, 16), channel 4*c + 2*by + bx = x[n, c, 2Y+by, 2X+bx] at (Y+2, X+2); zero in the
 * unused channels and in the (2 before, 1 after) border = the stride-1 filter's padding.  Written as two bf16 planes [hi | lo] of (n, h/2+3, w/2+3, 16) each (value = hi + lo): the operand
 * of scflow_conv_tc_fwd, which reads the 4x4 stride-1 stem as 4 filter rows of 64 contiguous channels
 * (four x-neighbours) through an overlapped-window view (x_stride_w = 16). */
int scflow_stem_s2d_split_fwd(const float* x, void* out, int n, int c, int h, int w, void* stream);

/* out (n, cout, h, w) planar = conv(in) + bias.  in: (n, h, w, c) channels-last, c % 4 == 0;
 * wpack: [c][k*k][cout] (Conv2d weight (cout,c,k,k) permuted 1,2,3,0); cout in {1,2}, ksize in {1,3},
 * (4 + 2*(ksize/2)) * w <= 1024. */
int scflow_conv_small_cout_fwd(const float* in, const float* wpack, const float* bias, float* out, int n,
                               int c, int h, int w, int cout, int ksize, void* stream);

/* ---------------------------------------------------------------------------------------------
 * f1-f3  The convolution family on the tensor cores (csrc/conv_tc.cu).  Replaces the cuDNN calls
 * behind nn.Conv2d / mmcv ConvModule in
 *   models/encoder/raft_encoder.py:286-314 + models/backbone/resnet.py:12-92   (RAFT encoders)
 *   models/decoder/raft_decoder.py:152-166                                     (MotionEncoder)
 *   models/decoder/raft_decoder.py:256-294                                     (flow / mask heads)
 *   models/decoder/scflow_decoder.py:101-106                                   (delta-flow / mask encoders)
 *   models/head/pose_head.py:108-150                                           (pose-head stride-2 convs)
 * Arithmetic: split-bf16 -- every operand is hi + lo (two bf16), a product is three tcgen05
 * kind::f16 MMAs into one fp32 TMEM accumulator (a_lo.b_hi + a_hi.b_lo + a_hi.b_hi): relative error
 * ~2^-17 per product, fp32 accumulation.  Activations are channels-last SPLIT PLANES: a pixel holds
 * [hi C | lo C] bf16.
 * ------------------------------------------------------------------------------------------- */
typedef struct scflow_conv_tc_desc {
  /* geometry: input (n, h, w, cin), filter (cout, cin, kh, kw), stride 1 or 2, zero padding */
  int n, h, w, cin, cout, kh, kw, stride, pad_h, pad_w;
  int bk;                       /* k-block of the weight pack: 64 or 32 channels               */
  /* input: hi plane at x, lo plane x_lo_off ELEMENTS further; element strides (bf16) between
   * x-neighbours, rows and images (all multiples of 8).  Channels-last split planes of C
   * channels: x_lo_off = C, x_stride_w = 2C, x_stride_h = w*2C, x_stride_n = h*w*2C.            */
  const void* x;
  long long x_lo_off, x_stride_w, x_stride_h, x_stride_n;
  const void* wpack;            /* scflow_conv_tc_pack_weights                                    */
  const float* bias;            /* [cout padded as the pack] or NULL                             */
  int relu;
  /* residual added before the ReLU: split planes (hi at res_split, lo res_lo elements further,
   * res_pitch elements per pixel) or fp32 channels-last (resf_pitch floats per pixel); or NULL  */
  const void* res_split; long long res_pitch; int res_lo;
  const float* res_f32; long long resf_pitch;
  /* outputs (any subset, at least one): split planes, fp32 channels-last, planar fp32 (n, cout, OH*OW)
   * optionally rounded to the nearest TF32 (the correlation kernel's operand format)           */
  void* out_split; long long out_pitch; int out_lo;
  float* out_f32; long long outf_pitch; int f32_tf32;      /* f32_tf32: round to the nearest TF32 */
  float* out_planar; int planar_tf32;
  /* optional: per-(image, channel) sum and sum of squares of the value BEFORE residual/ReLU
   * (after bias), added into stats[n][SCFLOW_STATS_REPLICAS][cout8][2] (zeroed by the caller) -- InstanceNorm
   * statistics; the replicas (picked per warp and tile) keep the fp64 atomics of one image from serialising
   * on a handful of L2 lines, their sum is the statistic */
  double* stats;
  /* tile width override: output channels per tile (multiple of 16 that divides the padded cout), 0 = the
   * widest tile (<= 256).  Narrower tiles trade operand re-reads for fuller waves. */
  int bn;
  /* epilogue: SCFLOW_EPI_PLAIN, or the two ConvGRU epilogues (models/decoder/raft_decoder.py:245-252) over
   * gru_h (n, OH, OW, gru_c) exact fp32 hidden state and gru_z (same shape) update gate:
   *   GRU_ZR    cout = 2 gru_c = [z | r] pre-activations (+ res_f32, 2 gru_c channels, if given): z = sigmoid(.)
   *             -> gru_z; r = sigmoid(.), r * gru_h -> out_split (gru_c channels)
   *   GRU_GATE  cout = gru_c, res_f32 = the x-part of the q convolution: q = tanh(acc + res + bias);
   *             gru_h = (1 - z) gru_h + z q in place, and as split planes -> out_split */
  int epi, gru_c;
  float* gru_h;
  float* gru_z;
  /* operand movement (ABI v6): SCFLOW_CONV_AUTO picks per layer; the others force one kernel (tests, tools):
   *   TAP_BOXES   one shifted TMA box per (tap, k-block), the filter slab re-read per tile (any stride)
   *   PATCH       stride 1, OW % 8 == 0, OH % 16 == 0, bk = 64: the input patch of a 16 x 8 tile is loaded
   *               once per k-block and every tap reads it through a shifted shared-memory descriptor;
   *               small filters stay resident in shared memory
   *   PATCH_PAIR  the same on CTA pairs (tcgen05 cta_group::2, M = 256): each CTA loads half the filter rows */
  int algo;
  /* optional second input tensor of the same n, h, w: its cin2 channels continue the filter's input channels
   * after the first tensor's cin (which must then be a multiple of bk).  What the reference concatenates in
   * front of a convolution (raft_decoder.py:250: cat([r*h, x])) is read from where it lies. */
  const void* x2;
  int cin2;
  long long x2_lo_off, x2_stride_w, x2_stride_h, x2_stride_n;
  /* optional projection output (ABI v7): dot_out[pixel][slot][k] = sum over the slot's channels c of
   * y[pixel][c] dot_w[c][k], k < dot_n <= dot_pitch <= SCFLOW_CONV_DOT_PITCH (dot_pitch % 4 == 0), y the finished
   * output (after bias / residual / ReLU); dot_w is [cout padded as the pack][SCFLOW_CONV_DOT_PITCH] fp32 (zero
   * rows past cout).  A slot is one (output-channel tile, epilogue warp half): dot_slots = 2 * cout_pad / bn with
   * bn given explicitly; every slot of every pixel is written (plain stores, no initialisation, run-to-run
   * identical bits) and their sum is the projection.  With dot_n = kh' kw' cout' and dot_w[c][tap cout' + co] =
   * W'[co][c][tap] these are the per-pixel tap products of a following small convolution W' (the 1-2 channel
   * predictors of the flow / mask heads, raft_decoder.py:256-294); scflow_tap_sum_fwd adds slots and taps up,
   * and the head's (n, 256, h, w) feature map is never written.  No other output is required with dot_out. */
  const float* dot_w;
  float* dot_out;
  int dot_n, dot_pitch, dot_slots;
  /* bf16 MMAs per product: 0 or 3 = a_lo.b_hi + a_hi.b_lo + a_hi.b_hi (fp32-grade: THE mode of the refiner, the
   * one the parity claims are made in).  Opt-in reduced modes, tolerance stated separately: 2 =
   * a_hi.(b_hi + b_lo): activations rounded to bf16, filters exact; 1 = a_hi.b_hi: plain bf16 operands. */
  int mma_terms;
} scflow_conv_tc_desc;

/* Output spatial size: OH = (h + 2 pad_h - kh) / stride + 1, likewise OW.  Supported: OW a multiple of
 * 128, or 128 % OW == 0 with OH a multiple of 128/OW or 128/OW a multiple of OH (whole small images per
 * tile); cout <= 256 or split evenly into <= 256-channel tiles; cin any (zero-filled to k-steps of 16). */
int scflow_conv_tc_fwd(const scflow_conv_tc_desc* desc, void* stream);

/* out (n, cout, h, w) planar = bias[co] + sum over slots and the ks x ks taps of dot[(n, y + ky - ks/2, x + kx -
 * ks/2)][slot][(ky ks + kx) cout + co] (zero outside the image): finishes a small convolution whose per-pixel tap
 * products scflow_conv_tc_fwd left in dot_out (slots x pitch floats per pixel).  ks 1 or 3, ks ks cout <= pitch. */
int scflow_tap_sum_fwd(const float* dot, int slots, int pitch, const float* bias, float* out, int n, int h, int w,
                       int ks, int cout, void* stream);

/* Weight pack: fp32 (cout, cin, kh, kw) [times scale[cout], e.g. a folded eval BatchNorm] ->
 * [tap][k-block][hi|lo][cout_pad][bk] bf16.  cout_pad = cout rounded up to 16 (<= 256) per tile. */
size_t scflow_conv_tc_packed_bytes(int cout, int cin, int kh, int kw, int bk);
int scflow_conv_tc_pack_weights(const float* w, const float* scale, void* packed, int cout, int cin, int kh,
                                int kw, int bk, void* stream);

/* fp32 channels-last (in_pitch floats per pixel) -> split planes (out_pitch bf16 per pixel, lo plane
 * out_lo elements after the hi plane); channels rounded up to 8 are written (zeros beyond). */
int scflow_split_bf16(const float* in, long long in_pitch, void* out, long long out_pitch, int out_lo,
                      int channels, long long pixels, void* stream);

/* planar fp32 (n, c, h, w), c * kw <= 16, kw odd -> split planes (n, h, w, [16 hi | 16 lo]) holding per pixel its kw
 * x-neighbours of every channel (index dx * c + ci; zero outside the image and in unused slots): the input side of the
 * 1-2 channel convolutions (flow_net[0], delta_flow_encoder[0], mask_encoder[0]; raft_decoder.py:152-160,
 * scflow_decoder.py:96-106) in their tensor-core form, a kh x 1 convolution on 16 channels (scflow_conv_tc_fwd). */
int scflow_expand_taps_fwd(const float* in, void* out, int n, int c, int kw, int h, int w, void* stream);

/* planar fp32 (n, cf, pixels) -> channels [0, cf) of a split-plane slice starting at `out` (hi plane; lo plane
 * out_lo elements further; out_pitch elements per pixel): the flow pair appended to the motion features
 * (raft_decoder.py:165). */
int scflow_split_planar(const float* in, void* out, long long out_pitch, int out_lo, int n, int cf,
                        long long pixels, void* stream);

/* InstanceNorm2d (+ residual) + ReLU of a raw convolution output (n, P, C) fp32 from the statistics
 * scflow_conv_tc_fwd accumulated (sums[n][SCFLOW_STATS_REPLICAS][C][2], C <= 256), written as split planes for the next convolution
 * (models/backbone/resnet.py:64-92): y = act((x - mean) rstd + r), r = nothing | res_split (split planes,
 * final values) | (res_raw - rmean) rrstd (the down-sampling branch: raw fp32 + its own statistics).
 * Biased variance, as F.instance_norm.  out_f32 (n, P, C), optional: the same values in fp32. */
int scflow_norm_apply_split(const float* x, const double* sums, const void* res_split, long long res_pitch,
                            int res_lo, const float* res_raw, const double* res_sums, void* out,
                            long long out_pitch, int out_lo, float* out_f32, int n, int P, int C, float eps,
                            int relu, void* stream);


/* ------------------------------------------------------------------ backward (training, cfg5)
 * The reference gets these from autograd through ATen (grid_sampler_2d_backward,
 * avg_pool2d_backward, bmm backward, upsample_bilinear2d_backward and the elementwise chain of
 * models/utils/pose.py:66-88).  The host wraps each forward/backward pair in a
 * torch.autograd.Function (scflow_b200/autograd.py). */

/* Gradient of scflow_corr_lookup_fwd wrt the pyramid (corr_lookup.py:67, grid_sample `input`).
 * grad_out: (n, 324, h, w) in `grad_layout` (SCFLOW_LAYOUT_*); flow as in the forward;
 * grad_levels[l]: (n*h*w, 1, h>>l, w>>l), fully OVERWRITTEN (no memset needed).  The flow carries
 * no gradient on the SCFlow path (detach_flow=True, scflow_decoder.py:193-194). */
int scflow_corr_lookup_bwd(const float* grad_out, int grad_layout, const float* flow,
                           float* const* grad_levels, int num_levels, int n, int h, int w, int radius,
                           int div_mode, void* stream);

/* First half of the backward of scflow_corr_pyramid_fwd (raft_decoder.py:48-56): folds the level
 * gradients (NULL entries = no gradient) into grad_corr (n*h*w, h, w) = d loss / d (f1^T f2), i.e.
 * including the 1/sqrt(c) scale and the 2x2-mean adjoints.  scflow_corr_pyramid_bwd finishes the job. */
int scflow_corr_fold_bwd(const float* const* grad_levels, int num_levels, float* grad_corr, int n,
                         int c, int h, int w, void* stream);

/* Second half: the two batched GEMMs autograd runs behind `torch.matmul(feat1^T, feat2)`
 * (raft_decoder.py:53), d f1[c][q] = sum_t f2[c][t] G[q][t] and d f2[c][t] = sum_q f1[c][q] G[q][t], as one
 * tcgen05 kind::tf32 launch.  grad_corr (n, P, P) from scflow_corr_fold_bwd; feat1/feat2/grad_f1/grad_f2
 * (n, c, h, w) fp32 planar; either gradient pointer may be NULL (not wanted).  precision as in the forward:
 * SCFLOW_PREC_TF32 runs on the tensor cores when c % 128 == 0, h*w % 256 == 0 and all pointers are 32-byte
 * aligned (the shipped shapes); SCFLOW_PREC_FP32 and every other shape run the fp32 FMA kernel. */
int scflow_corr_pyramid_bwd(const float* grad_corr, const float* feat1, const float* feat2, float* grad_f1,
                            float* grad_f2, int n, int c, int h, int w, int precision, void* stream);

/* Adjoint of scflow_upsample_fwd: grad_in (n, c, h, w) = mul * resize^T(grad_out (n, c, h*scale,
 * w*scale)); the gradient is the same for `in` and `add`. */
int scflow_upsample_bwd(const float* grad_out, float* grad_in, int n, int c, int h, int w, int scale,
                        float mul, void* stream);

/* ConvGRU gate arithmetic of the training path (models/decoder/raft_decoder.py:245-252) around the gate
 * convolutions (which stay nn.Conv2d under autograd): rh = sigmoid(a_r) h and h' = (1 - z) h + z q with
 * z = sigmoid(a_z), q = tanh(a_q), forward and adjoint, one launch each.  Contiguous fp32 tensors of `total`
 * elements. */
int scflow_gru_reset_fwd(const float* a_r, const float* h, float* rh, long long total, void* stream);
int scflow_gru_reset_bwd(const float* grad_rh, const float* a_r, const float* h, float* grad_a_r, float* grad_h,
                         long long total, void* stream);
int scflow_gru_blend_fwd(const float* a_z, const float* a_q, const float* h, float* h_new, long long total, void* stream);
int scflow_gru_blend_bwd(const float* grad_h_new, const float* a_z, const float* a_q, const float* h, float* grad_a_z,
                         float* grad_a_q, float* grad_h, long long total, void* stream);

/* get_pose_from_delta_pose (models/utils/pose.py:124-149, ortho6d rotation delta, depth_transform = 'exp') for the
 * training path and its adjoint, one launch each (under autograd the same algebra is ~60 element-wise launches per
 * iteration).  dR6 (n, 6), dt (n, 3), R_src (n, 3, 3) row-major, t_src (n, 3) -> R_dst = dR R_src, t_dst.
 * Backward: grad_R (n, 3, 3) and / or grad_t (n, 3) (either may be NULL = zero) -> grad_dR6, grad_dt, and, where
 * the pointers are not NULL, grad_R_src / grad_t_src (the shipped config detaches the pose: NULL).
 * detach_depth_for_xy as the reference's flag (tz' does not carry the gradient of tx', ty'). */
int scflow_pose_compose_fwd(const float* dR6, const float* dt, const float* R_src, const float* t_src, float* R_dst,
                            float* t_dst, int n, float weight, void* stream);
int scflow_pose_compose_bwd(const float* grad_R, const float* grad_t, const float* dR6, const float* dt,
                            const float* R_src, const float* t_src, float* grad_dR6, float* grad_dt, float* grad_R_src,
                            float* grad_t_src, int n, float weight, int detach_depth_for_xy, void* stream);

/* Gradient of the pose-induced flow (scflow_pose_flow_fwd's flow_full, depth path) wrt the pose it
 * was rendered with: grad_pose (n, 12) = [dL/dR row-major | dL/dt], overwritten.  R, t are the
 * (n,3,3)/(n,3) pose of the forward call; depth/K/R0/t0 as in the forward. */
int scflow_pose_flow_bwd(const float* grad_full, const float* depth, const float* K, const float* R0,
                         const float* t0, const float* R, const float* t, float* grad_pose, int n, int H,
                         int W, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SCFLOW_B200_H_ */
