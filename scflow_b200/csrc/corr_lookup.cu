// CorrLookup forward: radius-4 bilinear window gather from the 4-level correlation pyramid.
// Replaces models/utils/corr_lookup.py:102-136 (4x F.grid_sample + ~25 glue launches).
//
// HBM-bound gather (P*L*T*4 out + P*L*100*4 in + 8*P flow per crop and iteration).
// One CTA owns QB = 32 consecutive query pixels of one crop:
//   phase 0  128 threads compute, per (query, level), the window origin with the reference's
//            exact fp32 op sequence (see unnorm_coord);
//   phase 1  all threads copy the (<=12x12, zero padded) window footprints into shared memory,
//            one element per thread per step: 72 independent loads in flight per thread;
//   phase 2  warp = (level, half of the x taps), lane = query: every tap is 4 conflict-free LDS
//            + 4 FMA and each (level, a, b) store is one coalesced 128-byte line of the
//            (n, 324, h, w) output.
// grid = n*h*w/32 CTAs; 74 KB smem -> 3 CTAs / SM (24 warps).
#include "common.cuh"

namespace scflow {

constexpr int kR = 4;
constexpr int kD = 2 * kR + 1;     // 9 taps per axis
constexpr int kFoot = 12;          // staged footprint per axis (10 + 1 slack either side)
constexpr int kWinStride = kFoot * kFoot + 1;  // 145: odd stride -> lanes hit distinct banks
constexpr int kQB = 32;            // queries per CTA
constexpr int kLevels = 4;
constexpr int kThreads = 256;

struct LookupParams {
  const float* lvl[kLevels];
  const float* flow;
  float* out;
  short* windows;
  int n, h, w, div_mode;
};

// Un-normalised sample coordinate of tap offset `off` along one axis of a level whose extent is
// `size`.  Each step is one separately rounded fp32 operation, exactly as the reference issues
// them:  corr_lookup.py:128 (centroid + delta), :64-65 (* 2. / max(size-1,1) - 1.), then ATen
// grid_sampler_unnormalize(align_corners=True): ((g + 1) / 2) * (size - 1).
__device__ __forceinline__ float unnorm_coord(float centre_lvl, int off, float denom, float rcp,
                                              float sizem1, int div_mode) {
  float c = __fadd_rn(centre_lvl, (float)off);
  float g = __fmul_rn(c, 2.0f);
  g = (div_mode == SCFLOW_DIV_IEEE) ? __fdiv_rn(g, denom) : __fmul_rn(g, rcp);
  g = __fsub_rn(g, 1.0f);
  float u = __fadd_rn(g, 1.0f);
  u = __fmul_rn(u, 0.5f);
  return __fmul_rn(u, sizem1);
}

__device__ __forceinline__ int floor_to_int_clamped(float v, int lo, int hi) {
  float f = floorf(v);
  f = fminf(fmaxf(f, (float)lo), (float)hi);  // NaN -> lo
  return (int)f;
}

__global__ void __launch_bounds__(kThreads, 3)
corr_lookup_kernel(const LookupParams p) {
  extern __shared__ float smem[];
  float* win = smem;                                     // [kQB][kLevels][kWinStride]
  int* base = reinterpret_cast<int*>(smem + kQB * kLevels * kWinStride);  // [kQB*kLevels][2]

  const int P = p.h * p.w;
  const int tid = threadIdx.x;
  const long long q0 = (long long)blockIdx.x * kQB;      // flat query index over n*P
  const int n = (int)(q0 / P);
  const int qin0 = (int)(q0 - (long long)n * P);         // query index inside the crop

  // ---- phase 0: window origins
  if (tid < kQB * kLevels) {
    const int ql = tid & (kQB - 1), l = tid >> 5;
    const int q = qin0 + ql;
    const int y = q / p.w, x = q - y * p.w;
    const float fx = p.flow[((long long)n * 2 + 0) * P + q];
    const float fy = p.flow[((long long)n * 2 + 1) * P + q];
    const float sc = 1.0f / (float)(1 << l);
    const float cx = __fmul_rn(__fadd_rn((float)x, fx), sc);   // grid / 2**i  (exact scaling)
    const float cy = __fmul_rn(__fadd_rn((float)y, fy), sc);
    const int wl = p.w >> l, hl = p.h >> l;
    const float dx = (float)max(wl - 1, 1), dy = (float)max(hl - 1, 1);
    const float ix0 = unnorm_coord(cx, -kR, dx, __fdiv_rn(1.0f, dx), (float)(wl - 1), p.div_mode);
    const float iy0 = unnorm_coord(cy, -kR, dy, __fdiv_rn(1.0f, dy), (float)(hl - 1), p.div_mode);
    base[2 * tid + 0] = floor_to_int_clamped(ix0, -32, wl + 32) - 1;
    base[2 * tid + 1] = floor_to_int_clamped(iy0, -32, hl + 32) - 1;
  }
  __syncthreads();

  // ---- phase 1: stage footprints (zero outside the level)
  {
    constexpr int kElems = kQB * kLevels * kFoot * kFoot;
    for (int e = tid; e < kElems; e += kThreads) {
      const int col = e % kFoot;
      const int row = (e / kFoot) % kFoot;
      const int ql = (e / (kFoot * kFoot)) % kQB;
      const int l = e / (kFoot * kFoot * kQB);
      const int wl = p.w >> l, hl = p.h >> l;
      const int bx = base[2 * (l * kQB + ql) + 0], by = base[2 * (l * kQB + ql) + 1];
      const int gx = bx + col, gy = by + row;
      float v = 0.0f;
      if (gx >= 0 && gx < wl && gy >= 0 && gy < hl) {
        const float* src = p.lvl[l] + (q0 + ql) * (long long)(hl * wl);
        v = __ldg(src + gy * wl + gx);
      }
      win[(ql * kLevels + l) * kWinStride + row * kFoot + col] = v;
    }
  }
  __syncthreads();

  // ---- phase 2: interpolate; warp = (level, half), lane = query
  const int warp = tid >> 5, lane = tid & 31;
  const int l = warp >> 1, half = warp & 1;
  const int q = qin0 + lane;
  const int y = q / p.w, x = q - y * p.w;
  const float fx = p.flow[((long long)n * 2 + 0) * P + q];
  const float fy = p.flow[((long long)n * 2 + 1) * P + q];
  const float sc = 1.0f / (float)(1 << l);
  const float cx = __fmul_rn(__fadd_rn((float)x, fx), sc);
  const float cy = __fmul_rn(__fadd_rn((float)y, fy), sc);
  const int wl = p.w >> l, hl = p.h >> l;
  const float dx = (float)max(wl - 1, 1), dy = (float)max(hl - 1, 1);
  const float rx = __fdiv_rn(1.0f, dx), ry = __fdiv_rn(1.0f, dy);
  const int bx = base[2 * (l * kQB + lane) + 0], by = base[2 * (l * kQB + lane) + 1];
  const float* mywin = win + (lane * kLevels + l) * kWinStride;

  int rely[kD];
  float wy0[kD], wy1[kD];
#pragma unroll
  for (int b = 0; b < kD; ++b) {
    const float iy = unnorm_coord(cy, b - kR, dy, ry, (float)(hl - 1), p.div_mode);
    const float y0f = floorf(iy);
    const int y0 = floor_to_int_clamped(iy, -32, hl + 32);
    rely[b] = min(max(y0 - by, 0), kFoot - 2) * kFoot;
    wy1[b] = __fsub_rn(iy, y0f);                     // ATen: iy - iy_nw
    wy0[b] = __fsub_rn(__fadd_rn(y0f, 1.0f), iy);    // ATen: iy_se - iy
    if (p.windows != nullptr && half == 0)
      p.windows[(((q0 + lane) * kLevels + l) * 2 + 1) * kD + b] = (short)y0;
  }

  const int a_begin = half ? 5 : 0, a_end = half ? kD : 5;
  float* outq = p.out + ((long long)n * (kLevels * kD * kD) + l * (kD * kD)) * P + q;
  for (int a = a_begin; a < a_end; ++a) {
    const float ix = unnorm_coord(cx, a - kR, dx, rx, (float)(wl - 1), p.div_mode);
    const float x0f = floorf(ix);
    const int x0 = floor_to_int_clamped(ix, -32, wl + 32);
    const int relx = min(max(x0 - bx, 0), kFoot - 2);
    const float wx1 = __fsub_rn(ix, x0f);
    const float wx0 = __fsub_rn(__fadd_rn(x0f, 1.0f), ix);
    if (p.windows != nullptr)
      p.windows[(((q0 + lane) * kLevels + l) * 2 + 0) * kD + a] = (short)x0;
#pragma unroll
    for (int b = 0; b < kD; ++b) {
      const float* s = mywin + rely[b] + relx;
      const float v00 = s[0], v01 = s[1], v10 = s[kFoot], v11 = s[kFoot + 1];
      // ATen grid_sampler_2d: nw, ne, sw, se accumulated in this order
      float acc = v00 * (wx0 * wy0[b]);
      acc = fmaf(v01, wx1 * wy0[b], acc);
      acc = fmaf(v10, wx0 * wy1[b], acc);
      acc = fmaf(v11, wx1 * wy1[b], acc);
      outq[(long long)(a * kD + b) * P] = acc;
    }
  }
}

}  // namespace scflow

namespace scflow {
int launch_corr_lookup_nhwc(const float* const* levels, const float* flow, float* out, short* windows,
                            int n, int h, int w, int div_mode, bool gen1, bool split, cudaStream_t stream);
}

extern "C" int scflow_corr_lookup_fwd(const float* const* levels, int num_levels,
                                      const float* flow, float* out, short* windows, int n,
                                      int h, int w, int radius, int div_mode, int out_layout,
                                      void* stream) {
  using namespace scflow;
  if (!levels || !flow || !out) return SCFLOW_E_NULL;
  if (radius != kR || num_levels != kLevels) return SCFLOW_E_UNSUPPORTED;
  if (div_mode != SCFLOW_DIV_IEEE && div_mode != SCFLOW_DIV_RCP) return SCFLOW_E_UNSUPPORTED;
  if (out_layout != SCFLOW_LAYOUT_NCHW && out_layout != SCFLOW_LAYOUT_NHWC && out_layout != SCFLOW_LAYOUT_NHWC_GEN1 &&
      out_layout != SCFLOW_LAYOUT_NHWC_SPLIT)
    return SCFLOW_E_UNSUPPORTED;
  if (out_layout != SCFLOW_LAYOUT_NCHW) {
    if (n <= 0 || h < 8 || w < 8) return SCFLOW_E_BAD_SHAPE;
    for (int l = 0; l < kLevels; ++l)
      if (!levels[l]) return SCFLOW_E_NULL;
    return launch_corr_lookup_nhwc(levels, flow, out, windows, n, h, w, div_mode,
                                   out_layout == SCFLOW_LAYOUT_NHWC_GEN1, out_layout == SCFLOW_LAYOUT_NHWC_SPLIT,
                                   (cudaStream_t)stream);
  }
  if (n <= 0 || h < 8 || w < 8 || ((h * w) % kQB) != 0) return SCFLOW_E_BAD_SHAPE;
  LookupParams p;
  for (int l = 0; l < kLevels; ++l) {
    if (!levels[l]) return SCFLOW_E_NULL;
    p.lvl[l] = levels[l];
  }
  p.flow = flow; p.out = out; p.windows = windows;
  p.n = n; p.h = h; p.w = w; p.div_mode = div_mode;
  const size_t smem = (size_t)kQB * kLevels * kWinStride * sizeof(float) +
                      (size_t)kQB * kLevels * 2 * sizeof(int);
  cudaFuncSetAttribute(corr_lookup_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);   // per device
  const long long blocks = (long long)n * h * w / kQB;
  corr_lookup_kernel<<<(unsigned)blocks, kThreads, smem, (cudaStream_t)stream>>>(p);
  return check_launch();
}
