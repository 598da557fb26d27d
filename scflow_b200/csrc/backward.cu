// Backward kernels of the hot path (BASELINE.json configs[4]: training step forward+backward).
//
//   corr_lookup_bwd     adjoint of the radius-4 bilinear window gather wrt the pyramid
//                       (grid_sampler_2d_backward of corr_lookup.py:67 for the `input` argument)
//   corr_fold_bwd       folds the four pyramid-level gradients into one gradient of the scaled
//                       correlation volume (adjoint of 3x avg_pool2d + the 1/sqrt(C) scale,
//                       raft_decoder.py:48-56); the two remaining contractions are plain GEMMs
//   resize_bwd          adjoint of the align_corners=True bilinear up-sampling (scflow_decoder.py
//                       :223-228): gradient wrt the low-resolution flow / mask
//   pose_flow_bwd       gradient of the pose-induced flow (pose.py:66-88) wrt R_k and t_k
//
// All HBM-bound.  Every query / crop owns its output exclusively, so no kernel here needs global
// atomics except the 12-value-per-crop reduction of pose_flow_bwd.
#include "common.cuh"
#include <cstddef>

namespace scflow {
namespace bw {

constexpr int kR = 4, kD = 9, kTaps = 81, kLevels = 4;

// Same separately rounded fp32 op sequence as the forward (corr_lookup.cu: unnorm_coord), so the
// backward scatters into exactly the cells the forward read.
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

__device__ __forceinline__ int floor_clamped(float v, int lo, int hi) {
  float f = floorf(v);
  f = fminf(fmaxf(f, (float)lo), (float)hi);  // NaN -> lo
  return (int)f;
}

struct LookupBwdParams {
  const float* grad_out;      // (n, 324, h, w) with strides gs_q (query) / gs_c (channel)
  const float* flow;          // (n, 2, h, w)
  float* grad_lvl[kLevels];   // (n*h*w, 1, h>>l, w>>l), overwritten
  long long gs_q, gs_c;
  int n, h, w, div_mode;
  int plane_off[kLevels + 1];  // float offsets of the level planes inside a warp's staging area
};

constexpr int kLBWarps = 4;

// One warp per query: the four gradient planes of the query are assembled in shared memory (zero
// + 4 x 81 x 4 weighted corner adds) and streamed out whole, so the pyramid gradient is written
// exactly once with full-line stores and needs no memset.
__global__ void __launch_bounds__(kLBWarps * 32)
corr_lookup_bwd_kernel(const LookupBwdParams p) {
  extern __shared__ float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int plane_floats = p.plane_off[kLevels];
  float* pl = smem + (size_t)warp * plane_floats;
  const int P = p.h * p.w;
  const long long total_q = (long long)p.n * P;
  for (long long q = (long long)blockIdx.x * kLBWarps + warp; q < total_q; q += (long long)gridDim.x * kLBWarps) {
    for (int i = lane; i < plane_floats; i += 32) pl[i] = 0.0f;
    __syncwarp();
    const int n = (int)(q / P), qin = (int)(q - (long long)n * P);
    const int y = qin / p.w, x = qin - y * p.w;
    const float fx = __ldg(p.flow + ((long long)n * 2 + 0) * P + qin);
    const float fy = __ldg(p.flow + ((long long)n * 2 + 1) * P + qin);
    const float* g = p.grad_out + (long long)n * (kLevels * kTaps) * P + (long long)qin * p.gs_q;
#pragma unroll
    for (int l = 0; l < kLevels; ++l) {
      const float sc = 1.0f / (float)(1 << l);
      const float cx = __fmul_rn(__fadd_rn((float)x, fx), sc);
      const float cy = __fmul_rn(__fadd_rn((float)y, fy), sc);
      const int wl = p.w >> l, hl = p.h >> l;
      const float dx = (float)max(wl - 1, 1), dy = (float)max(hl - 1, 1);
      const float rx = __fdiv_rn(1.0f, dx), ry = __fdiv_rn(1.0f, dy);
      float* plane = pl + p.plane_off[l];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int t = lane + 32 * k;
        if (t < kTaps) {
          const int a = t / kD, b = t - a * kD;       // channel 81*l + 9*a + b: a offsets x, b offsets y
          const float gv = __ldg(g + (long long)(l * kTaps + t) * p.gs_c);
          const float ix = unnorm_coord(cx, a - kR, dx, rx, (float)(wl - 1), p.div_mode);
          const float iy = unnorm_coord(cy, b - kR, dy, ry, (float)(hl - 1), p.div_mode);
          const float x0f = floorf(ix), y0f = floorf(iy);
          const int x0 = floor_clamped(ix, -32, wl + 32), y0 = floor_clamped(iy, -32, hl + 32);
          const float wx1 = __fsub_rn(ix, x0f), wx0 = __fsub_rn(__fadd_rn(x0f, 1.0f), ix);
          const float wy1 = __fsub_rn(iy, y0f), wy0 = __fsub_rn(__fadd_rn(y0f, 1.0f), iy);
          const bool xin0 = x0 >= 0 && x0 < wl, xin1 = x0 + 1 >= 0 && x0 + 1 < wl;
          const bool yin0 = y0 >= 0 && y0 < hl, yin1 = y0 + 1 >= 0 && y0 + 1 < hl;
          if (yin0 && xin0) atomicAdd(plane + y0 * wl + x0, gv * (wx0 * wy0));
          if (yin0 && xin1) atomicAdd(plane + y0 * wl + x0 + 1, gv * (wx1 * wy0));
          if (yin1 && xin0) atomicAdd(plane + (y0 + 1) * wl + x0, gv * (wx0 * wy1));
          if (yin1 && xin1) atomicAdd(plane + (y0 + 1) * wl + x0 + 1, gv * (wx1 * wy1));
        }
      }
    }
    __syncwarp();
#pragma unroll
    for (int l = 0; l < kLevels; ++l) {
      const int cnt = (p.h >> l) * (p.w >> l);
      const float* src = pl + p.plane_off[l];
      float* dst = p.grad_lvl[l] + q * cnt;
      if ((cnt & 3) == 0) {
        for (int i = lane; i < cnt / 4; i += 32)
          reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src)[i];
      } else {
        for (int i = lane; i < cnt; i += 32) dst[i] = src[i];
      }
    }
    __syncwarp();
  }
}


// ---- shipped shapes (w % 32 == 0, h % 8 == 0): no atomics, no whole-plane staging ------------------------------
// fp32 shared-memory atomicAdd compiles to a compare-and-swap loop (ATOMS.CAST.SPIN), and that loop was the
// kernel above.  The gradient of a query on a level is nonzero only inside the window the forward read, and it is
// separable there: tap (a, b) adds g[a][b] wx(a, X) wy(b, Y) to its 2 x 2 cells, so
//     T[a][Y] = sum_b g[a][b] wy(b, Y)       G[Y][X] = sum_a T[a][Y] wx(a, X)
// with every output owned by one lane.  x0(a) and the weights come from the forward's exact op sequence per tap;
// relative to tap 0, x0(a) = x0(0) + a + e with e in {-1, 0, 1} (floors next to integer coordinates, e.g. at
// flow = 0, are irregular by an ulp), so tap a can only reach the cells x0(0) + a - 1 .. x0(0) + a + 2 and a cell
// only the taps a - 2 .. a + 1: four products per output, whatever the floors did.  The 1-D weights are stored
// skewed (us[k][cell] = weight of tap cell + 1 - k) in plane-aligned columns, the window (12 rows x 16 columns)
// starts on a 16-byte chunk of the plane and is written straight to global memory over a zero fill of the four
// planes (the overwrite contract; __syncwarp orders the two stores of a warp).  A warp stages the gradients of four
// consecutive queries through shared memory (16 bytes per channel row of a planar gradient) on its own.
constexpr int kWQ = 4, kGP = kLevels * kTaps + 5, kGuard = 4, kWinWarps = 8;   // 5 zero pads per query, 4 in front:
                                                                                 // phase B reads up to 3 taps outside a row
struct WinScratch {
  float usx[kLevels][4][16];    // x weights, skewed: [k][column of the window]
  float usy[kLevels][4][12];    // y weights: [k][row of the window]
  float t[kLevels][kD][12];     // T[a][row]
  int xb[kLevels], yb[kLevels], xoff[kLevels], pad[4];
};
static_assert(sizeof(WinScratch) % 16 == 0 && offsetof(WinScratch, usy) % 16 == 0 && offsetof(WinScratch, t) % 16 == 0,
              "scratch rows are read as float4");
struct WinWarp {
  float guard[kGuard];
  float g[kWQ * kGP];           // the gradients of the warp's queries, [query][324 + 5 zero pads]
  WinScratch sc;
};
static_assert(sizeof(WinWarp) % 16 == 0 && offsetof(WinWarp, sc) % 16 == 0, "scratch alignment");
constexpr size_t kWinSmem = kWinWarps * sizeof(WinWarp);

struct WinParams {
  const float* grad_out;
  const float* flow;
  float* grad_lvl[kLevels];
  int n, h, w, div_mode, nhwc;
  int chunks[kLevels];          // float4 chunks per level plane
  float rcp[2][kLevels];        // 1 / max(size - 1, 1) per axis and level (SCFLOW_DIV_RCP)
  int ctask[kLevels + 1];       // phase C: prefix sums of min(48 window chunks, plane chunks)
  float inv_wl4[kLevels];       // 1 / (chunks per plane row)
};

__device__ __forceinline__ float pick4(const float4& v, int t) { return t == 0 ? v.x : t == 1 ? v.y : t == 2 ? v.z : v.w; }

__global__ void __launch_bounds__(kWinWarps * 32, 3)
corr_lookup_bwd_win_kernel(const WinParams p) {
  extern __shared__ float4 smem4[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WinWarp& ws = reinterpret_cast<WinWarp*>(smem4)[warp];
  WinScratch& sc = ws.sc;
  float* gt = ws.g;
  constexpr int C = kLevels * kTaps;
  if (lane < 4) ws.guard[lane] = 0.f;
  for (int i = lane; i < kWQ * (kGP - C); i += 32) gt[(i / (kGP - C)) * kGP + C + i % (kGP - C)] = 0.f;
  const int P = p.h * p.w;
  const long long wtiles = ((long long)p.n * P) / kWQ;
  // a warp owns kWQ consecutive queries at a time: it stages their gradients itself (no block-wide barrier, so
  // the warps of an SM drift apart and one warp's loads overlap another's arithmetic)
  for (long long wt = (long long)blockIdx.x * kWinWarps + warp; wt < wtiles; wt += (long long)gridDim.x * kWinWarps) {
    const long long q0 = wt * kWQ;
    const int n = (int)(q0 / P), qin0 = (int)(q0 - (long long)n * P);
    __syncwarp();
    float fl = 0.f;                                 // lanes 0..3: fx of the four queries, 4..7: fy
    if (lane < 2 * kWQ) fl = __ldg(p.flow + ((long long)n * 2 + (lane >> 2)) * P + qin0 + (lane & 3));
    if (p.nhwc) {                                   // 4 x 324 contiguous floats
      const float4* src = reinterpret_cast<const float4*>(p.grad_out + q0 * C);
#pragma unroll 4
      for (int i = lane; i < kWQ * C / 4; i += 32) {
        const float4 v = __ldg(src + i);
        const int q = (4 * i) / C, ch = 4 * i - q * C;       // 324 % 4 == 0: a chunk never straddles two queries
        float* d = gt + q * kGP + ch;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
      }
    } else {                                        // 324 rows, 4 consecutive queries (16 bytes) of each
      const float* src = p.grad_out + (long long)n * C * P + qin0;
#pragma unroll 4
      for (int ch = lane; ch < C; ch += 32) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(src + (long long)ch * P));
        gt[ch] = v.x; gt[kGP + ch] = v.y; gt[2 * kGP + ch] = v.z; gt[3 * kGP + ch] = v.w;
      }
    }
    __syncwarp();
#pragma unroll 1
    for (int qi = 0; qi < kWQ; ++qi) {
      const int qin = qin0 + qi;
      const long long q = q0 + qi;
      const int y = qin / p.w, x = qin - y * p.w;
      const float fx = __shfl_sync(0xffffffffu, fl, qi), fy = __shfl_sync(0xffffffffu, fl, kWQ + qi);
      const float* g = gt + qi * kGP;
      // zeros: the weight tables of this query and the four planes
      {
        float4* z = reinterpret_cast<float4*>(sc.usx);
        for (int i = lane; i < (int)(offsetof(WinScratch, t) / 16); i += 32) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int l = 0; l < kLevels; ++l) {
        float4* dst = reinterpret_cast<float4*>(p.grad_lvl[l] + q * (long long)((p.h >> l) * (p.w >> l)));
        for (int c = lane; c < p.chunks[l]; c += 32) dst[c] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      __syncwarp();
      // A: one (level, axis, tap) per lane, three 9-lane groups per round: the tap's cell and weights by the forward's
      //    op sequence, placed into the skewed tables; tap 0 (shuffled across its group) fixes the window origin
#pragma unroll
      for (int rnd = 0; rnd < 3; ++rnd) {
        const int sub = lane / kD, idx = lane - sub * kD;
        const int grp = min(rnd * 3 + sub, kLevels * 2 - 1), l = grp >> 1, dir = grp & 1;
        const bool live = lane < 3 * kD && rnd * 3 + sub < kLevels * 2;
        const float scl = 1.0f / (float)(1 << l);
        const int size = (dir ? p.h : p.w) >> l;
        const float centre = __fmul_rn(__fadd_rn(dir ? (float)y : (float)x, dir ? fy : fx), scl);
        const float ic = unnorm_coord(centre, idx - kR, (float)max(size - 1, 1), p.rcp[dir][l], (float)(size - 1), p.div_mode);
        const float c0f = floorf(ic);
        const int c0 = floor_clamped(ic, -32, size + 32);
        const int d0 = __shfl_sync(0xffffffffu, c0, lane - idx);
        const float w1 = __fsub_rn(ic, c0f), w0 = __fsub_rn(__fadd_rn(c0f, 1.0f), ic);
        const int base = dir ? d0 - 1 : ((d0 - 1) >> 2) << 2;   // first row / first (chunk-aligned) column
        const int off = d0 - base;                              // 1 (rows), 1..4 (columns)
        const int e = c0 - d0 - idx;                            // -1, 0, 1 for every tap that can touch the plane
        const int k0 = e + 1, pos = idx - 1 + off;              // us[k][pos + k] belongs to tap idx
        float* tab = dir ? &sc.usy[l][0][0] : &sc.usx[l][0][0];
        const int pitch = dir ? 12 : 16;
        if (live && k0 >= 0 && k0 <= 3) tab[k0 * pitch + pos + k0] = w0;
        if (live && k0 + 1 >= 0 && k0 + 1 <= 3) tab[(k0 + 1) * pitch + pos + k0 + 1] = w1;
        if (live && idx == 0) {
          if (dir) sc.yb[l] = base; else { sc.xb[l] = base; sc.xoff[l] = off; }
        }
      }
      __syncwarp();
      // B: T[l][a][r] = sum_k g[l][a][r - k] usy[l][k][r]     (r = window row; 108 tasks of 4 rows)
#pragma unroll
      for (int task = lane; task < kLevels * kD * 3; task += 32) {
        const int l = task / (kD * 3), r = task - l * kD * 3, a = r / 3, rg = r - a * 3;
        const float* gp = g + l * kTaps + a * kD;
        float gv[7];
#pragma unroll
        for (int m = 0; m < 7; ++m) gv[m] = gp[4 * rg - 3 + m];    // taps outside 0..8: a neighbour or a zero pad, weight 0
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float4 u = reinterpret_cast<const float4*>(sc.usy[l][k])[rg];
          acc.x = fmaf(gv[3 - k], u.x, acc.x); acc.y = fmaf(gv[4 - k], u.y, acc.y);
          acc.z = fmaf(gv[5 - k], u.z, acc.z); acc.w = fmaf(gv[6 - k], u.w, acc.w);
        }
        reinterpret_cast<float4*>(sc.t[l][a])[rg] = acc;
      }
      __syncwarp();
      // C: G[l][r][col] = sum_k T[l][col - xoff + 1 - k][r] usx[l][k][col], one 16-byte chunk of the plane per task:
      //    the 48 chunks of the window, or the chunks of the plane where the plane is the smaller of the two
      for (int task = lane; task < p.ctask[kLevels]; task += 32) {
        const int l = (task >= p.ctask[1]) + (task >= p.ctask[2]) + (task >= p.ctask[3]);
        const int ci = task - p.ctask[l];
        const int wl = p.w >> l, hl = p.h >> l;
        int j, ch;
        if (p.ctask[l + 1] - p.ctask[l] == 48) { j = ci >> 2; ch = ci & 3; }
        else {
          const int Yp = __float2int_rz(((float)ci + 0.5f) * p.inv_wl4[l]), Xp = (ci - Yp * (wl >> 2)) * 4;
          j = Yp - sc.yb[l];
          ch = (Xp - sc.xb[l]) >> 2;
          if (j < 0 || j >= 12 || ch < 0 || ch >= 4) continue;
        }
        const int Y = sc.yb[l] + j, X4 = sc.xb[l] + 4 * ch;
        if (Y < 0 || Y >= hl || X4 < 0 || X4 >= wl) continue;
        const int a0 = 4 * ch - sc.xoff[l] - 2;                 // tap of (cell 0, k = 3)
        float tv[7];
#pragma unroll
        for (int m = 0; m < 7; ++m) tv[m] = sc.t[l][min(max(a0 + m, 0), kD - 1)][j];
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float4 u = reinterpret_cast<const float4*>(sc.usx[l][k])[ch];
          acc.x = fmaf(tv[3 - k], u.x, acc.x); acc.y = fmaf(tv[4 - k], u.y, acc.y);
          acc.z = fmaf(tv[5 - k], u.z, acc.z); acc.w = fmaf(tv[6 - k], u.w, acc.w);
        }
        *reinterpret_cast<float4*>(p.grad_lvl[l] + q * (long long)(hl * wl) + Y * wl + X4) = acc;
      }
      __syncwarp();
    }
  }
}

}  // namespace bw
}  // namespace scflow

extern "C" int scflow_corr_lookup_bwd(const float* grad_out, int grad_layout, const float* flow,
                                      float* const* grad_levels, int num_levels, int n, int h, int w,
                                      int radius, int div_mode, void* stream) {
  using namespace scflow;
  if (!grad_out || !flow || !grad_levels) return SCFLOW_E_NULL;
  if (radius != bw::kR || num_levels != bw::kLevels) return SCFLOW_E_UNSUPPORTED;
  if (div_mode != SCFLOW_DIV_IEEE && div_mode != SCFLOW_DIV_RCP) return SCFLOW_E_UNSUPPORTED;
  if (grad_layout != SCFLOW_LAYOUT_NCHW && grad_layout != SCFLOW_LAYOUT_NHWC) return SCFLOW_E_UNSUPPORTED;
  if (n <= 0 || h < 8 || w < 8) return SCFLOW_E_BAD_SHAPE;
  for (int l = 0; l < bw::kLevels; ++l) {
    if (!grad_levels[l]) return SCFLOW_E_NULL;
    if (!aligned16(grad_levels[l])) return SCFLOW_E_MISALIGNED;
  }
  if ((w % 32) == 0 && (h % 8) == 0 && aligned16(grad_out)) {
    bw::WinParams wp;
    wp.grad_out = grad_out; wp.flow = flow;
    wp.n = n; wp.h = h; wp.w = w; wp.div_mode = div_mode;
    wp.nhwc = grad_layout == SCFLOW_LAYOUT_NHWC;
    wp.ctask[0] = 0;
    for (int l = 0; l < bw::kLevels; ++l) {
      wp.grad_lvl[l] = grad_levels[l];
      wp.chunks[l] = ((h >> l) * (w >> l)) / 4;
      wp.rcp[0][l] = 1.0f / (float)((w >> l) - 1 > 1 ? (w >> l) - 1 : 1);
      wp.rcp[1][l] = 1.0f / (float)((h >> l) - 1 > 1 ? (h >> l) - 1 : 1);
      wp.ctask[l + 1] = wp.ctask[l] + (wp.chunks[l] < 48 ? wp.chunks[l] : 48);
      wp.inv_wl4[l] = 1.0f / (float)((w >> l) >> 2);
    }
    cudaFuncSetAttribute(bw::corr_lookup_bwd_win_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)bw::kWinSmem);
    const long long tiles = ((long long)n * h * w / bw::kWQ + bw::kWinWarps - 1) / bw::kWinWarps;
    const long long blocks = tiles < 3 * 148 ? tiles : 3 * 148;
    bw::corr_lookup_bwd_win_kernel<<<(unsigned)blocks, bw::kWinWarps * 32, bw::kWinSmem, (cudaStream_t)stream>>>(wp);
    return check_launch();
  }
  bw::LookupBwdParams p;
  p.grad_out = grad_out; p.flow = flow;
  p.n = n; p.h = h; p.w = w; p.div_mode = div_mode;
  const long long P = (long long)h * w;
  p.gs_q = grad_layout == SCFLOW_LAYOUT_NHWC ? bw::kLevels * bw::kTaps : 1;
  p.gs_c = grad_layout == SCFLOW_LAYOUT_NHWC ? 1 : P;
  int off = 0;
  for (int l = 0; l < bw::kLevels; ++l) {
    p.grad_lvl[l] = grad_levels[l];
    p.plane_off[l] = off;
    off += (((h >> l) * (w >> l)) + 3) & ~3;      // 16-byte aligned planes in the staging area
  }
  p.plane_off[bw::kLevels] = off;
  const size_t smem = (size_t)bw::kLBWarps * off * sizeof(float);
  if (smem > 200 * 1024) return SCFLOW_E_BAD_SHAPE;
  cudaFuncSetAttribute(bw::corr_lookup_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);   // per device
  const long long total_q = (long long)n * P;
  long long blocks = (total_q + bw::kLBWarps - 1) / bw::kLBWarps;
  const long long cap = 148LL * (smem > 64 * 1024 ? 2 : 8);
  if (blocks > cap) blocks = cap;
  bw::corr_lookup_bwd_kernel<<<(unsigned)blocks, bw::kLBWarps * 32, smem, (cudaStream_t)stream>>>(p);
  return check_launch();
}

// ---------------------------------------------------------------------------------------------
// corr_fold_bwd: G[q, y, x] = (g0[q,y,x] + g1[q,y/2,x/2]/4 + g2[q,y/4,x/4]/16 + g3[q,y/8,x/8]/64) / sqrt(C)
// One thread = 4 consecutive x of one row (w % 4 == 0); absent levels are skipped.
namespace scflow {
namespace bw {

struct FoldParams {
  const float* g[kLevels];
  float* out;
  long long rows;          // n*P*h rows of w floats
  int h, w, num_levels;
  float inv_scale;
};

__global__ void __launch_bounds__(256)
corr_fold_bwd_kernel(const FoldParams p) {
  const int qw = p.w >> 2;
  const long long total = p.rows * qw;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / qw;
    const int x4 = (int)(i - row * qw) * 4;
    const long long q = row / p.h;
    const int y = (int)(row - q * p.h);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.g[0]) v = __ldg(reinterpret_cast<const float4*>(p.g[0] + row * p.w + x4));
    if (p.num_levels > 1 && p.g[1]) {
      const int w1 = p.w >> 1, h1 = p.h >> 1;
      const float* s = p.g[1] + (q * h1 + (y >> 1)) * w1 + (x4 >> 1);
      const float a = 0.25f * __ldg(s), b = 0.25f * __ldg(s + 1);
      v.x += a; v.y += a; v.z += b; v.w += b;
    }
    if (p.num_levels > 2 && p.g[2]) {
      const int w2 = p.w >> 2, h2 = p.h >> 2;
      const float a = 0.0625f * __ldg(p.g[2] + (q * h2 + (y >> 2)) * w2 + (x4 >> 2));
      v.x += a; v.y += a; v.z += a; v.w += a;
    }
    if (p.num_levels > 3 && p.g[3]) {
      const int w3 = p.w >> 3, h3 = p.h >> 3;
      const float a = 0.015625f * __ldg(p.g[3] + (q * h3 + (y >> 3)) * w3 + (x4 >> 3));
      v.x += a; v.y += a; v.z += a; v.w += a;
    }
    v.x *= p.inv_scale; v.y *= p.inv_scale; v.z *= p.inv_scale; v.w *= p.inv_scale;
    *reinterpret_cast<float4*>(p.out + row * p.w + x4) = v;
  }
}

}  // namespace bw
}  // namespace scflow

extern "C" int scflow_corr_fold_bwd(const float* const* grad_levels, int num_levels, float* grad_corr,
                                    int n, int c, int h, int w, void* stream) {
  using namespace scflow;
  if (!grad_levels || !grad_corr) return SCFLOW_E_NULL;
  if (num_levels < 1 || num_levels > bw::kLevels) return SCFLOW_E_UNSUPPORTED;
  if (n <= 0 || c <= 0 || h <= 0 || w <= 0 || (w & 3) != 0 || (h % (1 << (num_levels - 1))) != 0 ||
      (w % (1 << (num_levels - 1))) != 0)
    return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(grad_corr)) return SCFLOW_E_MISALIGNED;
  bw::FoldParams p;
  for (int l = 0; l < bw::kLevels; ++l) p.g[l] = l < num_levels ? grad_levels[l] : nullptr;
  if (p.g[0] && !aligned16(p.g[0])) return SCFLOW_E_MISALIGNED;
  p.out = grad_corr;
  p.rows = (long long)n * h * w * h;
  p.h = h; p.w = w; p.num_levels = num_levels;
  p.inv_scale = 1.0f / sqrtf((float)c);
  long long blocks = (p.rows * (w >> 2) + 255) / 256;
  if (blocks > 148LL * 16) blocks = 148LL * 16;
  bw::corr_fold_bwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(p);
  return check_launch();
}

// ---------------------------------------------------------------------------------------------
// resize_bwd: adjoint of out = mul * bilinear_align_corners(in [+ add]) for an integer up-scale.
// One CTA per (plane, low-res row): the <= 2*scale+2 contributing high-res rows are first reduced
// along y into a shared row (coalesced reads), then along x.  Weights are recomputed with the
// forward's expressions (pose_flow.cu: resize_kernel), so the pair is an exact adjoint.
namespace scflow {
namespace bw {

__global__ void __launch_bounds__(256)
resize_bwd_kernel(const float* __restrict__ grad_out, float* __restrict__ grad_in, int ih, int iw,
                  int oh, int ow, float mul) {
  extern __shared__ float col[];                    // [ow]
  const int y = blockIdx.x, pl = blockIdx.y;
  const float sy = oh > 1 ? (float)(ih - 1) / (float)(oh - 1) : 0.0f;
  const float sx = ow > 1 ? (float)(iw - 1) / (float)(ow - 1) : 0.0f;
  const float* g = grad_out + (size_t)pl * oh * ow;
  // candidate high-res rows: src = sy*Y lies in (y-1, y+1)
  int Ylo = 0, Yhi = oh - 1;
  if (sy > 0.0f) {
    Ylo = max(0, (int)floorf((float)(y - 1) / sy) - 1);
    Yhi = min(oh - 1, (int)ceilf((float)(y + 1) / sy) + 1);
  }
  for (int X = threadIdx.x; X < ow; X += blockDim.x) {
    float acc = 0.0f;
    for (int Y = Ylo; Y <= Yhi; ++Y) {
      const float srcy = sy * (float)Y;
      const int y0 = (int)srcy;
      const int y1 = y0 + (y0 < ih - 1 ? 1 : 0);
      const float ly1 = srcy - (float)y0, ly0 = 1.0f - ly1;
      float wgt = 0.0f;
      if (y0 == y) wgt += ly0;
      if (y1 == y) wgt += ly1;
      if (wgt != 0.0f) acc = fmaf(wgt, __ldg(g + (size_t)Y * ow + X), acc);
    }
    col[X] = acc;
  }
  __syncthreads();
  for (int x = threadIdx.x; x < iw; x += blockDim.x) {
    int Xlo = 0, Xhi = ow - 1;
    if (sx > 0.0f) {
      Xlo = max(0, (int)floorf((float)(x - 1) / sx) - 1);
      Xhi = min(ow - 1, (int)ceilf((float)(x + 1) / sx) + 1);
    }
    float acc = 0.0f;
    for (int X = Xlo; X <= Xhi; ++X) {
      const float srcx = sx * (float)X;
      const int x0 = (int)srcx;
      const int x1 = x0 + (x0 < iw - 1 ? 1 : 0);
      const float lx1 = srcx - (float)x0, lx0 = 1.0f - lx1;
      float wgt = 0.0f;
      if (x0 == x) wgt += lx0;
      if (x1 == x) wgt += lx1;
      acc = fmaf(wgt, col[X], acc);
    }
    grad_in[((size_t)pl * ih + y) * iw + x] = mul * acc;
  }
}

}  // namespace bw
}  // namespace scflow

extern "C" int scflow_upsample_bwd(const float* grad_out, float* grad_in, int n, int c, int h, int w,
                                   int scale, float mul, void* stream) {
  using namespace scflow;
  if (!grad_out || !grad_in) return SCFLOW_E_NULL;
  if (n <= 0 || c <= 0 || h <= 0 || w <= 0 || scale <= 0 || (long long)n * c > 65535 ||
      (size_t)w * scale * sizeof(float) > 48 * 1024)
    return SCFLOW_E_BAD_SHAPE;
  dim3 grid(h, n * c);
  bw::resize_bwd_kernel<<<grid, 256, (size_t)w * scale * sizeof(float), (cudaStream_t)stream>>>(
      grad_out, grad_in, h, w, h * scale, w * scale, mul);
  return check_launch();
}

// ---------------------------------------------------------------------------------------------
// pose_flow_bwd: flow(x,y) = (K (R X + t)).xy / (.)z - (x,y) on foreground pixels (depth > 0),
// X = R0^-1 (K^-1 [x,y,1] d - t0) recomputed as in the forward.  Per crop:
//   dL/dt = sum_px K^T dL/du,   dL/dR = sum_px (K^T dL/du) X^T,
//   dL/du = (gx/uz, gy/uz, -(gx ux + gy uy)/uz^2).
namespace scflow {
namespace bw {

struct PoseBwdSmem { float R[9], t[3], K[9], Kinv[9], R0inv[9], t0[3]; float red[8][12]; };

__device__ __forceinline__ void mv3(const float* M, float a, float b, float c, float& x, float& y, float& z) {
  x = M[0] * a + M[1] * b + M[2] * c;
  y = M[3] * a + M[4] * b + M[5] * c;
  z = M[6] * a + M[7] * b + M[8] * c;
}

__global__ void __launch_bounds__(256)
pose_flow_bwd_kernel(const float* __restrict__ grad_full, const float* __restrict__ depth,
                     const float* __restrict__ K, const float* __restrict__ R0,
                     const float* __restrict__ t0, const float* __restrict__ R, const float* __restrict__ t,
                     float* __restrict__ grad_pose, int H, int W) {
  __shared__ PoseBwdSmem s;
  const int n = blockIdx.y;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 9; ++i) { s.R[i] = R[n * 9 + i]; s.K[i] = K[n * 9 + i]; }
    for (int i = 0; i < 3; ++i) { s.t[i] = t[n * 3 + i]; s.t0[i] = t0[n * 3 + i]; }
    inverse3x3(K + n * 9, s.Kinv);
    inverse3x3(R0 + n * 9, s.R0inv);
  }
  __syncthreads();
  const long long hw = (long long)H * W;
  float acc[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) acc[i] = 0.0f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (long long)gridDim.x * blockDim.x) {
    const float d = __ldg(depth + n * hw + i);
    if (d > 0.0f) {
      const int y = (int)(i / W), x = (int)(i - (long long)y * W);
      const float gx = __ldg(grad_full + ((long long)n * 2 + 0) * hw + i);
      const float gy = __ldg(grad_full + ((long long)n * 2 + 1) * hw + i);
      float cx, cy, cz, ox, oy, oz, rx, ry, rz, ux, uy, uz;
      mv3(s.Kinv, (float)x * d, (float)y * d, d, cx, cy, cz);
      mv3(s.R0inv, cx - s.t0[0], cy - s.t0[1], cz - s.t0[2], ox, oy, oz);
      mv3(s.R, ox, oy, oz, rx, ry, rz);
      mv3(s.K, rx + s.t[0], ry + s.t[1], rz + s.t[2], ux, uy, uz);
      const float iz = 1.0f / uz;
      const float du0 = gx * iz, du1 = gy * iz, du2 = -(gx * ux + gy * uy) * iz * iz;
      // dc = K^T du
      const float dc0 = s.K[0] * du0 + s.K[3] * du1 + s.K[6] * du2;
      const float dc1 = s.K[1] * du0 + s.K[4] * du1 + s.K[7] * du2;
      const float dc2 = s.K[2] * du0 + s.K[5] * du1 + s.K[8] * du2;
      acc[0] += dc0 * ox; acc[1] += dc0 * oy; acc[2] += dc0 * oz;
      acc[3] += dc1 * ox; acc[4] += dc1 * oy; acc[5] += dc1 * oz;
      acc[6] += dc2 * ox; acc[7] += dc2 * oy; acc[8] += dc2 * oz;
      acc[9] += dc0; acc[10] += dc1; acc[11] += dc2;
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < 12; ++i) {
    float v = acc[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s.red[warp][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < 12) {
    float v = 0.0f;
    for (int k = 0; k < 8; ++k) v += s.red[k][threadIdx.x];
    atomicAdd(grad_pose + n * 12 + threadIdx.x, v);
  }
}

}  // namespace bw
}  // namespace scflow

extern "C" int scflow_pose_flow_bwd(const float* grad_full, const float* depth, const float* K,
                                    const float* R0, const float* t0, const float* R, const float* t,
                                    float* grad_pose, int n, int H, int W, void* stream) {
  using namespace scflow;
  if (!grad_full || !depth || !K || !R0 || !t0 || !R || !t || !grad_pose) return SCFLOW_E_NULL;
  if (n <= 0 || H <= 0 || W <= 0 || n > 65535) return SCFLOW_E_BAD_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  if (cudaMemsetAsync(grad_pose, 0, (size_t)n * 12 * sizeof(float), st) != cudaSuccess) return check_launch();
  long long hw = (long long)H * W;
  int bx = (int)((hw + 256 * 8 - 1) / (256 * 8));
  if (bx > 16) bx = 16;
  if (bx < 1) bx = 1;
  bw::pose_flow_bwd_kernel<<<dim3(bx, n), 256, 0, st>>>(grad_full, depth, K, R0, t0, R, t, grad_pose, H, W);
  return check_launch();
}

// ---------------------------------------------------------------------------------------------
// pose compose (get_pose_from_delta_pose, models/utils/pose.py:124-149, ortho6d + 'exp') forward and backward for the
// training path: one thread per crop.  Under autograd the same algebra is ~20 element-wise launches forward and
// ~40 backward per iteration; here it is one each.
//   x = xr / |xr|, z = (x X yr) / |x X yr|, y = z X x, dR = [x y z] (columns), R' = dR R
//   tz' = tz / exp(dz), tx' = tz'* (dx / w + tx / tz), ty' likewise  (tz'* = tz' detached if detach_depth_for_xy)
namespace scflow {
namespace bw {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 cross3(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 add3(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 scale3(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
// F.normalize(v, eps = 1e-12): v / max(|v|, eps)
__device__ __forceinline__ V3 normalize3(V3 v, float* denom) {
  const float d = fmaxf(sqrtf(dot3(v, v)), 1e-12f);
  *denom = d;
  return scale3(v, 1.0f / d);
}
// adjoint of u = v / max(|v|, eps)
__device__ __forceinline__ V3 normalize3_bwd(V3 u, float denom, V3 gu) {
  if (denom <= 1e-12f) return scale3(gu, 1.0f / denom);
  return scale3(add3(gu, scale3(u, -dot3(u, gu))), 1.0f / denom);
}

__global__ void __launch_bounds__(128) pose_compose_fwd_kernel(const float* __restrict__ dR6, const float* __restrict__ dt,
                                                               const float* __restrict__ Rs, const float* __restrict__ ts,
                                                               float* __restrict__ Rd, float* __restrict__ td, int n,
                                                               float weight) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* r6 = dR6 + 6 * i;
  float dx, dz;
  const V3 x = normalize3({r6[0], r6[1], r6[2]}, &dx);
  const V3 yr = {r6[3], r6[4], r6[5]};
  const V3 z = normalize3(cross3(x, yr), &dz);
  const V3 y = cross3(z, x);
  const float D[3][3] = {{x.x, y.x, z.x}, {x.y, y.y, z.y}, {x.z, y.z, z.z}};      // columns x, y, z
  const float* R = Rs + 9 * i;
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) Rd[9 * i + 3 * a + b] = D[a][0] * R[b] + D[a][1] * R[3 + b] + D[a][2] * R[6 + b];
  const float tx = ts[3 * i], ty = ts[3 * i + 1], tz = ts[3 * i + 2];
  const float vz = tz / expf(dt[3 * i + 2]);
  td[3 * i] = vz * (dt[3 * i] / weight + tx / tz);
  td[3 * i + 1] = vz * (dt[3 * i + 1] / weight + ty / tz);
  td[3 * i + 2] = vz;
}

__global__ void __launch_bounds__(128) pose_compose_bwd_kernel(const float* __restrict__ gR, const float* __restrict__ gt,
                                                               const float* __restrict__ dR6, const float* __restrict__ dt,
                                                               const float* __restrict__ Rs, const float* __restrict__ ts,
                                                               float* __restrict__ g6, float* __restrict__ gdt,
                                                               float* __restrict__ gRs, float* __restrict__ gts, int n,
                                                               float weight, int detach_depth_for_xy) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* r6 = dR6 + 6 * i;
  float dnx, dnz;
  const V3 x = normalize3({r6[0], r6[1], r6[2]}, &dnx);
  const V3 yr = {r6[3], r6[4], r6[5]};
  const V3 z = normalize3(cross3(x, yr), &dnz);
  const V3 y = cross3(z, x);
  const float* R = Rs + 9 * i;
  float G[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
  if (gR) for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) G[a][b] = gR[9 * i + 3 * a + b];
  // R' = D R:  dL/dD = G R^T,  dL/dR = D^T G
  float gD[3][3];
  for (int a = 0; a < 3; ++a)
    for (int c = 0; c < 3; ++c) gD[a][c] = G[a][0] * R[3 * c] + G[a][1] * R[3 * c + 1] + G[a][2] * R[3 * c + 2];
  if (gRs) {
    const float D[3][3] = {{x.x, y.x, z.x}, {x.y, y.y, z.y}, {x.z, y.z, z.z}};
    for (int c = 0; c < 3; ++c)
      for (int b = 0; b < 3; ++b) gRs[9 * i + 3 * c + b] = D[0][c] * G[0][b] + D[1][c] * G[1][b] + D[2][c] * G[2][b];
  }
  const V3 gx0 = {gD[0][0], gD[1][0], gD[2][0]}, gy = {gD[0][1], gD[1][1], gD[2][1]}, gz0 = {gD[0][2], gD[1][2], gD[2][2]};
  // y = z X x
  const V3 gz = add3(gz0, cross3(x, gy));
  V3 gx = add3(gx0, cross3(gy, z));
  // z = normalize(zp), zp = x X yr
  const V3 gzp = normalize3_bwd(z, dnz, gz);
  gx = add3(gx, cross3(yr, gzp));
  const V3 gyr = cross3(gzp, x);
  const V3 gxr = normalize3_bwd(x, dnx, gx);
  g6[6 * i] = gxr.x; g6[6 * i + 1] = gxr.y; g6[6 * i + 2] = gxr.z;
  g6[6 * i + 3] = gyr.x; g6[6 * i + 4] = gyr.y; g6[6 * i + 5] = gyr.z;
  // translation
  const float tx = ts[3 * i], ty = ts[3 * i + 1], tz = ts[3 * i + 2];
  const float e = 1.0f / expf(dt[3 * i + 2]);
  const float vz = tz * e;
  const float ax = dt[3 * i] / weight + tx / tz, ay = dt[3 * i + 1] / weight + ty / tz;
  const float g0 = gt ? gt[3 * i] : 0.f, g1 = gt ? gt[3 * i + 1] : 0.f, g2 = gt ? gt[3 * i + 2] : 0.f;
  const float gvz = g2 + (detach_depth_for_xy ? 0.f : g0 * ax + g1 * ay);
  gdt[3 * i] = g0 * vz / weight;
  gdt[3 * i + 1] = g1 * vz / weight;
  gdt[3 * i + 2] = -vz * gvz;
  if (gts) {
    gts[3 * i] = g0 * vz / tz;
    gts[3 * i + 1] = g1 * vz / tz;
    gts[3 * i + 2] = gvz * e - vz * (g0 * tx + g1 * ty) / (tz * tz);
  }
}

}  // namespace bw
}  // namespace scflow

extern "C" int scflow_pose_compose_fwd(const float* dR6, const float* dt, const float* R_src, const float* t_src,
                                       float* R_dst, float* t_dst, int n, float weight, void* stream) {
  if (!dR6 || !dt || !R_src || !t_src || !R_dst || !t_dst) return SCFLOW_E_NULL;
  if (n <= 0 || weight == 0.0f) return SCFLOW_E_BAD_SHAPE;
  scflow::bw::pose_compose_fwd_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(dR6, dt, R_src, t_src, R_dst, t_dst, n,
                                                                                      weight);
  return scflow::check_launch();
}

extern "C" int scflow_pose_compose_bwd(const float* grad_R, const float* grad_t, const float* dR6, const float* dt,
                                       const float* R_src, const float* t_src, float* grad_dR6, float* grad_dt,
                                       float* grad_R_src, float* grad_t_src, int n, float weight,
                                       int detach_depth_for_xy, void* stream) {
  if (!dR6 || !dt || !R_src || !t_src || !grad_dR6 || !grad_dt || (!grad_R && !grad_t)) return SCFLOW_E_NULL;
  if (n <= 0 || weight == 0.0f) return SCFLOW_E_BAD_SHAPE;
  scflow::bw::pose_compose_bwd_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
      grad_R, grad_t, dR6, dt, R_src, t_src, grad_dR6, grad_dt, grad_R_src, grad_t_src, n, weight, detach_depth_for_xy);
  return scflow::check_launch();
}

// ---------------------------------------------------------------------------------------------
// ConvGRU gate arithmetic of the training path (models/decoder/raft_decoder.py:245-252) around the three gate
// convolutions, forward and adjoint, one launch each instead of the element-wise autograd chain (sigmoid, mul, tanh,
// 1 - z, mul, mul, add and their ~14 backward launches per pass):
//   reset:  rh = sigmoid(a_r) * h
//   blend:  h' = (1 - z) h + z q,  z = sigmoid(a_z), q = tanh(a_q)
// Contiguous fp32 tensors of one shape; four elements per thread where the count allows.
namespace scflow {
namespace bw {

__device__ __forceinline__ float sigmoidf_(float a) { return 1.0f / (1.0f + expf(-a)); }

template <int MODE>   // 0 reset fwd, 1 reset bwd, 2 blend fwd, 3 blend bwd
__global__ void __launch_bounds__(256) gru_gates_kernel(const float* __restrict__ g, const float* __restrict__ a0,
                                                        const float* __restrict__ a1, const float* __restrict__ h,
                                                        float* __restrict__ o0, float* __restrict__ o1,
                                                        float* __restrict__ o2, long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const float hv = h[i];
    if (MODE == 0) {
      o0[i] = sigmoidf_(a0[i]) * hv;
    } else if (MODE == 1) {
      const float r = sigmoidf_(a0[i]), gv = g[i];
      o0[i] = gv * hv * r * (1.0f - r);          // d a_r
      o1[i] = gv * r;                            // d h
    } else if (MODE == 2) {
      const float z = sigmoidf_(a0[i]), q = tanhf(a1[i]);
      o0[i] = (1.0f - z) * hv + z * q;
    } else {
      const float z = sigmoidf_(a0[i]), q = tanhf(a1[i]), gv = g[i];
      o0[i] = gv * (q - hv) * z * (1.0f - z);    // d a_z
      o1[i] = gv * z * (1.0f - q * q);           // d a_q
      o2[i] = gv * (1.0f - z);                   // d h
    }
  }
}

}  // namespace bw
}  // namespace scflow

static int gru_gates_launch(int mode, const float* g, const float* a0, const float* a1, const float* h, float* o0,
                            float* o1, float* o2, long long total, void* stream) {
  using namespace scflow::bw;
  if (total <= 0) return SCFLOW_E_BAD_SHAPE;
  long long blocks = (total + 255) / 256;
  if (blocks > 148LL * 16) blocks = 148LL * 16;
  cudaStream_t st = (cudaStream_t)stream;
  switch (mode) {
    case 0: gru_gates_kernel<0><<<(unsigned)blocks, 256, 0, st>>>(g, a0, a1, h, o0, o1, o2, total); break;
    case 1: gru_gates_kernel<1><<<(unsigned)blocks, 256, 0, st>>>(g, a0, a1, h, o0, o1, o2, total); break;
    case 2: gru_gates_kernel<2><<<(unsigned)blocks, 256, 0, st>>>(g, a0, a1, h, o0, o1, o2, total); break;
    default: gru_gates_kernel<3><<<(unsigned)blocks, 256, 0, st>>>(g, a0, a1, h, o0, o1, o2, total); break;
  }
  return scflow::check_launch();
}

extern "C" int scflow_gru_reset_fwd(const float* a_r, const float* h, float* rh, long long total, void* stream) {
  if (!a_r || !h || !rh) return SCFLOW_E_NULL;
  return gru_gates_launch(0, nullptr, a_r, nullptr, h, rh, nullptr, nullptr, total, stream);
}
extern "C" int scflow_gru_reset_bwd(const float* grad_rh, const float* a_r, const float* h, float* grad_a_r, float* grad_h,
                                    long long total, void* stream) {
  if (!grad_rh || !a_r || !h || !grad_a_r || !grad_h) return SCFLOW_E_NULL;
  return gru_gates_launch(1, grad_rh, a_r, nullptr, h, grad_a_r, grad_h, nullptr, total, stream);
}
extern "C" int scflow_gru_blend_fwd(const float* a_z, const float* a_q, const float* h, float* h_new, long long total,
                                    void* stream) {
  if (!a_z || !a_q || !h || !h_new) return SCFLOW_E_NULL;
  return gru_gates_launch(2, nullptr, a_z, a_q, h, h_new, nullptr, nullptr, total, stream);
}
extern "C" int scflow_gru_blend_bwd(const float* grad_h_new, const float* a_z, const float* a_q, const float* h,
                                    float* grad_a_z, float* grad_a_q, float* grad_h, long long total, void* stream) {
  if (!grad_h_new || !a_z || !a_q || !h || !grad_a_z || !grad_a_q || !grad_h) return SCFLOW_E_NULL;
  return gru_gates_launch(3, grad_h_new, a_z, a_q, h, grad_a_z, grad_a_q, grad_h, total, stream);
}
