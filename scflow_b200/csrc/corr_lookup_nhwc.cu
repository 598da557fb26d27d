// CorrLookup forward, channels-last output: one WARP per query, TMA-staged pyramid windows.
// Replaces models/utils/corr_lookup.py:102-136 (4x F.grid_sample + ~25 glue launches).
//
// HBM-bound gather (per query 324*4 B out + 4*100*4 B in + 8 B flow).  Design:
//   * out is (n, h, w, 324): a query's 324 results are one contiguous 1296-byte run, so a warp
//     stores full coalesced lines and the consumer (the 1x1 conv of the motion encoder) reads a
//     K-major operand.
//   * Each warp walks its queries with a 2-deep pipeline.  For query i+1 lanes 0..3 issue one
//     3-D TMA box per level, (<=16 x <=12 x 1) of the level viewed as (w_l, h_l, n*h*w); TMA's
//     out-of-bounds zero fill IS grid_sample's zeros padding, and the box origin is the window
//     origin computed with the reference's exact fp32 op sequence.  Completion lands on a
//     per-warp mbarrier; meanwhile the warp computes query i from the other buffer.
//   * Phase A: the 72 (level, axis, tap) sample coordinates of the query are computed once (lanes
//     in parallel), as {byte offset of the two corners inside the tile, two weights, zeroed where
//     the corner is outside the level} and parked in shared memory.
//   * Phase B: lane = 9*gb + a owns taps (a, b = gb + 3k), k = 0..2: 4 LDS + 4 FMA per tap
//     (channel = 81 l + 9 a + b; the three stores of a level fill one 324-byte run).
// Levels whose rows are not 16-byte multiples (w_l < 4, tiny test shapes) cannot be described to
// TMA and are staged with plain loads instead.
#include "common.cuh"
#include "tc_common.cuh"
#include <string.h>

namespace scflow {
namespace lk {

constexpr int kR = 4;
constexpr int kD = 2 * kR + 1;       // 9 taps per axis
constexpr int kLevels = 4;
constexpr int kTile = 12;            // window footprint per axis: 10 + one slack either side
constexpr int kTileW = 16;            // staged row: TMA's innermost coordinate must be 16-byte aligned,
                                     // so a windowed row is 16 floats starting at (origin & ~3)
constexpr int kTileFloats = kTile * kTileW;   // 192
constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;
constexpr int kCoords = kLevels * 2 * kD;   // 72
// per warp (sizes fixed by the host from the level shapes): 2 tile buffers, 72 x 16 B coordinate
// records, 2 mbarriers; every tile starts on a 128-byte boundary (TMA destination alignment)

struct Level {
  const float* ptr;
  int w, h;          // level extent
  int tw, th;        // staged tile extent: min(12, extent)
  int tma;           // 1: TMA box, 0: plain loads of the whole (tiny) level
  int off;           // tile offset inside a buffer, floats (multiple of 32)
  float dx, dy;      // max(extent - 1, 1) as float
  float rx, ry;      // 1 / dx, 1 / dy (IEEE)
};

struct Params {
  Level lv[kLevels];
  const float* flow;
  float* out;
  short* windows;
  int n, h, w, div_mode;
  long long total_q;
  unsigned tx_bytes;   // bytes all TMA boxes of one query deliver
  int buf_floats;      // one tile buffer (all levels)
  int warp_bytes;      // shared memory per warp
};

// Un-normalised sample coordinate: every step one separately rounded fp32 op, as the reference
// issues them: corr_lookup.py:128 (centroid + delta), :64-65 (*2. / max(size-1,1) - 1.), then
// ATen grid_sampler_unnormalize(align_corners=True): ((g + 1) / 2) * (size - 1).
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

// centre of the window at level l along one axis: (pixel + flow) / 2**l, exact scaling
__device__ __forceinline__ float centre_at(int pix, float fl, int l) {
  return __fmul_rn(__fadd_rn((float)pix, fl), 1.0f / (float)(1 << l));
}

__global__ void __launch_bounds__(kThreads, 4)
corr_lookup_nhwc_kernel(const __grid_constant__ CUtensorMap map0, const __grid_constant__ CUtensorMap map1,
                        const __grid_constant__ CUtensorMap map2, const __grid_constant__ CUtensorMap map3,
                        const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* mine = smem + warp * p.warp_bytes;
  float* tiles = reinterpret_cast<float*>(mine);                            // [2][buf_floats]
  uint4* info = reinterpret_cast<uint4*>(mine + 2 * p.buf_floats * 4);      // [72]
  uint64_t* bars = reinterpret_cast<uint64_t*>(mine + 2 * p.buf_floats * 4 + kCoords * 16);

  if (lane == 0) {
    tc::mbar_init(&bars[0], 1);
    tc::mbar_init(&bars[1], 1);
    tc::fence_barrier_init();
    if (warp == 0) {
      tc::tma_prefetch_desc(&map0); tc::tma_prefetch_desc(&map1);
      tc::tma_prefetch_desc(&map2); tc::tma_prefetch_desc(&map3);
    }
  }
  __syncwarp();

  const int P = p.h * p.w;
  const long long nwarps = (long long)gridDim.x * kWarps;
  const long long q_first = (long long)blockIdx.x * kWarps + warp;

  auto load_flow = [&](long long q, float& fx, float& fy) {
    fx = 0.0f; fy = 0.0f;
    if (q < p.total_q) {
      const long long n = q / P;
      const long long qi = q - n * P;
      fx = __ldg(p.flow + (n * 2 + 0) * P + qi);
      fy = __ldg(p.flow + (n * 2 + 1) * P + qi);
    }
  };

  // stage the four level windows of query q into buffer `b`
  auto stage = [&](long long q, int b, float fx, float fy) {
    const int qi = (int)(q % P);
    const int y = qi / p.w, x = qi - y * p.w;
    float* dst = tiles + b * p.buf_floats;
    // lanes 0..3 work out the window origin of "their" level; lane 0 alone talks to TMA
    int ox = 0, oy = 0;
    {
      const int l = lane & (kLevels - 1);
      const Level& L = p.lv[l];
      if (L.w > kTile)
        ox = floor_clamped(unnorm_coord(centre_at(x, fx, l), -kR, L.dx, L.rx, (float)(L.w - 1), p.div_mode),
                           -32, L.w + 32) - 1;
      ox &= ~3;
      if (L.h > kTile)
        oy = floor_clamped(unnorm_coord(centre_at(y, fy, l), -kR, L.dy, L.ry, (float)(L.h - 1), p.div_mode),
                           -32, L.h + 32) - 1;
    }
    if (lane == 0) tc::mbar_expect_tx(&bars[b], p.tx_bytes);
#pragma unroll
    for (int l = 0; l < kLevels; ++l) {
      const int oxl = __shfl_sync(0xffffffffu, ox, l), oyl = __shfl_sync(0xffffffffu, oy, l);
      if (lane == 0 && p.lv[l].tma) {
        const CUtensorMap* m = l == 0 ? &map0 : l == 1 ? &map1 : l == 2 ? &map2 : &map3;
        tc::tma_load_3d(m, &bars[b], dst + p.lv[l].off, oxl, oyl, (int)q);
      }
    }
#pragma unroll
    for (int l = 0; l < kLevels; ++l) {
      if (!p.lv[l].tma) {                       // tiny level: whole map, plain loads
        const int cnt = p.lv[l].w * p.lv[l].h;
        const float* src = p.lv[l].ptr + q * cnt;
        for (int e = lane; e < cnt; e += 32) dst[p.lv[l].off + e] = __ldg(src + e);
      }
    }
  };

  float fx, fy, fxn, fyn;
  load_flow(q_first, fx, fy);
  load_flow(q_first + nwarps, fxn, fyn);
  if (q_first < p.total_q) stage(q_first, 0, fx, fy);

  uint32_t it = 0;
  for (long long q = q_first; q < p.total_q; q += nwarps, ++it) {
    const int b = it & 1;
    const long long qn = q + nwarps;
    if (qn < p.total_q) stage(qn, b ^ 1, fxn, fyn);
    float fxnn, fynn;
    load_flow(qn + nwarps, fxnn, fynn);

    // ---- phase A: 72 sample coordinates -> {corner byte offsets, zero-masked weights}
    const int qi = (int)(q % P);
    const int y = qi / p.w, x = qi - y * p.w;
#pragma unroll
    for (int rr = 0; rr < 3; ++rr) {
      const int i = lane + 32 * rr;
      if (i < kCoords) {
        const int l = i / (2 * kD);
        const int rem = i - l * (2 * kD);
        const int axis = rem >= kD ? 1 : 0;
        const int k = rem - axis * kD;
        const Level& L = p.lv[l];
        const int size = axis ? L.h : L.w;
        const int tdim = axis ? L.th : L.tw;
        const float den = axis ? L.dy : L.dx, rcp = axis ? L.ry : L.rx;
        const float c = centre_at(axis ? y : x, axis ? fy : fx, l);
        const float iv = unnorm_coord(c, k - kR, den, rcp, (float)(size - 1), p.div_mode);
        const float fl = floorf(iv);
        const int i0 = floor_clamped(iv, -32, size + 32);
        int org = 0;
        if (size > kTile)
          org = floor_clamped(unnorm_coord(c, -kR, den, rcp, (float)(size - 1), p.div_mode), -32, size + 32) - 1;
        if (!axis) org &= ~3;
        const int r0 = i0 - org, r1 = r0 + 1;
        const bool ok0 = (unsigned)i0 < (unsigned)size && (unsigned)r0 < (unsigned)tdim;
        const bool ok1 = (unsigned)(i0 + 1) < (unsigned)size && (unsigned)r1 < (unsigned)tdim;
        const float w1 = __fsub_rn(iv, fl);                       // ATen: ix - ix_nw
        const float w0 = __fsub_rn(__fadd_rn(fl, 1.0f), iv);      // ATen: ix_se - ix
        const int mult = axis ? L.tw * 4 : 4;
        const int base = axis ? (b * p.buf_floats + L.off) * 4 : 0;
        uint4 rec;
        rec.x = (unsigned)(min(max(r0, 0), tdim - 1) * mult + base);
        rec.y = (unsigned)(min(max(r1, 0), tdim - 1) * mult + base);
        rec.z = __float_as_uint(ok0 ? w0 : 0.0f);
        rec.w = __float_as_uint(ok1 ? w1 : 0.0f);
        info[i] = rec;
        if (p.windows != nullptr) p.windows[q * kCoords + i] = (short)i0;
      }
    }
    __syncwarp();
    tc::mbar_wait(&bars[b], (it >> 1) & 1);

    // ---- phase B: lane = 9*gb + a owns taps (a, b = gb + 3k).  Rows are 16 floats apart, so a
    // warp-wide LDS sees columns a..a+1 of rows gb, gb+1, gb+2: at most a 2-way bank conflict.
    if (lane < 27) {
      const int gb = lane / kD, ta = lane - gb * kD;
      const uint8_t* tbase = reinterpret_cast<const uint8_t*>(tiles);
      float* oq = p.out + q * (kLevels * kD * kD) + kD * ta + gb;
#pragma unroll
      for (int l = 0; l < kLevels; ++l) {
        const uint4 xr = info[l * 2 * kD + ta];
        const float wx0 = __uint_as_float(xr.z), wx1 = __uint_as_float(xr.w);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const uint4 yr = info[l * 2 * kD + kD + gb + 3 * k];
          const float wy0 = __uint_as_float(yr.z), wy1 = __uint_as_float(yr.w);
          const float v00 = *reinterpret_cast<const float*>(tbase + yr.x + xr.x);
          const float v01 = *reinterpret_cast<const float*>(tbase + yr.x + xr.y);
          const float v10 = *reinterpret_cast<const float*>(tbase + yr.y + xr.x);
          const float v11 = *reinterpret_cast<const float*>(tbase + yr.y + xr.y);
          // ATen grid_sampler_2d: nw, ne, sw, se accumulated in this order
          float acc = v00 * (wx0 * wy0);
          acc = fmaf(v01, wx1 * wy0, acc);
          acc = fmaf(v10, wx0 * wy1, acc);
          acc = fmaf(v11, wx1 * wy1, acc);
          oq[l * (kD * kD) + 3 * k] = acc;
        }
      }
    }
    __syncwarp();          // tile buffer b and the records are free again
    fx = fxn; fy = fyn; fxn = fxnn; fyn = fynn;
  }
}

}  // namespace lk
}  // namespace scflow

namespace scflow {

int launch_corr_lookup_nhwc(const float* const* levels, const float* flow, float* out, short* windows,
                            int n, int h, int w, int div_mode, cudaStream_t stream) {
  using namespace lk;
  if (((h >> (kLevels - 1)) < 1) || ((w >> (kLevels - 1)) < 1)) return SCFLOW_E_BAD_SHAPE;
  const long long total_q = (long long)n * h * w;
  if (total_q >= (1ll << 31)) return SCFLOW_E_BAD_SHAPE;
  Params p;
  CUtensorMap maps[kLevels];
  p.tx_bytes = 0;
  int off = 0;
  for (int l = 0; l < kLevels; ++l) {
    Level& L = p.lv[l];
    L.ptr = levels[l];
    L.w = w >> l; L.h = h >> l;
    L.tw = L.w <= kTile ? L.w : kTileW;     // windowed rows: 16 floats from a 16-byte aligned origin
    L.th = L.h <= kTile ? L.h : kTile;
    L.dx = (float)(L.w - 1 > 1 ? L.w - 1 : 1);
    L.dy = (float)(L.h - 1 > 1 ? L.h - 1 : 1);
    L.rx = 1.0f / L.dx; L.ry = 1.0f / L.dy;
    L.tma = ((L.w % 4) == 0 && aligned16(L.ptr)) ? 1 : 0;
    L.off = off;
    off += ((L.tw * L.th + 31) / 32) * 32;
    if (!L.tma && L.w * L.h > kTileFloats) return SCFLOW_E_BAD_SHAPE;   // plain path holds whole levels only
    if (!L.tma && (L.w > kTile || L.h > kTile)) return SCFLOW_E_BAD_SHAPE;
    const uint64_t dims[3] = {(uint64_t)L.w, (uint64_t)L.h, (uint64_t)total_q};
    const uint64_t strides[2] = {(uint64_t)L.w * 4, (uint64_t)L.w * L.h * 4};
    const uint32_t box[3] = {(uint32_t)L.tw, (uint32_t)L.th, 1};
    if (L.tma) {
      if (!tc::encode_f32_map(&maps[l], L.ptr, 3, dims, strides, box, tc::kTmaNone, /*l2_promote=*/false))
        return SCFLOW_E_DRIVER;
      p.tx_bytes += (unsigned)(L.tw * L.th * 4);
    } else {
      // an unused but valid descriptor keeps the kernel signature uniform
      maps[l] = maps[0];
    }
  }
  if (!p.lv[0].tma) {     // maps[0] itself unused: fill every slot with zeros (never dereferenced)
    for (int l = 0; l < kLevels; ++l)
      if (!p.lv[l].tma) memset(&maps[l], 0, sizeof(CUtensorMap));
  }
  p.buf_floats = off;
  p.warp_bytes = ((2 * off * 4 + kCoords * 16 + 16 + 127) / 128) * 128;
  const int smem_bytes = kWarps * p.warp_bytes + 128;
  p.flow = flow; p.out = out; p.windows = windows;
  p.n = n; p.h = h; p.w = w; p.div_mode = div_mode; p.total_q = total_q;
  static int attr_bytes = 0;
  if (smem_bytes > attr_bytes) {
    cudaFuncSetAttribute(corr_lookup_nhwc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    attr_bytes = smem_bytes;
  }
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  long long blocks = (total_q + kWarps - 1) / kWarps;
  if (blocks > (long long)sms * 4) blocks = (long long)sms * 4;
  corr_lookup_nhwc_kernel<<<(unsigned)blocks, kThreads, smem_bytes, stream>>>(maps[0], maps[1], maps[2],
                                                                           maps[3], p);
  return check_launch();
}

}  // namespace scflow
