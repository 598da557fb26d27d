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
//   * Phase B: lane = 9*gb + a owns taps (a, b = 3gb + k), k = 0..2, sharing tile rows between
//     consecutive taps; the 324 results of the query are staged in shared memory and leave as
//     one 1296-byte bulk copy (cp.async.bulk shared -> global).
// Levels whose rows are not 16-byte multiples (w_l < 4, tiny test shapes) cannot be described to
// TMA and are staged with plain loads instead.
#include "common.cuh"
#include "tc_common.cuh"
#include <stdlib.h>
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
  int step_q, step_n, step_qi;   // a warp's stride through the queries: step_q = step_n*P + step_qi
  unsigned magic_w;    // ceil(2^32 / w): qi / w == umulhi(qi, magic_w) for qi < h*w
};

// Un-normalised sample coordinate: every step one separately rounded fp32 op, as the reference
// issues them: corr_lookup.py:128 (centroid + delta), :64-65 (*2. / max(size-1,1) - 1.), then
// ATen grid_sampler_unnormalize(align_corners=True): ((g + 1) / 2) * (size - 1).
// The IEEE quotient g / denom is formed as q0 = g*rcp; r = fma(-q0, denom, g); q = fma(r, rcp, q0)
// with rcp = RN(1/denom): correctly rounded for every finite g and the small integer denominators
// of a pyramid level (Markstein's theorem; checked against true division on 9e7 operands), at a
// third of the instructions of the generic division sequence.
template <int DIV>
__device__ __forceinline__ float unnorm_coord(float centre_lvl, float off, float denom, float rcp,
                                              float sizem1) {
  const float c = __fadd_rn(centre_lvl, off);
  float g = __fmul_rn(c, 2.0f);
  if (DIV == SCFLOW_DIV_IEEE) {
    const float q0 = __fmul_rn(g, rcp);
    const float r = __fmaf_rn(-q0, denom, g);
    g = __fmaf_rn(r, rcp, q0);
  } else {
    g = __fmul_rn(g, rcp);
  }
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

// A warp's position in the query stream: q = n*P + qi, advanced by a fixed stride without division.
struct Cursor {
  int q, n, qi;
};

template <int DIV, int MINB>
__global__ void __launch_bounds__(kThreads, MINB)
corr_lookup_nhwc_kernel(const __grid_constant__ CUtensorMap map0, const __grid_constant__ CUtensorMap map1,
                        const __grid_constant__ CUtensorMap map2, const __grid_constant__ CUtensorMap map3,
                        const Params p) {
  extern __shared__ uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // byte offsets from smem_raw (plain shared-memory indexing, so the compiler schedules the loads)
  const uint32_t sbase = tc::smem_u32(smem_raw);
  const uint32_t mine = (((sbase + 127u) & ~127u) - sbase) + warp * p.warp_bytes;
  const uint32_t tiles = mine;                                  // [2][buf_floats]
  const uint32_t info = mine + 2 * p.buf_floats * 4;            // [2][72] x 16 B
  const uint32_t bars = info + 2 * kCoords * 16;                // 2 mbarriers
  float* tiles_gen = reinterpret_cast<float*>(smem_raw + tiles);
  uint64_t* bars_gen = reinterpret_cast<uint64_t*>(smem_raw + bars);
  auto ldf = [&](uint32_t off) { return *reinterpret_cast<const float*>(smem_raw + off); };
  auto ld4 = [&](uint32_t off) { return *reinterpret_cast<const uint4*>(smem_raw + off); };

  if (lane == 0) {
    tc::mbar_init(&bars_gen[0], 1);
    tc::mbar_init(&bars_gen[1], 1);
    tc::fence_barrier_init();
    if (warp == 0) {
      tc::tma_prefetch_desc(&map0); tc::tma_prefetch_desc(&map1);
      tc::tma_prefetch_desc(&map2); tc::tma_prefetch_desc(&map3);
    }
  }
  __syncwarp();

  const int P = p.h * p.w;
  const int total = (int)p.total_q;
  auto advance = [&](Cursor c) {
    c.q += p.step_q; c.n += p.step_n; c.qi += p.step_qi;
    if (c.qi >= P) { c.qi -= P; c.n += 1; }
    return c;
  };
  auto load_flow = [&](const Cursor& c, float& fx, float& fy) {
    fx = 0.0f; fy = 0.0f;
    if (c.q < total) {
      const float* f = p.flow + (size_t)c.n * 2 * P + c.qi;
      fx = __ldg(f);
      fy = __ldg(f + P);
    }
  };

  // ---- per-lane constants of the coordinates this lane computes in phase A: coordinate
  // i = lane + 32*r (r < 3), i = 18*level + 9*axis + tap.  tap0 = where tap 0 of the same
  // (level, axis) lives (its floor fixes the window origin): lane | (round == r ? 32 : 0).
  float a_den[3], a_rcp[3], a_szm1[3], a_scale[3], a_off[3];
  int a_pack[3];   // tdim | mult << 8 | base << 16 | windowed << 30 | axis << 31
  int a_tap0[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int i = min(lane + 32 * r, kCoords - 1);
    const int l = i / (2 * kD);
    const int rem = i - l * (2 * kD);
    const int axis = rem >= kD ? 1 : 0;
    const int k = rem - axis * kD;
    const Level& L = p.lv[l];
    const int size = axis ? L.h : L.w;
    a_den[r] = axis ? L.dy : L.dx;
    a_rcp[r] = axis ? L.ry : L.rx;
    a_szm1[r] = (float)(size - 1);
    a_scale[r] = 1.0f / (float)(1 << l);
    a_off[r] = (float)(k - kR);
    const int tdim = axis ? L.th : L.tw, mult = axis ? L.tw * 4 : 4, base = axis ? L.off * 4 : 0;
    a_pack[r] = tdim | (mult << 8) | (base << 16) | ((size > kTile ? 1 : 0) << 30) | (axis << 31);
    const int t0 = i - k;
    a_tap0[r] = (t0 & 31) | ((t0 >> 5) == r ? 32 : 0);
  }
  // lanes 0..3 issue the TMA box of level `lane`: its x / y origins sit with tap 0 of coordinates
  // 18*lane and 18*lane + 9, i.e. in round (lane >> 1) at these lanes
  const int tx_lane = (18 * (lane & 3)) & 31, ty_lane = (18 * (lane & 3) + 9) & 31;
  // phase B: lane = 9*gb + a
  const int gb = lane / kD, ta = lane - gb * kD;
  const uint32_t xrec = info + ta * 16, yrec = info + (kD + 3 * gb) * 16;

  // Phase A + staging of the query at `c` (flow fx, fy) into record / tile buffer `b`.
  auto prepare = [&](const Cursor& c, float fx, float fy, int b) {
    const int y = (int)__umulhi((unsigned)c.qi, p.magic_w);
    const int x = c.qi - y * p.w;
    float iv[3];
    int i0[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const bool ax = a_pack[r] < 0;
      const float ctr = __fmul_rn(__fadd_rn((float)(ax ? y : x), ax ? fy : fx), a_scale[r]);
      iv[r] = unnorm_coord<DIV>(ctr, a_off[r], a_den[r], a_rcp[r], a_szm1[r]);
      // floor, clamped to [-32, size + 32] (far-out windows are all padding anyway; NaN -> -32)
      i0[r] = (int)fminf(fmaxf(floorf(iv[r]), -32.0f), a_szm1[r] + 33.0f);
    }
    int org[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int same = __shfl_sync(0xffffffffu, i0[r], a_tap0[r] & 31);
      const int prev = r > 0 ? __shfl_sync(0xffffffffu, i0[r - 1], a_tap0[r] & 31) : 0;
      const int t0 = (a_tap0[r] & 32) ? same : prev;
      const int al = a_pack[r] < 0 ? ~0 : ~3;                  // x origins: 16-byte aligned
      org[r] = ((a_pack[r] >> 30) & 1) ? ((t0 - 1) & al) : 0;
    }
    const uint32_t tilebase = tiles + b * p.buf_floats * 4;
    const uint32_t recbase = info + b * (kCoords * 16);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int i = lane + 32 * r;
      if (i < kCoords) {
        const int tdim = a_pack[r] & 0xff, mult = (a_pack[r] >> 8) & 0xff, base = (a_pack[r] >> 16) & 0x3fff;
        const float fl = floorf(iv[r]);
        const int r0 = i0[r] - org[r], r1 = r0 + 1;
        // a corner outside the tile is outside the level (the tile covers every tap of the window;
        // inside the tile TMA has zero-filled what lies outside the level): weight 0 = zeros padding
        const bool ok0 = (unsigned)r0 < (unsigned)tdim, ok1 = (unsigned)r1 < (unsigned)tdim;
        const float w1 = __fsub_rn(iv[r], fl);                       // ATen: ix - ix_nw
        const float w0 = __fsub_rn(__fadd_rn(fl, 1.0f), iv[r]);      // ATen: ix_se - ix
        const uint32_t rb = a_pack[r] < 0 ? tilebase + base : 0u;
        uint4 rec;
        rec.x = (unsigned)(min(max(r0, 0), tdim - 1) * mult) + rb;
        rec.y = (unsigned)(min(max(r1, 0), tdim - 1) * mult) + rb;
        rec.z = __float_as_uint(ok0 ? w0 : 0.0f);
        rec.w = __float_as_uint(ok1 ? w1 : 0.0f);
        *reinterpret_cast<uint4*>(smem_raw + recbase + i * 16) = rec;
        if (p.windows != nullptr) p.windows[(size_t)c.q * kCoords + i] = (short)i0[r];
      }
    }
    // ---- staging: the bulk store of two queries ago used tile buffer b as its staging line
    const int ox0 = __shfl_sync(0xffffffffu, org[0], tx_lane), ox1 = __shfl_sync(0xffffffffu, org[1], tx_lane);
    const int oy0 = __shfl_sync(0xffffffffu, org[0], ty_lane), oy1 = __shfl_sync(0xffffffffu, org[1], ty_lane);
    if (lane == 0) {
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      tc::mbar_expect_tx(&bars_gen[b], p.tx_bytes);
    }
    __syncwarp();
    float* dst = tiles_gen + b * p.buf_floats;
    if (lane < kLevels && p.lv[lane].tma) {
      const CUtensorMap* m = lane == 0 ? &map0 : lane == 1 ? &map1 : lane == 2 ? &map2 : &map3;
      tc::tma_load_3d(m, &bars_gen[b], dst + p.lv[lane].off, lane < 2 ? ox0 : ox1, lane < 2 ? oy0 : oy1, c.q);
    }
#pragma unroll
    for (int l = 0; l < kLevels; ++l) {
      if (!p.lv[l].tma) {                        // tiny level: whole map, plain loads
        const int cnt = p.lv[l].w * p.lv[l].h;
        const float* src = p.lv[l].ptr + (size_t)c.q * cnt;
        for (int e = lane; e < cnt; e += 32) dst[p.lv[l].off + e] = __ldg(src + e);
      }
    }
  };

  const int q_first = blockIdx.x * kWarps + warp;
  Cursor c0{q_first, q_first / P, q_first % P};
  Cursor c1 = advance(c0), c2 = advance(c1);
  float fx0, fy0, fx1, fy1;
  load_flow(c0, fx0, fy0);
  load_flow(c1, fx1, fy1);
  if (c0.q < total) prepare(c0, fx0, fy0, 0);

  for (uint32_t it = 0; c0.q < total; ++it) {
    const int b = it & 1;
    if (c1.q < total) prepare(c1, fx1, fy1, b ^ 1);      // records + TMA for the next query
    float fx2, fy2;
    load_flow(c2, fx2, fy2);
    __syncwarp();
    tc::mbar_wait(&bars_gen[b], (it >> 1) & 1);

    // ---- phase B: lane = 9*gb + a owns taps (a, b = 3*gb + k), k = 0..2.  Consecutive taps
    // share a tile row (bottom row of tap k = top row of tap k+1 unless the coordinate round trip
    // landed on the other side of an integer), so a level costs 8 LDS per lane, not 12.  Rows are
    // 16 floats apart: a warp-wide LDS sees columns a..a+1 of 3 rows, at most a 2-way conflict.
    float res[kLevels][3];
    if (lane < 27) {
      const uint32_t rb = b * (kCoords * 16);
#pragma unroll
      for (int l = 0; l < kLevels; ++l) {
        const uint4 xr = ld4(rb + xrec + l * (2 * kD * 16));
        const uint4 y0 = ld4(rb + yrec + l * (2 * kD * 16));
        const uint4 y1 = ld4(rb + yrec + l * (2 * kD * 16) + 16);
        const uint4 y2 = ld4(rb + yrec + l * (2 * kD * 16) + 32);
        const float wx0 = __uint_as_float(xr.z), wx1 = __uint_as_float(xr.w);
        // horizontal interpolation once per tile row, vertical per tap (separable form of the
        // four-corner sum: equal to ATen's up to fp32 rounding order).  Nominal rows first: all
        // eight loads are independent.
        const float h0 = fmaf(ldf(y0.x + xr.y), wx1, ldf(y0.x + xr.x) * wx0);
        const float h1 = fmaf(ldf(y0.y + xr.y), wx1, ldf(y0.y + xr.x) * wx0);
        const float h2 = fmaf(ldf(y1.y + xr.y), wx1, ldf(y1.y + xr.x) * wx0);
        const float h3 = fmaf(ldf(y2.y + xr.y), wx1, ldf(y2.y + xr.x) * wx0);
        float t1 = h1, t2 = h2;
        if (y1.x != y0.y) t1 = fmaf(ldf(y1.x + xr.y), wx1, ldf(y1.x + xr.x) * wx0);
        if (y2.x != y1.y) t2 = fmaf(ldf(y2.x + xr.y), wx1, ldf(y2.x + xr.x) * wx0);
        res[l][0] = fmaf(h1, __uint_as_float(y0.w), h0 * __uint_as_float(y0.z));
        res[l][1] = fmaf(h2, __uint_as_float(y1.w), t1 * __uint_as_float(y1.z));
        res[l][2] = fmaf(h3, __uint_as_float(y2.w), t2 * __uint_as_float(y2.z));
      }
    }
    __syncwarp();          // every lane is done reading tile buffer b: it becomes the staging line
    const uint32_t stage_line = tiles + b * p.buf_floats * 4;
    if (lane < 27) {
      float* so = reinterpret_cast<float*>(smem_raw + stage_line) + kD * ta + 3 * gb;
#pragma unroll
      for (int l = 0; l < kLevels; ++l) {
        so[l * (kD * kD) + 0] = res[l][0];
        so[l * (kD * kD) + 1] = res[l][1];
        so[l * (kD * kD) + 2] = res[l][2];
      }
    }
    tc::fence_proxy_async();        // written by the generic proxy, read by the bulk copy
    __syncwarp();
    if (lane == 0) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(
                       p.out + (size_t)c0.q * (kLevels * kD * kD)),
                   "r"(sbase + stage_line), "n"(kLevels * kD * kD * 4)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    c0 = c1; c1 = c2; c2 = advance(c2);
    fx0 = fx1; fy0 = fy1; fx1 = fx2; fy1 = fy2;
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace lk
}  // namespace scflow

namespace scflow {

int launch_corr_lookup_v2(const float* const* levels, const float* flow, float* out, short* windows,
                          int n, int h, int w, int div_mode, bool split, cudaStream_t stream);

int launch_corr_lookup_nhwc(const float* const* levels, const float* flow, float* out, short* windows,
                            int n, int h, int w, int div_mode, bool gen1, bool split, cudaStream_t stream) {
  using namespace lk;
  // second-generation kernel (corr_lookup_v2.cu) wherever every level row is a 16-byte multiple; this file's
  // kernel takes the other shapes, and any shape when the caller asks for it (SCFLOW_LAYOUT_NHWC_GEN1: the
  // tests hold the two independently written kernels to identical windows and values)
  if (!gen1) {
    const int rc = launch_corr_lookup_v2(levels, flow, out, windows, n, h, w, div_mode, split, stream);
    if (rc != SCFLOW_E_UNSUPPORTED) return rc;
  }
  if (split) return SCFLOW_E_UNSUPPORTED;      // split planes come from the second-generation kernel only
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
  if (off < kLevels * kD * kD) off = ((kLevels * kD * kD + 31) / 32) * 32;   // doubles as the staging line
  p.buf_floats = off;
  p.warp_bytes = ((2 * off * 4 + 2 * kCoords * 16 + 16 + 127) / 128) * 128;
  const int smem_bytes = kWarps * p.warp_bytes + 128;
  p.flow = flow; p.out = out; p.windows = windows;
  p.n = n; p.h = h; p.w = w; p.div_mode = div_mode; p.total_q = total_q;
  if ((double)h * w * w >= 4294967296.0) return SCFLOW_E_BAD_SHAPE;
  p.magic_w = (w & (w - 1)) == 0 ? (unsigned)(4294967296ull / (unsigned)w) : (unsigned)(4294967296ull / (unsigned)w) + 1u;
  constexpr int minb = 3;                           // 3 CTAs/SM (78 registers, no spills) measured faster than 4
  auto kern = div_mode == SCFLOW_DIV_IEEE ? corr_lookup_nhwc_kernel<SCFLOW_DIV_IEEE, 3>
                                          : corr_lookup_nhwc_kernel<SCFLOW_DIV_RCP, 3>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);   // per device: every launch
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  long long blocks = (total_q + kWarps - 1) / kWarps;
  if (blocks > (long long)sms * minb) blocks = (long long)sms * minb;
  const long long nwarps = blocks * kWarps;
  p.step_q = (int)nwarps; p.step_n = (int)(nwarps / (h * w)); p.step_qi = (int)(nwarps % (h * w));
  kern<<<(unsigned)blocks, kThreads, smem_bytes, stream>>>(maps[0], maps[1], maps[2], maps[3], p);
  return check_launch();
}

}  // namespace scflow
