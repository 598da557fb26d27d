// CorrLookup forward, channels-last output, second generation: one WARP per query, every level
// staged as a fixed 16 x 12 TMA box, compact sample records, immediate-offset taps.
// Replaces models/utils/corr_lookup.py:102-136 (4x F.grid_sample + ~25 glue launches).
//
// Same contract and the same fp32 op sequence as corr_lookup_nhwc.cu (window selection bit-exact,
// values identical), restructured for fewer issued instructions and fewer shared-memory
// wavefronts per query (that kernel was issue / L1-bound at 45 % of HBM peak):
//   * EVERY level is a windowed 16 x 12 box whose origin follows the query's window, also where
//     the level is smaller than the box: TMA zero-fills what lies outside the level (= grid_sample
//     zeros padding), so both corners of every tap are inside the tile by construction and the
//     records need no clamping and no validity flags.
//   * Phase A (coordinates): lane = 9*g + k computes tap k; round 0 = x of level g, round 1 = y of
//     level g (g = 0..2), round 2 = x | y of level 3.  A round is axis-uniform, the window origin
//     is one shuffle from tap 0 of the same group, and the lanes 0 / 9 / 18 / 27 that issue the
//     TMA boxes of levels 0..3 already hold their origins.
//   * Records: x tap = {byte offset in the tile row, w1} (8 B); y taps in groups of three =
//     {tile byte offset of the first tap's row | row deltas of taps 1, 2, w1 x 3} (16 B).
//     w0 = 1 - w1 is bit-identical to ATen's (ix_se - ix) wherever it multiplies a corner inside
//     the level (ix - floor(ix) is exact for ix >= 0), so it is not stored.
//   * Phase B: lane = 9*gb + a owns taps (a, 3gb..3gb+2).  When the three y floors are consecutive
//     (always, unless the normalise/un-normalise round trip dropped a coordinate just under an
//     integer: zero flow) the 8 corner loads sit at compile-time offsets from ONE address; a
//     warp-uniform slow path handles the other queries with per-tap rows.
//   * The 324 results are staged in the consumed tile buffer and leave as one 1296-byte bulk copy.
// Needs every level row to be a 16-byte multiple (w % 32 == 0); other shapes take the first-
// generation kernel.
#include "common.cuh"
#include "tc_common.cuh"
#include <stdlib.h>
#include <string.h>

namespace scflow {
namespace lk2 {

constexpr int kR = 4;
constexpr int kD = 2 * kR + 1;       // 9 taps per axis
constexpr int kLevels = 4;
constexpr int kTileH = 12;           // 10 rows of taps + one slack row either side
constexpr int kTileW = 16;           // 10 columns + slack, from a 16-byte aligned origin
constexpr int kRowBytes = kTileW * 4;                // 64
constexpr int kTileBytes = kTileH * kRowBytes;       // 768
constexpr int kBufBytes = kLevels * kTileBytes;      // 3072: one query's four tiles
constexpr int kXRecBytes = kLevels * kD * 8;         // 288
constexpr int kYRecBytes = kLevels * 3 * 16;         // 192
constexpr int kRecBytes = kXRecBytes + kYRecBytes;   // 480
constexpr int kWarpBytes = ((2 * kBufBytes + 2 * kRecBytes + 16 + 127) / 128) * 128;   // 7168
constexpr int kOutFloats = kLevels * kD * kD;        // 324
constexpr int kSplitC = (kOutFloats + 7) / 8 * 8;    // 328: channels per plane of the split-bf16 output

struct Maps {
  CUtensorMap m[kLevels];
};

struct LevelC {
  float dx, dy;      // max(extent - 1, 1)
  float rx, ry;      // 1 / dx, 1 / dy (IEEE)
  float sx, sy;      // extent - 1
};

struct Params {
  LevelC lv[kLevels];
  const float* flow;
  float* out;
  short* windows;
  int h, w, total_q;
  int step_q, step_n, step_qi;   // a warp's stride through the queries: step_q = step_n*P + step_qi
  unsigned magic_w;              // ceil(2^32 / w): qi / w == umulhi(qi, magic_w) for qi < h*w
};

// Un-normalised sample coordinate, one separately rounded fp32 op per reference op
// (corr_lookup.py:128, :64-65, ATen grid_sampler_unnormalize with align_corners=True); the IEEE
// quotient through Markstein's correction, see corr_lookup_nhwc.cu.
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

struct Cursor {
  int q, n, qi;
};

template <int DIV, int WARPS, int MINB, bool WIN, bool SPLIT>
__global__ void __launch_bounds__(WARPS * 32, MINB)
corr_lookup_v2_kernel(const __grid_constant__ Maps maps, const __grid_constant__ Params p) {
  // no static shared memory in this kernel: the dynamic window starts 1 KiB-aligned, which the
  // 128-byte TMA destination alignment relies on (checked once below)
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sbase = tc::smem_u32(smem_raw);
  if ((sbase & 127u) != 0u) __trap();
  const uint32_t mine = warp * kWarpBytes;
  const uint32_t tiles = mine;                       // [2][kBufBytes]
  const uint32_t recs = mine + 2 * kBufBytes;        // [2][x records | y records]
  const uint32_t bars = recs + 2 * kRecBytes;        // 2 mbarriers
  uint64_t* bars_gen = reinterpret_cast<uint64_t*>(smem_raw + bars);
  auto ldf = [&](uint32_t off) { return *reinterpret_cast<const float*>(smem_raw + off); };

  if (lane == 0) {
    tc::mbar_init(&bars_gen[0], 1);
    tc::mbar_init(&bars_gen[1], 1);
    tc::fence_barrier_init();
  }
  if (warp == 0 && lane < kLevels) tc::tma_prefetch_desc(&maps.m[lane]);
  __syncwarp();

  const int P = p.h * p.w;
  const int total = p.total_q;
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

  // ---- lane roles.  lane = 9*g + k; lanes 27..31 (g = 3) idle in phase A / B, lane 27 issues
  // the TMA box of level 3.
  const int g = lane / kD, k = lane - g * kD;
  const int grp0 = lane - k;                       // lane of tap 0 of this group
  const int k3 = k % 3, kq = k / 3;
  const float offk = (float)(k - kR);
  const int gl = g < 3 ? g : 2;
  const float sc01 = 1.0f / (float)(1 << gl);
  const float dx0 = p.lv[gl].dx, rx0 = p.lv[gl].rx, sx0 = p.lv[gl].sx;
  const float dy0 = p.lv[gl].dy, ry0 = p.lv[gl].ry, sy0 = p.lv[gl].sy;
  const bool yl3 = (g == 1);                       // round 2: lanes 0..8 = x, 9..17 = y of level 3
  const float d3 = yl3 ? p.lv[3].dy : p.lv[3].dx, r3 = yl3 ? p.lv[3].ry : p.lv[3].rx;
  const float s3 = yl3 ? p.lv[3].sy : p.lv[3].sx;
  // record addresses of this lane (buffer 0): phase A writes, phase B reads the same slots
  const uint32_t xrec01 = recs + (g * kD + k) * 8;                        // level g, tap k
  const uint32_t yrec01 = recs + kXRecBytes + (g * 3 + kq) * 16;          // level g, group k/3
  const uint32_t xrec3 = recs + (3 * kD + k) * 8;
  const uint32_t yrec3 = recs + kXRecBytes + (3 * 3 + kq) * 16;
  // phase B: lane = 9*gb + a with gb = g, a = k
  const uint32_t xrecB = recs + k * 8, yrecB = recs + kXRecBytes + g * 16;

  // floor clamped to [-32, size + 32] (far-out windows are all padding anyway; NaN -> -32)
  auto iclamp = [](float fl, float sizem1) { return (int)fminf(fmaxf(fl, -32.0f), sizem1 + 33.0f); };

  // y-tap record of one round: w1 of every tap; the group's first lane adds the row offset of its
  // tap and the row deltas of the next two.  Returns true where a group is not consecutive.
  auto put_y = [&](bool act, uint32_t rec, int rel, float w1, uint32_t tile_off) {
    const int r1 = __shfl_down_sync(0xffffffffu, rel, 1);
    const int r2 = __shfl_down_sync(0xffffffffu, rel, 2);
    bool bad = false;
    if (act) {
      *reinterpret_cast<float*>(smem_raw + rec + 4 + 4 * k3) = w1;
      if (k3 == 0) {
        const int d1 = r1 - rel, d2 = r2 - rel;
        *reinterpret_cast<uint32_t*>(smem_raw + rec) =
            (tile_off + (uint32_t)rel * kRowBytes) | ((uint32_t)d1 << 16) | ((uint32_t)d2 << 20);
        bad = (d1 != 1) | (d2 != 2);
      }
    }
    return bad;
  };

  // Phase A + TMA issue for the query at `c` (flow fx, fy) into buffer b.  Returns (warp-uniform)
  // the levels (bit l) whose y taps are not consecutive and need the per-tap form of phase B.
  auto prepare = [&](const Cursor& c, float fx, float fy, int b) -> unsigned {
    const int y = (int)__umulhi((unsigned)c.qi, p.magic_w);
    const int x = c.qi - y * p.w;
    const float xf = __fadd_rn((float)x, fx), yf = __fadd_rn((float)y, fy);
    const uint32_t rb = b * kRecBytes;
    short* win = WIN ? p.windows + (size_t)c.q * (kLevels * 2 * kD) : nullptr;
    // round 0: x taps of level g
    const float ivx = unnorm_coord<DIV>(__fmul_rn(xf, sc01), offk, dx0, rx0, sx0);
    const float flx = floorf(ivx);
    const int ix = iclamp(flx, sx0);
    const int orgx = (__shfl_sync(0xffffffffu, ix, grp0) - 1) & ~3;     // 16-byte aligned column
    if (lane < 27) {
      uint2 rec;
      rec.x = (uint32_t)(ix - orgx) * 4u;
      rec.y = __float_as_uint(__fsub_rn(ivx, flx));                     // ATen: ix - ix_nw
      *reinterpret_cast<uint2*>(smem_raw + xrec01 + rb) = rec;
      if (WIN) win[18 * g + k] = (short)ix;
    }
    // round 1: y taps of level g
    const float ivy = unnorm_coord<DIV>(__fmul_rn(yf, sc01), offk, dy0, ry0, sy0);
    const float fly = floorf(ivy);
    const int iy = iclamp(fly, sy0);
    const int orgy = __shfl_sync(0xffffffffu, iy, grp0) - 1;
    const bool bad01 = put_y(lane < 27, yrec01 + rb, iy - orgy, __fsub_rn(ivy, fly), (uint32_t)g * kTileBytes);
    if (WIN && lane < 27) win[18 * g + kD + k] = (short)iy;
    // round 2: level 3, lanes 0..8 x, lanes 9..17 y
    const float iv3 = unnorm_coord<DIV>(__fmul_rn(yl3 ? yf : xf, 0.125f), offk, d3, r3, s3);
    const float fl3 = floorf(iv3);
    const int i3 = iclamp(fl3, s3);
    const int t3 = __shfl_sync(0xffffffffu, i3, grp0) - 1;
    const int org3 = yl3 ? t3 : (t3 & ~3);
    const float w13 = __fsub_rn(iv3, fl3);
    if (lane < kD) {
      uint2 rec;
      rec.x = (uint32_t)(i3 - org3) * 4u;
      rec.y = __float_as_uint(w13);
      *reinterpret_cast<uint2*>(smem_raw + xrec3 + rb) = rec;
    }
    const bool bad3 = put_y(lane >= kD && lane < 2 * kD, yrec3 + rb, i3 - org3, w13, 3u * kTileBytes);
    if (WIN && lane < 2 * kD) win[18 * 3 + lane] = (short)i3;
    // lanes 9g .. 9g+8 hold level g in round 1 (zero flow = integer coordinates makes level 0 slow
    // for every background query, level l only where y % 2^l == 0)
    const unsigned b01 = __ballot_sync(0xffffffffu, bad01);
    const unsigned slow = ((b01 & 0x1ffu) ? 1u : 0u) | ((b01 & (0x1ffu << 9)) ? 2u : 0u) | ((b01 & (0x1ffu << 18)) ? 4u : 0u) |
                          (__any_sync(0xffffffffu, bad3) ? 8u : 0u);
    // ---- staging: the bulk store of two queries ago used tile buffer b as its staging line
    const int ox3 = __shfl_sync(0xffffffffu, org3, 0), oy3 = __shfl_sync(0xffffffffu, org3, kD);
    if (lane == 0) {
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      tc::mbar_expect_tx(&bars_gen[b], (uint32_t)kBufBytes);
    }
    __syncwarp();
    if (k == 0 && lane <= 27) {
      uint8_t* dst = smem_raw + tiles + b * kBufBytes + g * kTileBytes;
      tc::tma_load_3d(&maps.m[g], &bars_gen[b], dst, g < 3 ? orgx : ox3, g < 3 ? orgy : oy3, c.q);
    }
    return slow;
  };

  const int q_first = blockIdx.x * WARPS + warp;
  Cursor c0{q_first, q_first / P, q_first % P};
  Cursor c1 = advance(c0), c2 = advance(c1);
  float fx0, fy0, fx1, fy1;
  load_flow(c0, fx0, fy0);
  load_flow(c1, fx1, fy1);
  unsigned slow0 = 0, slow1 = 0;
  if (c0.q < total) slow0 = prepare(c0, fx0, fy0, 0);

  for (uint32_t it = 0; c0.q < total; ++it) {
    const int b = it & 1;
    if (c1.q < total) slow1 = prepare(c1, fx1, fy1, b ^ 1);      // records + TMA for the next query
    float fx2, fy2;
    load_flow(c2, fx2, fy2);
    __syncwarp();
    tc::mbar_wait(&bars_gen[b], (it >> 1) & 1);

    // ---- phase B: lane = 9*gb + a owns taps (a, b = 3*gb + j), j = 0..2.  Horizontal
    // interpolation once per tile row, vertical per tap (separable form of the four-corner sum:
    // equal to ATen's up to fp32 rounding order; same order as the first-generation kernel).
    float res[kLevels][3];
    const uint32_t tb = tiles + b * kBufBytes, rb = b * kRecBytes;
    if (lane < 27) {
#pragma unroll
      for (int l = 0; l < kLevels; ++l) {
        const uint2 xr = *reinterpret_cast<const uint2*>(smem_raw + xrecB + rb + l * (kD * 8));
        const uint4 yr = *reinterpret_cast<const uint4*>(smem_raw + yrecB + rb + l * (3 * 16));
        const uint32_t a0 = tb + (yr.x & 0xffffu) + xr.x;
        const float wx1 = __uint_as_float(xr.y), wx0 = __fsub_rn(1.0f, wx1);
        const float v0 = __uint_as_float(yr.y), v1 = __uint_as_float(yr.z), v2 = __uint_as_float(yr.w);
        if (!((slow0 >> l) & 1u)) {                   // consecutive floors: four shared tile rows
          const float h0 = fmaf(ldf(a0 + 4), wx1, ldf(a0) * wx0);
          const float h1 = fmaf(ldf(a0 + kRowBytes + 4), wx1, ldf(a0 + kRowBytes) * wx0);
          const float h2 = fmaf(ldf(a0 + 2 * kRowBytes + 4), wx1, ldf(a0 + 2 * kRowBytes) * wx0);
          const float h3 = fmaf(ldf(a0 + 3 * kRowBytes + 4), wx1, ldf(a0 + 3 * kRowBytes) * wx0);
          res[l][0] = fmaf(h1, v0, h0 * __fsub_rn(1.0f, v0));
          res[l][1] = fmaf(h2, v1, h1 * __fsub_rn(1.0f, v1));
          res[l][2] = fmaf(h3, v2, h2 * __fsub_rn(1.0f, v2));
        } else {                                      // per-tap rows
          const float wy[3] = {v0, v1, v2};
          const uint32_t dr[3] = {0u, (yr.x >> 16) & 15u, (yr.x >> 20) & 15u};
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            const uint32_t a = a0 + dr[j] * kRowBytes;
            const float top = fmaf(ldf(a + 4), wx1, ldf(a) * wx0);
            const float bot = fmaf(ldf(a + kRowBytes + 4), wx1, ldf(a + kRowBytes) * wx0);
            res[l][j] = fmaf(bot, wy[j], top * __fsub_rn(1.0f, wy[j]));
          }
        }
      }
    }
    __syncwarp();          // every lane is done reading tile buffer b: it becomes the staging line
    if (SPLIT) {
      // split planes for the tensor-core 1x1 convolution that consumes the lookup (raft_decoder.py:152-155):
      // 328 bf16 hi, then 328 bf16 lo (324 values + 4 zero channels each), two 656-byte bulk copies
      // Channel order of the planes: LANE-MAJOR, c' = 12 * lane + 3 * level + j for the value of channel
      // 81 * level + 9 * a + 3 * g + j (lane = 9 g + a).  The consumer is a 1x1 convolution whose input channels
      // are permuted to match when its filter is packed (ops.LOOKUP_SPLIT_ORDER), so a lane's twelve values leave
      // as three 8-byte stores per plane with packed conversions instead of 24 two-byte stores (the natural order
      // cost 7 us of the 55).
      uint8_t* so = smem_raw + tb;
      if (lane < 27) {
        uint32_t hi[6], lo[6];
#pragma unroll
        for (int t = 0; t < 6; ++t) {
          const float v0 = res[(2 * t) / 3][(2 * t) % 3], v1 = res[(2 * t + 1) / 3][(2 * t + 1) % 3];
          hi[t] = cvt_bf16x2(v0, v1);
          lo[t] = cvt_bf16x2(v0 - __uint_as_float(hi[t] << 16), v1 - __uint_as_float(hi[t] & 0xffff0000u));
        }
        uint2* ph = reinterpret_cast<uint2*>(so + 24 * lane);
        uint2* pl = reinterpret_cast<uint2*>(so + 2 * kSplitC + 24 * lane);
        ph[0] = make_uint2(hi[0], hi[1]); ph[1] = make_uint2(hi[2], hi[3]); ph[2] = make_uint2(hi[4], hi[5]);
        pl[0] = make_uint2(lo[0], lo[1]); pl[1] = make_uint2(lo[2], lo[3]); pl[2] = make_uint2(lo[4], lo[5]);
      } else if (lane == 27) {
        *reinterpret_cast<uint2*>(so + 2 * kOutFloats) = make_uint2(0u, 0u);               // 4 zero channels
        *reinterpret_cast<uint2*>(so + 2 * kSplitC + 2 * kOutFloats) = make_uint2(0u, 0u);
      }
    } else if (lane < 27) {
      float* so = reinterpret_cast<float*>(smem_raw + tb) + kD * k + 3 * g;
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
      if (SPLIT) {
        const uint8_t* o = reinterpret_cast<const uint8_t*>(p.out) + (size_t)c0.q * (4 * kSplitC);
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(o), "r"(sbase + tb),
                     "n"(2 * kSplitC) : "memory");
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(o + 2 * kSplitC),
                     "r"(sbase + tb + 2 * kSplitC), "n"(2 * kSplitC) : "memory");
      } else {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(
                         p.out + (size_t)c0.q * kOutFloats),
                     "r"(sbase + tb), "n"(kOutFloats * 4)
                     : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    c0 = c1; c1 = c2; c2 = advance(c2);
    fx0 = fx1; fy0 = fy1; fx1 = fx2; fy1 = fy2;
    slow0 = slow1;
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace lk2
}  // namespace scflow

namespace scflow {

// Returns SCFLOW_E_UNSUPPORTED when the shape / alignment is outside what this kernel takes (the
// caller then uses the first-generation kernel), otherwise the launch status.
int launch_corr_lookup_v2(const float* const* levels, const float* flow, float* out, short* windows,
                          int n, int h, int w, int div_mode, bool split, cudaStream_t stream) {
  using namespace lk2;
  if ((w % 32) != 0 || (h >> (kLevels - 1)) < 1) return SCFLOW_E_UNSUPPORTED;
  const long long total_q = (long long)n * h * w;
  if (total_q >= (1ll << 31) || (double)h * w * w >= 4294967296.0) return SCFLOW_E_UNSUPPORTED;
  Maps maps;
  Params p;
  for (int l = 0; l < kLevels; ++l) {
    const int wl = w >> l, hl = h >> l;
    if (!aligned16(levels[l])) return SCFLOW_E_UNSUPPORTED;
    LevelC& L = p.lv[l];
    L.dx = (float)(wl - 1 > 1 ? wl - 1 : 1);
    L.dy = (float)(hl - 1 > 1 ? hl - 1 : 1);
    L.rx = 1.0f / L.dx; L.ry = 1.0f / L.dy;
    L.sx = (float)(wl - 1); L.sy = (float)(hl - 1);
    // the box may exceed the level (4 x 4 at 256^2 crops): the excess is out-of-bounds zero fill
    const uint64_t dims[3] = {(uint64_t)wl, (uint64_t)hl, (uint64_t)total_q};
    const uint64_t strides[2] = {(uint64_t)wl * 4, (uint64_t)wl * hl * 4};
    const uint32_t box[3] = {(uint32_t)kTileW, (uint32_t)kTileH, 1};
    if (!tc::encode_f32_map(&maps.m[l], levels[l], 3, dims, strides, box, tc::kTmaNone, /*l2_promote=*/false))
      return SCFLOW_E_UNSUPPORTED;
  }
  p.flow = flow; p.out = out; p.windows = windows;
  p.h = h; p.w = w; p.total_q = (int)total_q;
  p.magic_w = (w & (w - 1)) == 0 ? (unsigned)(4294967296ull / (unsigned)w) : (unsigned)(4294967296ull / (unsigned)w) + 1u;
  // 8 warps per CTA x 3 CTAs per SM.  Measured at cfg2 on B200: 8 x 3 47.7 us, 7 x 4 51.8 us, 6 x 5 54.9 us,
  // 8 x 2 / 6 x 3 / 12 x 2 / 4 x 6 all 47-48 us: more resident warps lose, fewer change nothing
  constexpr int warps = 8, minb = 3;
  typedef void (*Kern)(const Maps, const Params);
  const bool ieee = div_mode == SCFLOW_DIV_IEEE, win = windows != nullptr;
#define SCFLOW_LK2(D, W_) (split ? (Kern)corr_lookup_v2_kernel<D, warps, minb, W_, true> \
                                : (Kern)corr_lookup_v2_kernel<D, warps, minb, W_, false>)
  Kern kern = ieee ? (win ? SCFLOW_LK2(SCFLOW_DIV_IEEE, true) : SCFLOW_LK2(SCFLOW_DIV_IEEE, false))
                   : (win ? SCFLOW_LK2(SCFLOW_DIV_RCP, true) : SCFLOW_LK2(SCFLOW_DIV_RCP, false));
#undef SCFLOW_LK2
  const int smem_bytes = warps * kWarpBytes;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);   // per device: every launch
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  long long blocks = (total_q + warps - 1) / warps;
  if (blocks > (long long)sms * minb) blocks = (long long)sms * minb;
  const long long nwarps = blocks * warps;
  p.step_q = (int)nwarps; p.step_n = (int)(nwarps / (h * w)); p.step_qi = (int)(nwarps % (h * w));
  kern<<<(unsigned)blocks, warps * 32, smem_bytes, stream>>>(maps, p);
  return check_launch();
}

}  // namespace scflow
