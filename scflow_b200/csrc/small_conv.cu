// The decoder's degenerate convolutions (flow/mask heads, delta-flow and mask
// encoders), which cuDNN serves with generic engines at 65-75 us per call:
//
//   conv_small_cin   Cin in {1,2}, k in {3,7}, Cout % 32 == 0, stride 1, + bias + ReLU
//                    MotionEncoder.flow_net[0] / delta_flow_encoder[0] (7x7, 2 -> 128) and
//                    mask_encoder[0] (3x3, 1 -> 64): raft_decoder.py:152-160, scflow_decoder.py:96-106.
//                    FP32-FMA bound (98 MAC per output): lane = output channel, 8 pixels of a row in
//                    registers, weights [ci][ky][kx][Cout] and the input patch in shared memory.
//   conv_small_cout  Cout in {1,2}, k in {1,3}, channels-last input, + bias
//                    XHead.predict_layer (raft_decoder.py:281-287): flow 3x3 256 -> 2, mask 1x1 256 -> 1.
//                    HBM-bound (reads the (N,h,w,C) feature map once): a CTA owns a band of rows, stages
//                    32-channel slabs through shared memory (coalesced), every thread reduces its own
//                    pixel against all taps, then the taps are gathered from the neighbours.
//
// Exact fp32 accumulation (tighter than the TF32 cuDNN kernels they replace).
#include "common.cuh"
#include <cuda_bf16.h>
#include <stdlib.h>

namespace scflow {
namespace sc {

constexpr int kSegPx = 8;          // minimum pixels per warp task (one row segment); 16 where the width allows
constexpr int kTileW = 32;         // pixels per CTA tile (one row segment of <= 32)

struct CinParams {
  const float* in;        // element (n, ci, y, x) at n*sn + ci*sc + (y*w + x)*sp
  const float* wpack;     // [CIN][KS][KS][Cout]
  const float* bias;      // [Cout] or nullptr
  float* out;             // (n, h, w, Cout) channels-last fp32, or nullptr
  long long sn, sc, sp;
  int n, h, w, cout, relu;
  uint16_t* out_split;    // or/and split-bf16 planes: pixel pitch out_pitch bf16, lo plane out_lo further
  long long out_pitch;
  int out_lo;
};

// SEG pixels of a row per lane: the inputs are broadcast LDS.128, the filter taps one LDS.32 per
// lane; with 8 pixels that is 11 shared-memory reads per 56 FMAs and the two pipes tie (43 % of the
// FP32 rate); 16 pixels halve the reads per FMA.
template <int CIN, int KS, int SEG>
__global__ void __launch_bounds__(256)
conv_small_cin_kernel(const CinParams p) {
  constexpr int PAD = KS / 2;
  constexpr int ROWLEN = kTileW + 8;                 // >= kTileW + KS - 1, multiple of 4
  constexpr int PATCH = CIN * KS * ROWLEN;
  constexpr int PER_THREAD = (PATCH + 255) / 256;
  extern __shared__ __align__(16) float smem[];
  float* wsm = smem;                                 // CIN*KS*KS*cout
  float* patch0 = smem + CIN * KS * KS * p.cout;     // 2 x [CIN][KS][ROWLEN]
  __shared__ __align__(8) uint64_t wbar;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t wbytes = (uint32_t)(CIN * KS * KS * p.cout * sizeof(float));
  if (threadIdx.x == 0) {
    // the filter bank (up to 50 KB) arrives as ONE bulk async copy while the first patch is fetched
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&wbar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(wbytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(wsm)),
                 "l"(p.wpack), "r"(wbytes), "r"(bar)
                 : "memory");
  }
  const int xblocks = (p.w + kTileW - 1) / kTileW;
  const long long tiles = (long long)p.n * p.h * xblocks;
  const int groups = p.cout >> 5;
  float pre[PER_THREAD];
  auto fetch = [&](long long tile) {                 // this thread's share of a tile's input patch
    const int xb = (int)(tile % xblocks);
    const int y = (int)((tile / xblocks) % p.h);
    const int n = (int)(tile / ((long long)xblocks * p.h));
#pragma unroll
    for (int k = 0; k < PER_THREAD; ++k) {
      const int i = threadIdx.x + 256 * k;
      const int col = i % ROWLEN, ky = (i / ROWLEN) % KS, ci = i / (ROWLEN * KS);
      const int yy = y + ky - PAD, xx = xb * kTileW + col - PAD;
      float v = 0.0f;
      if (i < PATCH && col < kTileW + KS - 1 && yy >= 0 && yy < p.h && xx >= 0 && xx < p.w)
        v = __ldg(p.in + n * p.sn + ci * p.sc + ((long long)yy * p.w + xx) * p.sp);
      pre[k] = v;
    }
  };
  auto stash = [&](float* dst) {
#pragma unroll
    for (int k = 0; k < PER_THREAD; ++k)
      if (threadIdx.x + 256 * k < PATCH) dst[threadIdx.x + 256 * k] = pre[k];
  };
  int buf = 0;
  if (blockIdx.x < tiles) { fetch(blockIdx.x); stash(patch0); }
  __syncthreads();                                   // barrier initialised, first patch in place
  {
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&wbar);
    uint32_t done;
    do {
      asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], 0;\n\t"
                   "selp.b32 %0, 1, 0, P1;\n\t}" : "=r"(done) : "r"(bar) : "memory");
    } while (!done);
  }
  for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int xb = (int)(tile % xblocks);
    const int y = (int)((tile / xblocks) % p.h);
    const int n = (int)(tile / ((long long)xblocks * p.h));
    const int x0 = xb * kTileW;
    const float* patch = patch0 + buf * PATCH;
    const bool more = tile + gridDim.x < tiles;
    if (more) fetch(tile + gridDim.x);               // next tile's loads fly under this tile's FMAs
    const int segs = min(kTileW, p.w - x0) / SEG;
    for (int task = warp; task < segs * groups; task += 8) {
      const int seg = task % segs, grp = task / segs;
      const int c = grp * 32 + lane;
      float acc[SEG];
      const float b = p.bias ? __ldg(p.bias + c) : 0.0f;
#pragma unroll
      for (int i = 0; i < SEG; ++i) acc[i] = b;
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci) {
#pragma unroll
        for (int ky = 0; ky < KS; ++ky) {
          float xr[SEG + 8];
          const float4* src = reinterpret_cast<const float4*>(patch + (ci * KS + ky) * ROWLEN + seg * SEG);
#pragma unroll
          for (int j = 0; j < (SEG + KS - 1 + 3) / 4; ++j) {
            const float4 v = src[j];                 // same address in every lane: broadcast
            xr[4 * j] = v.x; xr[4 * j + 1] = v.y; xr[4 * j + 2] = v.z; xr[4 * j + 3] = v.w;
          }
          const float* wrow = wsm + ((ci * KS + ky) * KS) * p.cout + c;
#pragma unroll
          for (int kx = 0; kx < KS; ++kx) {
            const float wv = wrow[kx * p.cout];
#pragma unroll
            for (int i = 0; i < SEG; ++i) acc[i] = fmaf(xr[kx + i], wv, acc[i]);
          }
        }
      }
      const long long pix0 = ((long long)n * p.h + y) * p.w + x0 + seg * SEG;
      if (p.relu) {
#pragma unroll
        for (int i = 0; i < SEG; ++i) acc[i] = fmaxf(acc[i], 0.0f);
      }
      if (p.out) {
        float* o = p.out + pix0 * p.cout + c;
#pragma unroll
        for (int i = 0; i < SEG; ++i) o[(long long)i * p.cout] = acc[i];
      }
      if (p.out_split) {        // the next layer is a tensor-core convolution: its operand, no fp32 round trip
        uint16_t* o = p.out_split + pix0 * p.out_pitch + c;
#pragma unroll
        for (int i = 0; i < SEG; i += 2) {
          uint32_t hi, lo;
          split2(acc[i], acc[i + 1], hi, lo);
          o[(long long)i * p.out_pitch] = (uint16_t)hi;
          o[(long long)i * p.out_pitch + p.out_lo] = (uint16_t)lo;
          o[(long long)(i + 1) * p.out_pitch] = (uint16_t)(hi >> 16);
          o[(long long)(i + 1) * p.out_pitch + p.out_lo] = (uint16_t)(lo >> 16);
        }
      }
    }
    if (more) stash(patch0 + (buf ^ 1) * PATCH);
    __syncthreads();                                 // next patch visible; this one free for reuse
    buf ^= 1;
  }
}

template <int CIN, int KS>
static int launch_cin(const CinParams& p, cudaStream_t st) {
  const size_t smem = ((size_t)CIN * KS * KS * p.cout + 2 * (size_t)CIN * KS * (kTileW + 8)) * sizeof(float);
  if (smem > 48 * 1024) {                              // per device: set on every launch
    cudaFuncSetAttribute(conv_small_cin_kernel<CIN, KS, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(conv_small_cin_kernel<CIN, KS, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  const long long tiles = (long long)p.n * p.h * ((p.w + kTileW - 1) / kTileW);
  const int per_sm = smem > 100 * 1024 ? 2 : smem > 56 * 1024 ? 3 : 4;
  long long grid = tiles < 148LL * per_sm ? tiles : 148LL * per_sm;
  const bool wide = (p.w % 16) == 0;                  // 16 pixels per lane: 55 -> 49 us on the 7x7 2->128 layer
  if (wide) conv_small_cin_kernel<CIN, KS, 16><<<(unsigned)grid, 256, smem, st>>>(p);
  else conv_small_cin_kernel<CIN, KS, 8><<<(unsigned)grid, 256, smem, st>>>(p);
  return check_launch();
}

}  // namespace sc
}  // namespace scflow

extern "C" int scflow_conv_small_cin_fwd(const float* in, long long stride_n, long long stride_c,
                                         long long stride_px, const float* wpack, const float* bias,
                                         float* out, void* out_split, long long out_pitch, int out_lo, int n,
                                         int cin, int h, int w, int cout, int ksize, int relu, void* stream) {
  using namespace scflow;
  if (!in || !wpack || (!out && !out_split)) return SCFLOW_E_NULL;
  if (n <= 0 || h <= 0 || w <= 0 || (w % sc::kSegPx) != 0 || cout <= 0 || (cout & 31) != 0 || cout > 256)
    return SCFLOW_E_BAD_SHAPE;
  sc::CinParams p{in, wpack, bias, out, stride_n, stride_c, stride_px, n, h, w, cout, relu,
                  static_cast<uint16_t*>(out_split), out_pitch, out_lo};
  cudaStream_t st = (cudaStream_t)stream;
  if (cin == 2 && ksize == 7) return sc::launch_cin<2, 7>(p, st);
  if (cin == 1 && ksize == 3) return sc::launch_cin<1, 3>(p, st);
  if (cin == 2 && ksize == 3) return sc::launch_cin<2, 3>(p, st);
  if (cin == 1 && ksize == 7) return sc::launch_cin<1, 7>(p, st);
  return SCFLOW_E_UNSUPPORTED;
}

// ---------------------------------------------------------------------------------------------
namespace scflow {
namespace sc {

constexpr int kBandRows = 4;       // output rows per CTA
constexpr int kSlab = 32;          // channels staged per step

struct CoutParams {
  const float* in;        // (n, h, w, C) channels-last
  const float* wpack;     // [C][KS*KS*COUT]
  const float* bias;      // [COUT] or nullptr
  float* out;             // (n, COUT, h, w) planar
  int n, h, w, c;
};

template <int KS, int COUT>
__global__ void __launch_bounds__(1024)
conv_small_cout_kernel(const CoutParams p) {
  constexpr int PAD = KS / 2;
  constexpr int NT = KS * KS * COUT;                 // partial sums per pixel
  constexpr int NT4 = (NT + 3) & ~3;                 // padded so a channel's taps are float4 loads
  extern __shared__ __align__(16) float smem[];
  float* wsm = smem;                                 // [C][NT4]
  float* T = wsm + p.c * NT4;                        // [pixels][NT]
  const int bands = (p.h + kBandRows - 1) / kBandRows;
  const int n = blockIdx.x / bands, band = blockIdx.x % bands;
  const int y0 = band * kBandRows - PAD;             // first (halo) row of the band
  const int tid = threadIdx.x;
  for (int i = tid; i < p.c * NT4; i += blockDim.x) {
    const int ch = i / NT4, j = i - ch * NT4;
    wsm[i] = j < NT ? __ldg(p.wpack + ch * NT + j) : 0.0f;
  }
  __syncthreads();
  const int r = tid / p.w, x = tid - r * p.w;
  const int y = y0 + r;
  float acc[NT4];
#pragma unroll
  for (int j = 0; j < NT4; ++j) acc[j] = 0.0f;
  if (y >= 0 && y < p.h) {
    // a thread walks its own pixel's channels (16 B per step; the 128-byte lines stay in L1 for the
    // 8 steps that consume them), weights are broadcast float4 reads
    const float4* src = reinterpret_cast<const float4*>(p.in + (((long long)n * p.h + y) * p.w + x) * p.c);
#pragma unroll 2
    for (int c4 = 0; c4 < p.c / 4; ++c4) {
      const float4 v = __ldg(src + c4);
      const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float4* wv = reinterpret_cast<const float4*>(wsm + (c4 * 4 + k) * NT4);
#pragma unroll
        for (int j = 0; j < NT4 / 4; ++j) {
          const float4 w4 = wv[j];
          acc[4 * j] = fmaf(vv[k], w4.x, acc[4 * j]);
          acc[4 * j + 1] = fmaf(vv[k], w4.y, acc[4 * j + 1]);
          acc[4 * j + 2] = fmaf(vv[k], w4.z, acc[4 * j + 2]);
          acc[4 * j + 3] = fmaf(vv[k], w4.w, acc[4 * j + 3]);
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < NT; ++j) T[tid * NT + j] = acc[j];
  __syncthreads();
  if (r >= PAD && r < PAD + kBandRows && y < p.h) {
#pragma unroll
    for (int co = 0; co < COUT; ++co) {
      float s = p.bias ? __ldg(p.bias + co) : 0.0f;
#pragma unroll
      for (int ky = 0; ky < KS; ++ky)
#pragma unroll
        for (int kx = 0; kx < KS; ++kx) {
          const int xx = x + kx - PAD;
          if (xx >= 0 && xx < p.w) s += T[((r + ky - PAD) * p.w + xx) * NT + (ky * KS + kx) * COUT + co];
        }
      p.out[(((long long)n * COUT + co) * p.h + y) * p.w + x] = s;
    }
  }
}

template <int KS, int COUT>
static int launch_cout(const CoutParams& p, cudaStream_t st) {
  constexpr int NT = KS * KS * COUT;
  const int threads = (kBandRows + 2 * (KS / 2)) * p.w;
  if (threads > 1024) return SCFLOW_E_BAD_SHAPE;
  const size_t smem = ((size_t)p.c * ((NT + 3) & ~3) + (size_t)threads * NT) * sizeof(float);
  if (smem > 200 * 1024) return SCFLOW_E_BAD_SHAPE;
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(conv_small_cout_kernel<KS, COUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int bands = (p.h + kBandRows - 1) / kBandRows;
  conv_small_cout_kernel<KS, COUT><<<(unsigned)(p.n * bands), threads, smem, st>>>(p);
  return check_launch();
}

}  // namespace sc
}  // namespace scflow

// ---------------------------------------------------------------------------------------------
// TMA-staged form of conv_small_cout_kernel.  The plain kernel reads its pixel's channels with
// 16-byte loads at a 1 KB lane stride: 32 tag look-ups per load instruction, and it is bound by
// the L1/TEX pipe (ncu: 69 %, 73 us for a 67 MB input; neither register-blocking nor
// line-at-a-time loads moved it).  Here the band's activations arrive as [pixels][32 channels]
// slabs through TMA with the 128-byte swizzle -- a thread reads its pixel's 128-byte row with
// eight conflict-free LDS.128 (chunk j of row r sits at chunk j ^ (r & 7)) -- the halo rows
// outside the image are TMA's zero fill, and PX pixels per thread share every filter read.
#include "tc_common.cuh"
namespace scflow {
namespace sc {

constexpr int kSlabCh = 32;        // channels per slab: one 128-byte swizzle row per pixel
constexpr int kSlabStages = 3;

template <int KS, int COUT, int PX>
__global__ void __launch_bounds__(256)
conv_small_cout_tma_kernel(const __grid_constant__ CUtensorMap map_in, const CoutParams p) {
  constexpr int PAD = KS / 2;
  constexpr int NT = KS * KS * COUT;                 // partial sums per pixel
  constexpr int NT4 = (NT + 3) & ~3;
  constexpr int ROWS = kBandRows + 2 * PAD;          // band rows incl. halo
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int npx = ROWS * p.w;                        // pixels (tile rows) per slab
  const int slab_bytes = npx * 128;
  float* wsm = reinterpret_cast<float*>(base + kSlabStages * slab_bytes);     // [C][NT4]
  float* T = wsm + p.c * NT4;                                                  // [pixels][NT]
  uint64_t* full = reinterpret_cast<uint64_t*>(T + npx * NT + 4);             // [kSlabStages], 8-byte aligned below
  full = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(full) + 7) & ~(uintptr_t)7);
  const int bands = (p.h + kBandRows - 1) / kBandRows;
  const int n = blockIdx.x / bands, band = blockIdx.x % bands;
  const int y0 = band * kBandRows - PAD;             // first (halo) row of the band
  const int tid = threadIdx.x, nthreads = blockDim.x;
  const int nslabs = p.c / kSlabCh;
  if (tid == 0) {
    for (int i = 0; i < kSlabStages; ++i) tc::mbar_init(&full[i], 1);
    tc::fence_barrier_init();
    tc::tma_prefetch_desc(&map_in);
  }
  __syncthreads();
  if (tid == 0) {
    for (int s = 0; s < kSlabStages && s < nslabs; ++s) {
      tc::mbar_expect_tx(&full[s], (uint32_t)slab_bytes);
      tc::tma_load_4d(&map_in, &full[s], base + s * slab_bytes, s * kSlabCh, 0, y0, n);
    }
  }
  for (int i = tid; i < p.c * NT4; i += nthreads) {
    const int ch = i / NT4, j = i - ch * NT4;
    wsm[i] = j < NT ? __ldg(p.wpack + ch * NT + j) : 0.0f;
  }
  __syncthreads();
  float acc[PX][NT4];
#pragma unroll
  for (int u = 0; u < PX; ++u)
#pragma unroll
    for (int j = 0; j < NT4; ++j) acc[u][j] = 0.0f;
  // thread rows: tid, tid + nthreads, ...  (npx == PX * nthreads)
  for (int s = 0; s < nslabs; ++s) {
    const int st = s % kSlabStages;
    tc::mbar_wait(&full[st], (s / kSlabStages) & 1);
    const uint8_t* slab = base + st * slab_bytes;
    float4 a[PX][8];
#pragma unroll
    for (int u = 0; u < PX; ++u) {
      const int row = tid + u * nthreads;
#pragma unroll
      for (int q = 0; q < 8; ++q)
        a[u][q] = *reinterpret_cast<const float4*>(slab + row * 128 + ((q ^ (row & 7)) << 4));
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float4* wv = reinterpret_cast<const float4*>(wsm + (s * kSlabCh + q * 4 + k) * NT4);
#pragma unroll
        for (int j = 0; j < NT4 / 4; ++j) {
          const float4 w4 = wv[j];
#pragma unroll
          for (int u = 0; u < PX; ++u) {
            const float v = k == 0 ? a[u][q].x : k == 1 ? a[u][q].y : k == 2 ? a[u][q].z : a[u][q].w;
            acc[u][4 * j] = fmaf(v, w4.x, acc[u][4 * j]);
            acc[u][4 * j + 1] = fmaf(v, w4.y, acc[u][4 * j + 1]);
            acc[u][4 * j + 2] = fmaf(v, w4.z, acc[u][4 * j + 2]);
            acc[u][4 * j + 3] = fmaf(v, w4.w, acc[u][4 * j + 3]);
          }
        }
      }
    }
    __syncthreads();                                  // every thread is done with this stage's slab
    if (tid == 0 && s + kSlabStages < nslabs) {
      tc::mbar_expect_tx(&full[st], (uint32_t)slab_bytes);
      tc::tma_load_4d(&map_in, &full[st], base + st * slab_bytes, (s + kSlabStages) * kSlabCh, 0, y0, n);
    }
  }
#pragma unroll
  for (int u = 0; u < PX; ++u) {
    const int row = tid + u * nthreads;
#pragma unroll
    for (int j = 0; j < NT; ++j) T[row * NT + j] = acc[u][j];
  }
  __syncthreads();
#pragma unroll
  for (int u = 0; u < PX; ++u) {
    const int row = tid + u * nthreads;
    const int r = row / p.w, x = row - r * p.w, y = y0 + r;
    if (r >= PAD && r < PAD + kBandRows && y < p.h) {
#pragma unroll
      for (int co = 0; co < COUT; ++co) {
        float sum = p.bias ? __ldg(p.bias + co) : 0.0f;
#pragma unroll
        for (int ky = 0; ky < KS; ++ky)
#pragma unroll
          for (int kx = 0; kx < KS; ++kx) {
            const int xx = x + kx - PAD;
            if (xx >= 0 && xx < p.w) sum += T[((r + ky - PAD) * p.w + xx) * NT + (ky * KS + kx) * COUT + co];
          }
        p.out[(((long long)n * COUT + co) * p.h + y) * p.w + x] = sum;
      }
    }
  }
}

// SCFLOW_E_UNSUPPORTED: shape outside what the staged kernel takes (the caller uses the plain kernel)
template <int KS, int COUT>
static int launch_cout_tma(const CoutParams& p, cudaStream_t st) {
  constexpr int NT = KS * KS * COUT, NT4 = (NT + 3) & ~3, PX = 2;
  constexpr int ROWS = kBandRows + 2 * (KS / 2);
  const int npx = ROWS * p.w;
  if ((p.c % kSlabCh) != 0 || p.w > 256 || (npx % PX) != 0 || npx / PX > 256 || !aligned16(p.in))
    return SCFLOW_E_UNSUPPORTED;
  const size_t smem = 1024 + (size_t)kSlabStages * npx * 128 + ((size_t)p.c * NT4 + (size_t)npx * NT + 4) * sizeof(float) +
                      8 * (kSlabStages + 1);
  if (smem > 200 * 1024) return SCFLOW_E_UNSUPPORTED;
  CUtensorMap map;
  const uint64_t dims[4] = {(uint64_t)p.c, (uint64_t)p.w, (uint64_t)p.h, (uint64_t)p.n};
  const uint64_t strides[3] = {(uint64_t)p.c * 4, (uint64_t)p.w * p.c * 4, (uint64_t)p.h * p.w * p.c * 4};
  const uint32_t box[4] = {(uint32_t)kSlabCh, (uint32_t)p.w, (uint32_t)ROWS, 1};
  if (!tc::encode_f32_map(&map, p.in, 4, dims, strides, box, tc::kTmaSw128)) return SCFLOW_E_UNSUPPORTED;
  cudaFuncSetAttribute(conv_small_cout_tma_kernel<KS, COUT, PX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int bands = (p.h + kBandRows - 1) / kBandRows;
  conv_small_cout_tma_kernel<KS, COUT, PX><<<(unsigned)(p.n * bands), npx / PX, smem, st>>>(map, p);
  return check_launch();
}

}  // namespace sc
}  // namespace scflow

extern "C" int scflow_conv_small_cout_fwd(const float* in, const float* wpack, const float* bias,
                                          float* out, int n, int c, int h, int w, int cout, int ksize,
                                          void* stream) {
  using namespace scflow;
  if (!in || !wpack || !out) return SCFLOW_E_NULL;
  if (n <= 0 || h <= 0 || w <= 0 || c <= 0 || (c & 3) != 0) return SCFLOW_E_BAD_SHAPE;
  sc::CoutParams p{in, wpack, bias, out, n, h, w, c};
  cudaStream_t st = (cudaStream_t)stream;
  int rc = SCFLOW_E_UNSUPPORTED;
  if (ksize == 3 && cout == 2) rc = sc::launch_cout_tma<3, 2>(p, st);
  else if (ksize == 1 && cout == 1) rc = sc::launch_cout_tma<1, 1>(p, st);
  if (rc != SCFLOW_E_UNSUPPORTED) return rc;
  if (ksize == 3 && cout == 2) return sc::launch_cout<3, 2>(p, st);
  if (ksize == 3 && cout == 1) return sc::launch_cout<3, 1>(p, st);
  if (ksize == 1 && cout == 1) return sc::launch_cout<1, 1>(p, st);
  if (ksize == 1 && cout == 2) return sc::launch_cout<1, 2>(p, st);
  return SCFLOW_E_UNSUPPORTED;
}

// ---------------------------------------------------------------------------------------------
// Stem input of the RAFT encoders (raft_encoder.py:286-288, conv1 = 7x7 stride 2 on 3 channels)
// rewritten as a 4x4 stride-1 convolution over the 2x2 space-to-depth image: this kernel
// produces that image in one pass -- planar (n, c, H, W) in, channels-last (n, H/2+3, W/2+3, 16)
// out with channel = 4*c + 2*by + bx for pixel (2Y+by, 2X+bx) at (Y+2, X+2), zero in the unused
// channels and in the (2, 1) border that is the stride-1 filter's padding.  Replaces torch's
// pixel_unshuffle + pad + layout copy (four passes).
namespace scflow {
namespace sc {
// The image is written as two bf16 planes (hi = rn(v), lo = rn(v - hi)), each (n, H/2+3, W/2+3, 16): the operand
// format of the tensor-core convolution family (conv_tc.cu), which reads four x-neighbours (64 contiguous
// channels) per filter row through an overlapped-window tensor map
__global__ void __launch_bounds__(256)
stem_s2d_split_kernel(const float* __restrict__ x, uint4* __restrict__ out, int n, int c, int H, int W) {
  const int Wo = W / 2 + 3, Ho = H / 2 + 3;
  const long long total = (long long)n * Ho * Wo;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int xo = (int)(i % Wo);
    const long long r = i / Wo;
    const int yo = (int)(r % Ho), img = (int)(r / Ho);
    const int Y = yo - 2, X = xo - 2;
    float v[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = 0.0f;
    if (Y >= 0 && Y < H / 2 && X >= 0 && X < W / 2) {
      for (int ch = 0; ch < c; ++ch) {
        const float* p = x + (((size_t)img * c + ch) * H + 2 * Y) * W + 2 * X;
        const float2 a = __ldg(reinterpret_cast<const float2*>(p));
        const float2 b = __ldg(reinterpret_cast<const float2*>(p + W));
        v[4 * ch + 0] = a.x; v[4 * ch + 1] = a.y; v[4 * ch + 2] = b.x; v[4 * ch + 3] = b.y;
      }
    }
    uint32_t hi[8], lo[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const __nv_bfloat16 ah = __float2bfloat16_rn(v[2 * k]), bh = __float2bfloat16_rn(v[2 * k + 1]);
      const __nv_bfloat16 al = __float2bfloat16_rn(v[2 * k] - __bfloat162float(ah));
      const __nv_bfloat16 bl = __float2bfloat16_rn(v[2 * k + 1] - __bfloat162float(bh));
      hi[k] = (uint32_t)__bfloat16_as_ushort(ah) | ((uint32_t)__bfloat16_as_ushort(bh) << 16);
      lo[k] = (uint32_t)__bfloat16_as_ushort(al) | ((uint32_t)__bfloat16_as_ushort(bl) << 16);
    }
    uint4* oh = out + i * 2;
    uint4* ol = out + (total + i) * 2;
    oh[0] = make_uint4(hi[0], hi[1], hi[2], hi[3]); oh[1] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
    ol[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]); ol[1] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
  }
}
}  // namespace sc
}  // namespace scflow

extern "C" int scflow_stem_s2d_split_fwd(const float* x, void* out, int n, int c, int h, int w, void* stream) {
  using namespace scflow;
  if (!x || !out) return SCFLOW_E_NULL;
  if (n <= 0 || c <= 0 || c > 4 || h <= 0 || w <= 0 || (h & 1) || (w & 1)) return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(out) || (reinterpret_cast<uintptr_t>(x) & 7u)) return SCFLOW_E_MISALIGNED;
  const long long total = (long long)n * (h / 2 + 3) * (w / 2 + 3);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  sc::stem_s2d_split_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, reinterpret_cast<uint4*>(out), n, c, h, w);
  return check_launch();
}
