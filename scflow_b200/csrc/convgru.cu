// ConvGRU (SeqConv) pass as ONE fused tcgen05 implicit-GEMM kernel.
// Replaces models/decoder/raft_decoder.py:245-252 for one (1,5) or (5,1) pass: 3 cuDNN convs,
// 3 activation kernels, 2 cat and 5 elementwise kernels of the reference.
//
//   hx = cat[h, x];  z = sigmoid(Wz*hx + bz);  r = sigmoid(Wr*hx + br)
//   q  = tanh(Wq*cat[r.h, x] + bq);            h' = (1 - z).h + z.q
//
// Data layout: channels-last fp32 activations (n, H, W, C): a pixel's channels are contiguous, so
// an activation tile is a K-major GEMM operand that TMA loads as 128 pixels x 32 channels with
// SWIZZLE_128B.  Weights are pre-packed to [tap][k-block][384 rows = z|r|q][32] (K-major too).
//
// A tile is 128 pixels that span the WHOLE convolution axis (pass 0: 128/W full rows; pass 1:
// 128/H full columns).  The 5 taps are 5 shifted TMA loads of the same tile, the zero padding is
// TMA's out-of-bounds fill, and -- because no tap ever leaves the tile's lines -- r.h of the
// tile is all the second convolution needs, so both convolutions and all gating fuse:
//
//   stage 1  D[:,   0:256] = sum_{tap,kb<12} hx_tap . [Wz|Wr]^T       (N = 256)
//            D[:, 256:384] = sum_{tap,kb>=4}  x_tap . Wq[:, x part]^T  (N = 128)
//   epi 1    r = sigmoid(D[:,128:256] + br);  r.h -> scratch (64 KB per tile, stays in L2)
//   stage 2  D[:, 256:384] += sum_{tap,kb<4} (r.h)_tap . Wq[:, h part]^T
//   epi 2    z = sigmoid(D[:,0:128] + bz); q = tanh(D[:,256:384] + bq); h' = (1-z)h + zq
//
// Accumulators live in TMEM (384 of 512 columns); tcgen05.mma kind::tf32 reads the fp32 smem
// tiles directly.  Warp roles: warp 0 TMA producer, warp 1 MMA issuer, warps 2-9 epilogue (two per TMEM lane quarter).
// Tensor-bound: 3 convs x 2*pixels*128*1920 FLOP per pass (96.6 GFLOP for 64 crops at 32x32).
#include "common.cuh"
#include "tc_common.cuh"
#include <stdlib.h>

namespace scflow {
namespace gru {

constexpr int CH = 128;                 // hidden channels
constexpr int CX = 256;                 // x channels (context 128 + motion 126 + flow 2)
constexpr int TAPS = 5;
constexpr int BM = 128;                 // pixels per tile
constexpr int BK = 32;                  // channels per k-block (128-byte swizzle row)
constexpr int KB_H = CH / BK;           // 4 k-blocks from h (or r.h)
constexpr int KB_ALL = (CH + CX) / BK;  // 12 k-blocks of hx
constexpr int ROWS = 3 * CH;            // 384 packed weight rows per (tap, k-block): z | r | q
constexpr int STAGES = 3;
constexpr int A_BYTES = BM * BK * 4;    // 16 KB
constexpr int B_BYTES = ROWS * BK * 4;  // 48 KB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int EPI_WARPS = 8;            // two warps per TMEM lane quarter, 64 channels each
constexpr int THREADS = 64 + 32 * EPI_WARPS;   // TMA warp, MMA warp, epilogue warps
constexpr int TMEM_COLS = 512;
constexpr int S1_ITERS = TAPS * KB_ALL;   // 60
constexpr int S2_ITERS = TAPS * KB_H;     // 20
constexpr size_t WPACK_FLOATS = (size_t)TAPS * KB_ALL * ROWS * BK + ROWS;
constexpr size_t SMEM_BYTES = 1024 + (size_t)STAGES * STAGE_BYTES + ROWS * 4 + 256;

// wz/wr/wq: (CH, CH+CX, 1, 5) or (CH, CH+CX, 5, 1) -- both are [o][ci][tap] in memory.
__global__ void pack_weights_kernel(const float* __restrict__ wz, const float* __restrict__ wr,
                                    const float* __restrict__ wq, const float* __restrict__ bz,
                                    const float* __restrict__ br, const float* __restrict__ bq,
                                    float* __restrict__ out) {
  const int cin = CH + CX;
  const size_t total = (size_t)TAPS * KB_ALL * ROWS * BK;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total + ROWS;
       i += (size_t)gridDim.x * blockDim.x) {
    if (i >= total) {
      const int r = (int)(i - total);
      out[i] = r < CH ? bz[r] : r < 2 * CH ? br[r - CH] : bq[r - 2 * CH];
      continue;
    }
    const int c = (int)(i % BK);
    const int row = (int)((i / BK) % ROWS);
    const int kb = (int)((i / ((size_t)BK * ROWS)) % KB_ALL);
    const int t = (int)(i / ((size_t)BK * ROWS * KB_ALL));
    const float* w = row < CH ? wz : row < 2 * CH ? wr : wq;
    const int o = row % CH, ci = kb * BK + c;
    // round-to-nearest TF32 here (free): tcgen05 would otherwise truncate the low 13 mantissa bits
    float v = w[((size_t)o * cin + ci) * TAPS + t];
    uint32_t u;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
    out[i] = __uint_as_float(u);
  }
}

struct Params {
  const float* h_in;
  const float* bias;     // [384] z | r | q (tail of wpack)
  float* h_out;
  float* h_out_tf32;     // optional copy of h_out rounded to nearest TF32: the next pass's MMA operand
  float* rh;
  int n, H, W, pass;
  int tiles_per_img, total_tiles, G;   // G = lines per tile (128 / conv-axis length)
  int ni;                              // images per tile (>1 only when H*W < 128: whole-image tiles)
  long long total_pix;
};

// sigmoid / tanh from one ex2.approx and one rcp.approx each (a few ulp: ~1e-6, three orders of
// magnitude below the TF32 rounding of the pre-activations they are applied to); the epilogue
// is MUFU/issue-bound, so the libm forms (IEEE division, tanhf) cost 3x the time for nothing
__device__ __forceinline__ float sigmoidf_(float v) { return __fdividef(1.0f, 1.0f + __expf(-v)); }
__device__ __forceinline__ float tanhf_(float v) {
  const float e = __expf(-2.0f * fabsf(v));                 // in (0, 1]: no overflow
  return copysignf(__fdividef(1.0f - e, 1.0f + e), v);
}
// tcgen05 kind::tf32 ignores the low 13 mantissa bits (truncation, a biased error that adds up
// coherently over K = 1920); operands we produce ourselves are rounded to nearest instead
__device__ __forceinline__ float rn_tf32(float v) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
  return __uint_as_float(u);
}

// Both epilogues of one 128-pixel tile for one warp: 32 pixels (= its TMEM lane quarter) x 64
// channels (`half` selects which).  The pixel's hidden-state values are fetched ONCE, before the
// wait for stage 1, so the strided loads hide behind the MMA main loop and serve both the r.h
// product and the final gate.
template <bool CLUSTER>
__device__ __forceinline__ void epilogue_tile(const Params& p, const float* sbias, uint32_t taddr, int off,
                                              bool live, int half, int lane, uint64_t* s1_done,
                                              uint64_t* s2_done, uint64_t* rh_ready, uint32_t parity) {
  constexpr int HC = CH / 2;
  const int cbase = half * HC;
  float h[HC];
  {
    const float4* hp = reinterpret_cast<const float4*>(p.h_in + (live ? off : 0) + cbase);
#pragma unroll
    for (int k = 0; k < HC / 4; ++k) {
      const float4 t = live ? __ldg(hp + k) : make_float4(0.f, 0.f, 0.f, 0.f);
      h[4 * k + 0] = t.x; h[4 * k + 1] = t.y; h[4 * k + 2] = t.z; h[4 * k + 3] = t.w;
    }
  }
  // ---- epilogue 1: r = sigmoid(r_pre + br); r.h (rounded to nearest TF32) -> scratch
  if (CLUSTER) tc::mbar_wait_cluster(s1_done, parity); else tc::mbar_wait(s1_done, parity);
  tc::tc_fence_after();
  float* rhp = p.rh + off + cbase;
#pragma unroll
  for (int c = 0; c < HC / 32; ++c) {
    float v[32];
    tc::tmem_ld_32x32(taddr + CH + cbase + c * 32, v);
    tc::tmem_ld_wait();
    if (live) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float4 o;
        o.x = rn_tf32(sigmoidf_(v[4 * k + 0] + sbias[CH + cbase + c * 32 + 4 * k + 0]) * h[c * 32 + 4 * k + 0]);
        o.y = rn_tf32(sigmoidf_(v[4 * k + 1] + sbias[CH + cbase + c * 32 + 4 * k + 1]) * h[c * 32 + 4 * k + 1]);
        o.z = rn_tf32(sigmoidf_(v[4 * k + 2] + sbias[CH + cbase + c * 32 + 4 * k + 2]) * h[c * 32 + 4 * k + 2]);
        o.w = rn_tf32(sigmoidf_(v[4 * k + 3] + sbias[CH + cbase + c * 32 + 4 * k + 3]) * h[c * 32 + 4 * k + 3]);
        *reinterpret_cast<float4*>(rhp + c * 32 + 4 * k) = o;
      }
    }
  }
  __threadfence();                       // r.h visible device-wide (L2) ...
  tc::fence_proxy_async_all();           // ... and ordered before the async-proxy (TMA) reads
  __syncwarp();
  if (lane == 0) tc::mbar_arrive(rh_ready);

  // ---- epilogue 2: z, q, gate
  if (CLUSTER) tc::mbar_wait_cluster(s2_done, parity); else tc::mbar_wait(s2_done, parity);
  tc::tc_fence_after();
  float* op = p.h_out + off + cbase;
  float* opr = p.h_out_tf32 ? p.h_out_tf32 + off + cbase : nullptr;
#pragma unroll
  for (int c = 0; c < HC / 32; ++c) {
    float zv[32], qv[32];
    tc::tmem_ld_32x32(taddr + cbase + c * 32, zv);
    tc::tmem_ld_32x32(taddr + 2 * CH + cbase + c * 32, qv);
    tc::tmem_ld_wait();
    if (live) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int ch = c * 32 + 4 * k + e;
          const float z = sigmoidf_(zv[4 * k + e] + sbias[cbase + ch]);
          const float q = tanhf_(qv[4 * k + e] + sbias[2 * CH + cbase + ch]);
          o[e] = (1.0f - z) * h[ch] + z * q;
        }
        *reinterpret_cast<float4*>(op + c * 32 + 4 * k) = make_float4(o[0], o[1], o[2], o[3]);
        if (opr)
          *reinterpret_cast<float4*>(opr + c * 32 + 4 * k) =
              make_float4(rn_tf32(o[0]), rn_tf32(o[1]), rn_tf32(o[2]), rn_tf32(o[3]));
      }
    }
  }
}

__global__ void __launch_bounds__(THREADS, 1)
convgru_pass_kernel(const __grid_constant__ CUtensorMap map_h, const __grid_constant__ CUtensorMap map_x,
                    const __grid_constant__ CUtensorMap map_rh, const __grid_constant__ CUtensorMap map_w,
                    const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* sbias = reinterpret_cast<float*>(smem + STAGES * STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES + ROWS * 4);
  uint64_t* full = bars;                    // [STAGES]
  uint64_t* empty = bars + STAGES;          // [STAGES]
  uint64_t* s1_done = bars + 2 * STAGES;    // stage-1 accumulators complete
  uint64_t* s2_done = s1_done + 1;          // stage-2 accumulation complete
  uint64_t* rh_ready = s1_done + 2;         // r.h tile written (8 epilogue warps)
  uint64_t* tmem_free = s1_done + 3;        // epilogue finished reading TMEM (8 warps)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s1_done + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&map_h); tc::tma_prefetch_desc(&map_x);
    tc::tma_prefetch_desc(&map_rh); tc::tma_prefetch_desc(&map_w);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
      tc::mbar_init(s1_done, 1); tc::mbar_init(s2_done, 1);
      tc::mbar_init(rh_ready, EPI_WARPS); tc::mbar_init(tmem_free, EPI_WARPS);
      tc::fence_barrier_init();
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  for (int i = threadIdx.x; i < ROWS; i += THREADS) sbias[i] = p.bias[i];
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0; uint32_t it_tile = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it_tile) {
        const int n = (tile / p.tiles_per_img) * p.ni, tl = tile % p.tiles_per_img;
        for (int it = 0; it < S1_ITERS + S2_ITERS; ++it) {
          const bool s2 = it >= S1_ITERS;
          if (it == S1_ITERS) tc::mbar_wait(rh_ready, it_tile & 1);   // r.h of this tile is in L2
          const int j = s2 ? it - S1_ITERS : it;
          const int t = s2 ? j / KB_H : j / KB_ALL;
          const int kb = s2 ? j % KB_H : j % KB_ALL;
          tc::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* a_s = smem + stage * STAGE_BYTES;
          uint8_t* b_s = a_s + A_BYTES;
          const int nb = s2 ? 1 : (kb < KB_H ? 2 : 3);                // 128-row weight boxes
          tc::mbar_expect_tx(&full[stage], A_BYTES + nb * (CH * BK * 4));
          const CUtensorMap* am = s2 ? &map_rh : (kb < KB_H ? &map_h : &map_x);
          const int c0 = (s2 || kb < KB_H) ? kb * BK : (kb - KB_H) * BK;
          // tile origin (x, y) shifted by the tap along the convolution axis; TMA zero-fills
          // everything outside [0,W) x [0,H): that is the conv's zero padding
          const int cx = p.pass == 0 ? (t - 2) : tl * p.G;
          const int cy = p.pass == 0 ? tl * p.G : (t - 2);
          tc::tma_load_4d(am, &full[stage], a_s, c0, cx, cy, n);
          const int wrow = (t * KB_ALL + kb) * ROWS;
          if (s2) {
            tc::tma_load_2d(&map_w, &full[stage], b_s, 0, wrow + 2 * CH);
          } else {
            for (int b = 0; b < nb; ++b)
              tc::tma_load_2d(&map_w, &full[stage], b_s + b * (CH * BK * 4), 0, wrow + b * CH);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t id256 = tc::make_idesc_tf32(BM, 256, 0, 0);   // K-major A and B
      constexpr uint32_t id128 = tc::make_idesc_tf32(BM, 128, 0, 0);
      int stage = 0; uint32_t phase = 0; uint32_t it_tile = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it_tile) {
        tc::mbar_wait(tmem_free, (it_tile & 1) ^ 1);
        tc::tc_fence_after();
        for (int it = 0; it < S1_ITERS + S2_ITERS; ++it) {
          const bool s2 = it >= S1_ITERS;
          const int kb = s2 ? 0 : it % KB_ALL;
          tc::mbar_wait(&full[stage], phase);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
          for (int ks = 0; ks < BK / 8; ++ks) {
            // K-major SWIZZLE_128B: 8-row groups 1024 B apart; a k-step (8 tf32) is +32 B
            const uint64_t da = tc::make_smem_desc(a_addr + ks * 32, 16, 1024, tc::kLayoutSw128);
            if (s2) {
              const uint64_t dq = tc::make_smem_desc(b_addr + ks * 32, 16, 1024, tc::kLayoutSw128);
              tc::mma_tf32_ss(tmem_base + 2 * CH, da, dq, id128, 1u);
            } else {
              const uint64_t dzr = tc::make_smem_desc(b_addr + ks * 32, 16, 1024, tc::kLayoutSw128);
              tc::mma_tf32_ss(tmem_base, da, dzr, id256, (it | ks) != 0 ? 1u : 0u);
              if (kb >= KB_H) {
                const uint64_t dq = tc::make_smem_desc(b_addr + 2 * CH * BK * 4 + ks * 32, 16, 1024, tc::kLayoutSw128);
                tc::mma_tf32_ss(tmem_base + 2 * CH, da, dq, id128, (it != KB_H || ks != 0) ? 1u : 0u);
              }
            }
          }
          tc::tc_commit(&empty[stage]);
          if (it == S1_ITERS - 1) tc::tc_commit(s1_done);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        tc::tc_commit(s2_done);
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue (8 warps)
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;                     // TMEM lane = pixel of the tile
    uint32_t it_tile = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it_tile) {
      const int n = (tile / p.tiles_per_img) * p.ni, tl = tile % p.tiles_per_img;
      long long pix;
      if (p.pass == 0 || p.ni > 1) {
        pix = (long long)n * p.H * p.W + (long long)tl * BM + m;                 // G full rows
      } else {
        const int y = m / p.G, x = tl * p.G + (m % p.G);                          // G full columns
        pix = ((long long)n * p.H + y) * p.W + x;
      }
      const bool live = pix < p.total_pix;       // whole-image tiles may run past the last image
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
      epilogue_tile<false>(p, sbias, taddr, (int)(pix * CH), live, (warp - 2) >> 2, lane, s1_done, s2_done,
                           rh_ready, it_tile & 1);
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(tmem_free);
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace gru
}  // namespace scflow

// ================================================================================================
// 2-CTA variant: a cluster of two CTAs refines a PAIR of 128-pixel tiles as one M = 256 MMA unit
// (tcgen05 cta_group::2).  The single-CTA kernel above streams the full 48 KB weight slab of every
// k-block into every SM and is bound by L2 -> SM bandwidth (6.6 TB/s of TMA traffic, tensor pipe
// 32 % busy); here each CTA loads only HALF of the weight rows (z or r, and half of q) next to
// its own 16 KB activation tile: 40 KB instead of 64 KB per k-block per SM.
//   CTA rank r:  A  = its 128 pixels x 32 channels                     (16 KB)
//                Bzr = rows [128 r, 128 r + 128) of z|r               (16 KB)  -> D[:,   0:256]
//                Bq  = rows [256 + 64 r, 256 + 64 r + 64) of q        ( 8 KB)  -> D[:, 256:384]
// Both CTAs' TMA loads complete on rank 0's `full` barrier; rank 0's MMA warp issues the
// cta_group::2 MMAs and its commits arrive on the `empty` / `done` barriers of BOTH CTAs.  Each
// CTA's epilogue warps read their own TMEM (their 128 pixels) exactly as in the 1-CTA kernel.
namespace scflow {
namespace gru {

constexpr int STAGES2 = 5;
constexpr int BZR_BYTES = CH * BK * 4;        // 16 KB: 128 weight rows
constexpr int BQ_BYTES = (CH / 2) * BK * 4;   //  8 KB: 64 weight rows
constexpr int STAGE2_BYTES = A_BYTES + BZR_BYTES + BQ_BYTES;   // 40 KB
constexpr size_t SMEM2_BYTES = 1024 + (size_t)STAGES2 * STAGE2_BYTES + ROWS * 4 + 256;

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
convgru_pass_2cta_kernel(const __grid_constant__ CUtensorMap map_h, const __grid_constant__ CUtensorMap map_x,
                         const __grid_constant__ CUtensorMap map_rh, const __grid_constant__ CUtensorMap map_w,
                         const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* sbias = reinterpret_cast<float*>(smem + STAGES2 * STAGE2_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES2 * STAGE2_BYTES + ROWS * 4);
  uint64_t* full = bars;                    // [STAGES2]  used in rank 0 only: both CTAs' bytes
  uint64_t* empty = bars + STAGES2;         // [STAGES2]  per CTA, armed by rank 0's commits
  uint64_t* s1_done = bars + 2 * STAGES2;   // per CTA
  uint64_t* s2_done = s1_done + 1;          // per CTA
  uint64_t* rh_ready = s1_done + 2;         // per CTA: its r.h tile is written (8 epilogue warps)
  uint64_t* tmem_free = s1_done + 3;        // rank 0 only: 8 epilogue warps of the pair
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s1_done + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = tc::cluster_ctarank();
  const int pair0 = blockIdx.x >> 1, npairs_grid = gridDim.x >> 1;
  const int total_pairs = (p.total_tiles + 1) >> 1;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&map_h); tc::tma_prefetch_desc(&map_x);
    tc::tma_prefetch_desc(&map_rh); tc::tma_prefetch_desc(&map_w);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES2; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
      tc::mbar_init(s1_done, 1); tc::mbar_init(s2_done, 1);
      tc::mbar_init(rh_ready, EPI_WARPS); tc::mbar_init(tmem_free, 2 * EPI_WARPS);
      tc::fence_barrier_init();
    }
    __syncwarp();
    tc::tmem_alloc_2cta(tmem_slot, TMEM_COLS);
  }
  for (int i = threadIdx.x; i < ROWS; i += THREADS) sbias[i] = p.bias[i];
  tc::tc_fence_before();
  __syncthreads();
  tc::cluster_sync();                 // the peer's barriers exist before anything signals them
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer (both CTAs)
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0; uint32_t it_tile = 0;
      for (int pair = pair0; pair < total_pairs; pair += npairs_grid, ++it_tile) {
        const int tile = 2 * pair + (int)rank;
        const int n = (tile / p.tiles_per_img) * p.ni, tl = tile % p.tiles_per_img;
        for (int it = 0; it < S1_ITERS + S2_ITERS; ++it) {
          const bool s2 = it >= S1_ITERS;
          if (it == S1_ITERS) tc::mbar_wait(rh_ready, it_tile & 1);   // this CTA's r.h tile is in L2
          const int j = s2 ? it - S1_ITERS : it;
          const int t = s2 ? j / KB_H : j / KB_ALL;
          const int kb = s2 ? j % KB_H : j % KB_ALL;
          tc::mbar_wait_cluster(&empty[stage], phase ^ 1);
          uint8_t* a_s = smem + stage * STAGE2_BYTES;
          uint8_t* bzr_s = a_s + A_BYTES;
          uint8_t* bq_s = bzr_s + BZR_BYTES;
          const bool has_zr = !s2, has_q = s2 || kb >= KB_H;
          const uint32_t fullbar = tc::smem_u32(&full[stage]) & tc::kPeerMask;       // rank 0's barrier
          if (rank == 0)                                                             // bytes of BOTH CTAs
            tc::mbar_expect_tx(&full[stage], 2 * (A_BYTES + (has_zr ? BZR_BYTES : 0) + (has_q ? BQ_BYTES : 0)));
          const CUtensorMap* am = s2 ? &map_rh : (kb < KB_H ? &map_h : &map_x);
          const int c0 = (s2 || kb < KB_H) ? kb * BK : (kb - KB_H) * BK;
          const int cx = p.pass == 0 ? (t - 2) : tl * p.G;
          const int cy = p.pass == 0 ? tl * p.G : (t - 2);
          tc::tma_load_4d_2cta(am, fullbar, a_s, c0, cx, cy, n);
          const int wrow = (t * KB_ALL + kb) * ROWS;
          if (has_zr) {                     // 128 rows of z (rank 0) or r (rank 1): two 64-row boxes
            tc::tma_load_2d_2cta(&map_w, fullbar, bzr_s, 0, wrow + (int)rank * CH);
            tc::tma_load_2d_2cta(&map_w, fullbar, bzr_s + BQ_BYTES, 0, wrow + (int)rank * CH + CH / 2);
          }
          if (has_q) tc::tma_load_2d_2cta(&map_w, fullbar, bq_s, 0, wrow + 2 * CH + (int)rank * (CH / 2));
          if (++stage == STAGES2) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer (rank 0 only)
    if (lane == 0 && rank == 0) {
      constexpr uint32_t id256 = tc::make_idesc_tf32(2 * BM, 256, 0, 0);   // M = 256 over the pair
      constexpr uint32_t id128 = tc::make_idesc_tf32(2 * BM, 128, 0, 0);
      int stage = 0; uint32_t phase = 0; uint32_t it_tile = 0;
      for (int pair = pair0; pair < total_pairs; pair += npairs_grid, ++it_tile) {
        tc::mbar_wait_cluster(tmem_free, (it_tile & 1) ^ 1);
        tc::tc_fence_after();
        for (int it = 0; it < S1_ITERS + S2_ITERS; ++it) {
          const bool s2 = it >= S1_ITERS;
          const int kb = s2 ? 0 : it % KB_ALL;
          tc::mbar_wait_cluster(&full[stage], phase);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(smem + stage * STAGE2_BYTES);
          const uint32_t bzr_addr = a_addr + A_BYTES, bq_addr = bzr_addr + BZR_BYTES;
#pragma unroll
          for (int ks = 0; ks < BK / 8; ++ks) {
            const uint64_t da = tc::make_smem_desc(a_addr + ks * 32, 16, 1024, tc::kLayoutSw128);
            const uint64_t dq = tc::make_smem_desc(bq_addr + ks * 32, 16, 1024, tc::kLayoutSw128);
            if (s2) {
              tc::mma_tf32_ss_2cta(tmem_base + 2 * CH, da, dq, id128, 1u);
            } else {
              const uint64_t dzr = tc::make_smem_desc(bzr_addr + ks * 32, 16, 1024, tc::kLayoutSw128);
              tc::mma_tf32_ss_2cta(tmem_base, da, dzr, id256, (it | ks) != 0 ? 1u : 0u);
              if (kb >= KB_H) tc::mma_tf32_ss_2cta(tmem_base + 2 * CH, da, dq, id128, (it != KB_H || ks != 0) ? 1u : 0u);
            }
          }
          tc::tc_commit_2cta(&empty[stage]);
          if (it == S1_ITERS - 1) tc::tc_commit_2cta(s1_done);
          if (++stage == STAGES2) { stage = 0; phase ^= 1; }
        }
        tc::tc_commit_2cta(s2_done);
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue (8 warps per CTA)
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;
    const uint32_t free_bar = tc::smem_u32(tmem_free) & tc::kPeerMask;     // rank 0's barrier
    uint32_t it_tile = 0;
    for (int pair = pair0; pair < total_pairs; pair += npairs_grid, ++it_tile) {
      const int tile = 2 * pair + (int)rank;
      const int n = (tile / p.tiles_per_img) * p.ni, tl = tile % p.tiles_per_img;
      long long pix;
      if (p.pass == 0 || p.ni > 1) {
        pix = (long long)n * p.H * p.W + (long long)tl * BM + m;
      } else {
        const int y = m / p.G, x = tl * p.G + (m % p.G);
        pix = ((long long)n * p.H + y) * p.W + x;
      }
      const bool live = tile < p.total_tiles && pix < p.total_pix;
      if (!live) pix = 0;                                    // keep addresses valid; nothing is stored
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
      epilogue_tile<true>(p, sbias, taddr, (int)(pix * CH), live, (warp - 2) >> 2, lane, s1_done, s2_done,
                          rh_ready, it_tile & 1);
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive_remote(free_bar);
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  tc::cluster_sync();                  // nobody leaves while the pair's MMAs may touch its memory
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc_2cta(tmem_base, TMEM_COLS);
  }
}

}  // namespace gru
}  // namespace scflow

extern "C" size_t scflow_convgru_packed_floats(int ch, int cx) {
  if (ch != scflow::gru::CH || cx != scflow::gru::CX) return 0;
  return scflow::gru::WPACK_FLOATS;
}

extern "C" int scflow_convgru_pack_weights(const float* wz, const float* wr, const float* wq,
                                           const float* bz, const float* br, const float* bq,
                                           float* wpack, int ch, int cx, int pass, void* stream) {
  using namespace scflow;
  if (!wz || !wr || !wq || !bz || !br || !bq || !wpack) return SCFLOW_E_NULL;
  if (ch != gru::CH || cx != gru::CX) return SCFLOW_E_BAD_SHAPE;
  if (pass != 0 && pass != 1) return SCFLOW_E_UNSUPPORTED;
  gru::pack_weights_kernel<<<148 * 4, 256, 0, (cudaStream_t)stream>>>(wz, wr, wq, bz, br, bq, wpack);
  return check_launch();
}

extern "C" int scflow_convgru_pass_fwd(const float* h_in, const float* h_in_tf32, const float* x,
                                       const float* wpack, float* h_out, float* h_out_tf32,
                                       float* rh_scratch, int n, int hgt, int wid, int ch, int cx,
                                       int pass, void* stream) {
  using namespace scflow;
  using namespace scflow::gru;
  if (!h_in || !x || !wpack || !h_out || !rh_scratch) return SCFLOW_E_NULL;
  if (ch != CH || cx != CX) return SCFLOW_E_BAD_SHAPE;
  if (pass != 0 && pass != 1) return SCFLOW_E_UNSUPPORTED;
  const int axis = pass == 0 ? wid : hgt;           // convolution axis length
  if (n <= 0 || hgt <= 0 || wid <= 0 || axis > BM || (BM % axis) != 0) return SCFLOW_E_BAD_SHAPE;
  int G = BM / axis;
  const int lines = pass == 0 ? hgt : wid;
  int ni = 1;
  if (hgt * wid < BM) {                             // small maps: a tile is BM/(H*W) whole images
    if (BM % (hgt * wid) != 0) return SCFLOW_E_BAD_SHAPE;
    ni = BM / (hgt * wid);
    G = lines;
  } else if ((lines % G) != 0) {
    return SCFLOW_E_BAD_SHAPE;
  }
  if (!aligned16(h_in) || !aligned16(x) || !aligned16(wpack) || !aligned16(h_out) || !aligned16(rh_scratch) ||
      !aligned16(h_in_tf32) || !aligned16(h_out_tf32))
    return SCFLOW_E_MISALIGNED;
  const float* h_mma = h_in_tf32 ? h_in_tf32 : h_in;    // what the tensor core reads

  // activation maps: dims (c, x, y, n); box = 32 channels x the tile's pixel rectangle
  auto act_map = [&](CUtensorMap* mp, const float* base, int c) {
    const uint64_t dims[4] = {(uint64_t)c, (uint64_t)wid, (uint64_t)hgt, (uint64_t)n};
    const uint64_t strides[3] = {(uint64_t)c * 4, (uint64_t)wid * c * 4, (uint64_t)hgt * wid * c * 4};
    const uint32_t box[4] = {BK, (uint32_t)(pass == 0 ? wid : G), (uint32_t)(pass == 0 ? G : hgt), (uint32_t)ni};
    return tc::encode_f32_map(mp, base, 4, dims, strides, box, tc::kTmaSw128);
  };
  CUtensorMap map_h, map_x, map_rh, map_w;
  if (!act_map(&map_h, h_mma, CH) || !act_map(&map_x, x, CX) || !act_map(&map_rh, rh_scratch, CH))
    return SCFLOW_E_DRIVER;
  // the 2-CTA (cta_group::2) variant halves the weight bytes per SM but measures 10 % slower at
  // cfg2 (cross-CTA signalling on a latency-bound ring): kept selectable, not the default
  static const bool one_cta = getenv("SCFLOW_GRU_2CTA") == nullptr;
  {
    const uint64_t dims[2] = {BK, (uint64_t)TAPS * KB_ALL * ROWS};
    const uint64_t strides[1] = {BK * 4};
    const uint32_t box[2] = {BK, (uint32_t)(one_cta ? CH : CH / 2)};
    if (!tc::encode_f32_map(&map_w, wpack, 2, dims, strides, box, tc::kTmaSw128)) return SCFLOW_E_DRIVER;
  }
  Params p;
  p.h_in = h_in; p.bias = wpack + (size_t)TAPS * KB_ALL * ROWS * BK; p.h_out = h_out; p.h_out_tf32 = h_out_tf32; p.rh = rh_scratch;
  p.n = n; p.H = hgt; p.W = wid; p.pass = pass; p.G = G;
  p.ni = ni;
  p.total_pix = (long long)n * hgt * wid;
  if (p.total_pix * CH >= (1ll << 31)) return SCFLOW_E_BAD_SHAPE;      // epilogue uses 32-bit offsets
  p.tiles_per_img = lines / G;
  p.total_tiles = ((n + ni - 1) / ni) * p.tiles_per_img;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(convgru_pass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    attr_set = true;
  }
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  if (one_cta) {
    const int grid = p.total_tiles < sms ? p.total_tiles : sms;
    convgru_pass_kernel<<<grid, THREADS, SMEM_BYTES, (cudaStream_t)stream>>>(map_h, map_x, map_rh, map_w, p);
    return check_launch();
  }
  static int max_clusters = 0;
  if (max_clusters == 0) {
    cudaFuncSetAttribute(convgru_pass_2cta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM2_BYTES);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * sms, 1, 1); cfg.blockDim = dim3(THREADS, 1, 1); cfg.dynamicSmemBytes = SMEM2_BYTES;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr; cfg.numAttrs = 1;
    int nc = 0;
    if (cudaOccupancyMaxActiveClusters(&nc, convgru_pass_2cta_kernel, &cfg) != cudaSuccess || nc <= 0) {
      cudaGetLastError();
      nc = sms / 2;
    }
    max_clusters = nc;
  }
  const int pairs = (p.total_tiles + 1) / 2;
  const int clusters = pairs < max_clusters ? pairs : max_clusters;
  convgru_pass_2cta_kernel<<<2 * clusters, THREADS, SMEM2_BYTES, (cudaStream_t)stream>>>(map_h, map_x, map_rh,
                                                                                      map_w, p);
  return check_launch();
}

namespace scflow {
namespace gru {
__global__ void pack_nhwc_kernel(const float* __restrict__ a, int ca, const float* __restrict__ b, int cb,
                                 const float* __restrict__ f, int cf, float* __restrict__ dst, int cdst,
                                 int dst_off, long long npix, int P, int round_tf32) {
  const int ctot = ca + cb + cf;
  const long long total = npix * ctot;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long pix = i / ctot;
    const int c = (int)(i - pix * ctot);
    float v;
    if (c < ca) {
      v = a[pix * ca + c];
    } else if (c < ca + cb) {
      v = b[pix * cb + (c - ca)];
    } else {
      const long long n = pix / P;
      v = f[(n * cf + (c - ca - cb)) * P + (pix - n * P)];
    }
    dst[pix * cdst + dst_off + c] = round_tf32 ? rn_tf32(v) : v;
  }
}

// one dense channels-last source whose channel counts / offsets are multiples of 4: a strided float4 copy
__global__ void __launch_bounds__(256)
pack4_nhwc_kernel(const float4* __restrict__ a, int ca4, float4* __restrict__ dst, int cdst4, int off4,
                  unsigned total, int round_tf32) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned pix = i / (unsigned)ca4;
    const unsigned c = i - pix * (unsigned)ca4;
    float4 v = __ldg(a + i);
    if (round_tf32) { v.x = rn_tf32(v.x); v.y = rn_tf32(v.y); v.z = rn_tf32(v.z); v.w = rn_tf32(v.w); }
    dst[(size_t)pix * cdst4 + off4 + c] = v;
  }
}
}  // namespace gru
}  // namespace scflow

extern "C" int scflow_pack_nhwc(const float* a, int ca, const float* b, int cb, const float* f, int cf,
                                float* dst, int cdst, int dst_off, long long npix, int P,
                                int round_tf32, void* stream) {
  using namespace scflow;
  if (!dst || (ca > 0 && !a) || (cb > 0 && !b) || (cf > 0 && !f)) return SCFLOW_E_NULL;
  if (ca < 0 || cb < 0 || cf < 0 || npix <= 0 || P <= 0 || dst_off < 0 || dst_off + ca + cb + cf > cdst)
    return SCFLOW_E_BAD_SHAPE;
  const long long total = npix * (ca + cb + cf);
  if (total == 0) return SCFLOW_OK;
  if (cb == 0 && cf == 0 && ((ca | cdst | dst_off) & 3) == 0 && total / 4 < (1ll << 32) && aligned16(a) &&
      aligned16(dst)) {
    long long b4 = (total / 4 + 255) / 256;
    if (b4 > 148 * 16) b4 = 148 * 16;
    gru::pack4_nhwc_kernel<<<(unsigned)b4, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4*>(a), ca / 4, reinterpret_cast<float4*>(dst), cdst / 4, dst_off / 4,
        (unsigned)(total / 4), round_tf32);
    return check_launch();
  }
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  gru::pack_nhwc_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a, ca, b, cb, f, cf, dst, cdst,
                                                                          dst_off, npix, P, round_tf32);
  return check_launch();
}

// ================================================================================================
// Both passes in ONE persistent launch.  The (1,5) pass and the (5,1) pass are 512 tiles each at
// cfg2: as two launches on 148 SMs that is 2 x 4 waves at 86 % fill; as one list of 1024 work
// items it is 7 waves at 99 %.  A pass-1 item (columns of image group g) needs every pass-0 item
// of g: pass-0 epilogue warps publish h with threadfence + proxy fence + a release add on
// flags[g]; the pass-1 TMA producer and epilogue warps acquire it before touching h.  Items are
// taken in order (pass 0 first) by co-resident CTAs (grid <= SM count, 1 CTA / SM), so a wait
// only ever points at work that is running or done: no deadlock.  The same kernel runs a single
// pass when asked to (npass = 1, no flags).
namespace scflow {
namespace gru {

struct GruMaps {
  CUtensorMap h[2], x[2], rh[2], w[2];   // per slot (slot = position in the pass list)
};

struct FParams {
  const float* h;        // exact fp32 state, read by the gate (in place with h_out for npass = 2)
  float* h_out;
  float* h_out_tf32;     // nearest-TF32 copy = the tensor-core operand of the next pass / iteration
  float* rh;
  const float* bias[2];  // per slot: [384] z | r | q
  int* flags;            // npass = 2: one counter per image group, zeroed before the launch
  int n, H, W, ni;
  int npass;             // 1 or 2
  int pass[2];           // conv axis of each slot: 0 = along x, 1 = along y
  int G[2], tiles_per_img[2], tiles[2];
  int need;              // flag value that releases a slot-1 item: tiles_per_img[0] * EPI_WARPS
  int dbg_max_boxes;     // timing experiments only (SCFLOW_GRU_DBG_BOXES): cap on weight boxes per k-block
  long long* dbg_ts;     // timing experiments only (SCFLOW_GRU_DBG_TS): clock64 stamps of CTA 0, 8 per item
  int dbg_mma;           // timing experiments only (SCFLOW_GRU_DBG_MMA): 1 = no stage-1 q MMAs, 2 = no MMAs at all
  long long total_pix;
};

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add(int* p, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Both epilogues of one item for one warp (see epilogue_tile): hidden state through L2 (ld.cg:
// another SM may have written it during this launch), optional publish of the result.
__device__ __forceinline__ void epilogue_item(const FParams& p, const float* sbias, uint32_t taddr, int off,
                                              bool live, int half, int lane, uint64_t* s1_done,
                                              uint64_t* s2_done, uint64_t* rh_ready, uint32_t parity,
                                              const int* wait_flag, int* post_flag, long long* ts) {
  constexpr int HC = CH / 2;
  const int cbase = half * HC;
  if (wait_flag != nullptr) {
    if (lane == 0) while (ld_acquire(wait_flag) < p.need) __nanosleep(64);
    __syncwarp();
  }
  float h[HC];
  {
    const float4* hp = reinterpret_cast<const float4*>(p.h + (live ? off : 0) + cbase);
#pragma unroll
    for (int k = 0; k < HC / 4; ++k) {
      const float4 t = live ? __ldcg(hp + k) : make_float4(0.f, 0.f, 0.f, 0.f);
      h[4 * k + 0] = t.x; h[4 * k + 1] = t.y; h[4 * k + 2] = t.z; h[4 * k + 3] = t.w;
    }
  }
  // ---- epilogue 1: r = sigmoid(r_pre + br); r.h (rounded to nearest TF32) -> scratch
  tc::mbar_wait(s1_done, parity);
  tc::tc_fence_after();
  if (ts) ts[2] = clock64();
  float* rhp = p.rh + off + cbase;
#pragma unroll
  for (int c = 0; c < HC / 32; ++c) {
    float v[32];
    tc::tmem_ld_32x32(taddr + CH + cbase + c * 32, v);
    tc::tmem_ld_wait();
    if (live) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float4 o;
        o.x = rn_tf32(sigmoidf_(v[4 * k + 0] + sbias[CH + cbase + c * 32 + 4 * k + 0]) * h[c * 32 + 4 * k + 0]);
        o.y = rn_tf32(sigmoidf_(v[4 * k + 1] + sbias[CH + cbase + c * 32 + 4 * k + 1]) * h[c * 32 + 4 * k + 1]);
        o.z = rn_tf32(sigmoidf_(v[4 * k + 2] + sbias[CH + cbase + c * 32 + 4 * k + 2]) * h[c * 32 + 4 * k + 2]);
        o.w = rn_tf32(sigmoidf_(v[4 * k + 3] + sbias[CH + cbase + c * 32 + 4 * k + 3]) * h[c * 32 + 4 * k + 3]);
        *reinterpret_cast<float4*>(rhp + c * 32 + 4 * k) = o;
      }
    }
  }
  __threadfence();                       // r.h visible device-wide (L2) ...
  tc::fence_proxy_async_all();           // ... and ordered before the async-proxy (TMA) reads
  __syncwarp();
  if (lane == 0) tc::mbar_arrive(rh_ready);
  if (ts) ts[3] = clock64();

  // ---- epilogue 2: z, q, gate
  tc::mbar_wait(s2_done, parity);
  tc::tc_fence_after();
  if (ts) ts[6] = clock64();
  float* op = p.h_out + off + cbase;
  float* opr = p.h_out_tf32 ? p.h_out_tf32 + off + cbase : nullptr;
#pragma unroll
  for (int c = 0; c < HC / 32; ++c) {
    float zv[32], qv[32];
    tc::tmem_ld_32x32(taddr + cbase + c * 32, zv);
    tc::tmem_ld_32x32(taddr + 2 * CH + cbase + c * 32, qv);
    tc::tmem_ld_wait();
    if (live) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int ch = c * 32 + 4 * k + e;
          const float z = sigmoidf_(zv[4 * k + e] + sbias[cbase + ch]);
          const float q = tanhf_(qv[4 * k + e] + sbias[2 * CH + cbase + ch]);
          o[e] = (1.0f - z) * h[ch] + z * q;
        }
        *reinterpret_cast<float4*>(op + c * 32 + 4 * k) = make_float4(o[0], o[1], o[2], o[3]);
        if (opr)
          *reinterpret_cast<float4*>(opr + c * 32 + 4 * k) =
              make_float4(rn_tf32(o[0]), rn_tf32(o[1]), rn_tf32(o[2]), rn_tf32(o[3]));
      }
    }
  }
  if (ts) ts[7] = clock64();
  if (post_flag != nullptr) {
    __threadfence();                     // h, h_tf32 of this warp's pixels visible device-wide ...
    tc::fence_proxy_async_all();         // ... also to the other SMs' TMA reads
    __syncwarp();
    if (lane == 0) red_release_add(post_flag, 1);
  }
}

// Ring geometry per k-block width: 32 channels (128-byte rows, SWIZZLE_128B, 3 x 64 KB stages) or
// 16 channels (64-byte rows, SWIZZLE_64B, 6 x 32 KB stages).  The ring is LATENCY-bound, not
// bandwidth-bound (dropping every weight load changes nothing: 203.8 -> 201.6 us): a stage's
// refill round trip is ~1.5 us whatever it carries, so twice the stages in the same 192 KB keep
// the tensor core fed for longer.  Both read the same packed weights ([..][384][32] rows): the
// narrow ring loads each row block as two 16-float column halves.
template <int BKT>
struct Ring {
  static constexpr int kStages = BKT == 32 ? 3 : 6;
  static constexpr int kABytes = BM * BKT * 4;
  static constexpr int kBoxBytes = CH * BKT * 4;                 // one 128-row weight box
  static constexpr int kStageBytes = kABytes + 3 * kBoxBytes;
  static constexpr int kKbH = CH / BKT, kKbAll = (CH + CX) / BKT;
  static constexpr int kS1 = TAPS * kKbAll, kS2 = TAPS * kKbH;
  static constexpr uint32_t kLayout = BKT == 32 ? 2u : 4u;        // UMMA: SWIZZLE_128B / SWIZZLE_64B
  static constexpr uint32_t kSbo = 8 * BKT * 4;                   // 8-row group pitch
  static constexpr size_t kSmem = 1024 + (size_t)kStages * kStageBytes + 2 * ROWS * 4 + 256;
};

struct Item {
  int slot, tile, n, tl, grp;   // grp = image group = flag index
};
__device__ __forceinline__ Item decode_item(const FParams& p, int w) {
  Item it;
  it.slot = w >= p.tiles[0] ? 1 : 0;
  it.tile = w - (it.slot ? p.tiles[0] : 0);
  it.grp = it.tile / p.tiles_per_img[it.slot];
  it.tl = it.tile - it.grp * p.tiles_per_img[it.slot];
  it.n = it.grp * p.ni;
  return it;
}

template <int BKT>
__global__ void __launch_bounds__(THREADS, 1)
convgru_fused_kernel(const __grid_constant__ GruMaps maps, const __grid_constant__ FParams p) {
  using R = Ring<BKT>;
  constexpr int STAGES = R::kStages, STAGE_BYTES = R::kStageBytes, A_BYTES = R::kABytes;
  constexpr int S1_ITERS = R::kS1, S2_ITERS = R::kS2, KB_H = R::kKbH, KB_ALL = R::kKbAll;
  constexpr int BOX_BYTES = R::kBoxBytes;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* sbias = reinterpret_cast<float*>(smem + STAGES * STAGE_BYTES);          // [2][ROWS]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES + 2 * ROWS * 4);
  uint64_t* full = bars;                    // [STAGES]
  uint64_t* empty = bars + STAGES;          // [STAGES]
  uint64_t* s1_done = bars + 2 * STAGES;    // stage-1 accumulators complete
  uint64_t* s2_done = s1_done + 1;          // stage-2 accumulation complete
  uint64_t* rh_ready = s1_done + 2;         // r.h tile written (8 epilogue warps)
  uint64_t* tmem_free = s1_done + 3;        // epilogue finished reading TMEM (8 warps)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s1_done + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_items = p.tiles[0] + (p.npass == 2 ? p.tiles[1] : 0);

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < p.npass; ++s) {
      tc::tma_prefetch_desc(&maps.h[s]); tc::tma_prefetch_desc(&maps.x[s]);
      tc::tma_prefetch_desc(&maps.rh[s]); tc::tma_prefetch_desc(&maps.w[s]);
    }
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
      tc::mbar_init(s1_done, 1); tc::mbar_init(s2_done, 1);
      tc::mbar_init(rh_ready, EPI_WARPS); tc::mbar_init(tmem_free, EPI_WARPS);
      tc::fence_barrier_init();
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  for (int i = threadIdx.x; i < p.npass * ROWS; i += THREADS) sbias[i] = p.bias[i / ROWS][i % ROWS];
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0; uint32_t it_tile = 0;
      for (int w = blockIdx.x; w < total_items; w += gridDim.x, ++it_tile) {
        const Item item = decode_item(p, w);
        const int s = item.slot, pass = p.pass[s], G = p.G[s];
        if (s == 1) {                       // h of this image group comes from slot-0 items
          while (ld_acquire(p.flags + item.grp) < p.need) __nanosleep(64);
          tc::fence_proxy_async_all();
        }
        for (int it = 0; it < S1_ITERS + S2_ITERS; ++it) {
          const bool s2 = it >= S1_ITERS;
          if (it == S1_ITERS) {
            tc::mbar_wait(rh_ready, it_tile & 1);   // r.h of this tile is in L2
            if (p.dbg_ts && blockIdx.x == 0 && it_tile < 4) p.dbg_ts[it_tile * 8 + 4] = clock64();
          }
          const int j = s2 ? it - S1_ITERS : it;
          const int t = s2 ? j / KB_H : j / KB_ALL;
          const int kb = s2 ? j % KB_H : j % KB_ALL;
          tc::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* a_s = smem + stage * STAGE_BYTES;
          uint8_t* b_s = a_s + A_BYTES;
          int nb = s2 ? 1 : (kb < KB_H ? 2 : 3);                      // 128-row weight boxes
          if (nb > p.dbg_max_boxes) nb = p.dbg_max_boxes;
          tc::mbar_expect_tx(&full[stage], (p.dbg_mma == 4 ? 0 : A_BYTES) + nb * BOX_BYTES);
          const CUtensorMap* am = s2 ? &maps.rh[s] : (kb < KB_H ? &maps.h[s] : &maps.x[s]);
          const int c0 = (s2 || kb < KB_H) ? kb * BKT : (kb - KB_H) * BKT;
          const int cx = pass == 0 ? (t - 2) : item.tl * G;
          const int cy = pass == 0 ? item.tl * G : (t - 2);
          if (p.dbg_mma < 3) tc::tma_load_4d(am, &full[stage], a_s, c0, cx, cy, item.n);
          else if (p.dbg_mma == 3) tc::tma_load_2d(&maps.w[s], &full[stage], a_s, 0, (it * 7 + item.tile) % 1000 * CH);
          // packed weights are [tap][32-channel block][384 rows][32]: a narrower k-block is a column slice
          const int wrow = (t * ((CH + CX) / BK) + (kb * BKT) / BK) * ROWS, wcol = (kb * BKT) % BK;
          if (s2) {
            if (nb > 0) tc::tma_load_2d(&maps.w[s], &full[stage], b_s, wcol, wrow + 2 * CH);
          } else {
            for (int b = 0; b < nb; ++b)
              tc::tma_load_2d(&maps.w[s], &full[stage], b_s + b * BOX_BYTES, wcol, wrow + b * CH);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t id256 = tc::make_idesc_tf32(BM, 256, 0, 0);   // K-major A and B
      constexpr uint32_t id128 = tc::make_idesc_tf32(BM, 128, 0, 0);
      int stage = 0; uint32_t phase = 0; uint32_t it_tile = 0;
      for (int w = blockIdx.x; w < total_items; w += gridDim.x, ++it_tile) {
        tc::mbar_wait(tmem_free, (it_tile & 1) ^ 1);
        tc::tc_fence_after();
        long long* ts = (p.dbg_ts && blockIdx.x == 0 && it_tile < 4) ? p.dbg_ts + it_tile * 8 : nullptr;
        if (ts) ts[0] = clock64();
        for (int it = 0; it < S1_ITERS + S2_ITERS; ++it) {
          const bool s2 = it >= S1_ITERS;
          const int kb = s2 ? 0 : it % KB_ALL;
          tc::mbar_wait(&full[stage], phase);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
          for (int ks = 0; ks < BKT / 8; ++ks) {
            // K-major swizzled rows: 8-row groups kSbo apart; a k-step (8 tf32) is +32 B
            const uint64_t da = tc::make_smem_desc(a_addr + ks * 32, 16, R::kSbo, R::kLayout);
            if (s2) {
              const uint64_t dq = tc::make_smem_desc(b_addr + ks * 32, 16, R::kSbo, R::kLayout);
              tc::mma_tf32_ss(tmem_base + 2 * CH, da, dq, id128, 1u);
            } else {
              const uint64_t dzr = tc::make_smem_desc(b_addr + ks * 32, 16, R::kSbo, R::kLayout);
              if (p.dbg_mma < 2) tc::mma_tf32_ss(tmem_base, da, dzr, id256, (it | ks) != 0 ? 1u : 0u);
              if (kb >= KB_H && p.dbg_mma == 0) {
                const uint64_t dq = tc::make_smem_desc(b_addr + 2 * BOX_BYTES + ks * 32, 16, R::kSbo, R::kLayout);
                tc::mma_tf32_ss(tmem_base + 2 * CH, da, dq, id128, (it != KB_H || ks != 0) ? 1u : 0u);
              }
            }
          }
          tc::tc_commit(&empty[stage]);
          if (it == S1_ITERS - 1) { tc::tc_commit(s1_done); if (ts) ts[1] = clock64(); }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        tc::tc_commit(s2_done);
        if (ts) ts[5] = clock64();
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue (8 warps)
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;                     // TMEM lane = pixel of the tile
    uint32_t it_tile = 0;
    for (int w = blockIdx.x; w < total_items; w += gridDim.x, ++it_tile) {
      const Item item = decode_item(p, w);
      const int s = item.slot, G = p.G[s];
      long long pix;
      if (p.pass[s] == 0 || p.ni > 1) {
        pix = (long long)item.n * p.H * p.W + (long long)item.tl * BM + m;       // G full rows
      } else {
        const int y = m / G, x = item.tl * G + (m % G);                            // G full columns
        pix = ((long long)item.n * p.H + y) * p.W + x;
      }
      const bool live = pix < p.total_pix;       // whole-image tiles may run past the last image
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
      const bool fused = p.npass == 2;
      epilogue_item(p, sbias + s * ROWS, taddr, (int)(pix * CH), live, (warp - 2) >> 2, lane, s1_done, s2_done,
                    rh_ready, it_tile & 1, fused && s == 1 ? p.flags + item.grp : nullptr,
                    fused && s == 0 ? p.flags + item.grp : nullptr,
                    (p.dbg_ts && blockIdx.x == 0 && it_tile < 4 && warp == 2 && lane == 0) ? p.dbg_ts + it_tile * 8 : nullptr);
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(tmem_free);
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace gru
}  // namespace scflow

namespace scflow {
namespace gru {

// shape of one pass: lines per tile, images per tile, tiles
struct PassShape {
  int G, ni, tiles_per_img, tiles;
};
static int pass_shape(int n, int hgt, int wid, int pass, PassShape* s) {
  const int axis = pass == 0 ? wid : hgt;           // convolution axis length
  if (n <= 0 || hgt <= 0 || wid <= 0 || axis > BM || (BM % axis) != 0) return SCFLOW_E_BAD_SHAPE;
  s->G = BM / axis;
  const int lines = pass == 0 ? hgt : wid;
  s->ni = 1;
  if (hgt * wid < BM) {                             // small maps: a tile is BM/(H*W) whole images
    if (BM % (hgt * wid) != 0) return SCFLOW_E_BAD_SHAPE;
    s->ni = BM / (hgt * wid);
    s->G = lines;
  } else if ((lines % s->G) != 0) {
    return SCFLOW_E_BAD_SHAPE;
  }
  s->tiles_per_img = lines / s->G;
  s->tiles = ((n + s->ni - 1) / s->ni) * s->tiles_per_img;
  return SCFLOW_OK;
}

static int launch_fused(const float* h, const float* h_tf32, const float* x, const float* const* wpack,
                        float* h_out, float* h_out_tf32, float* rh, int* flags, int n, int hgt, int wid,
                        int npass, const int* passes, cudaStream_t stream) {
  GruMaps maps;
  FParams p;
  const float* h_mma = h_tf32 ? h_tf32 : h;
  // k-block width of the operand ring (see Ring<>): 16 = six 32 KB stages (default), 32 = three 64 KB
  const char* ebk = getenv("SCFLOW_GRU_BK");
  const int bk = (ebk && atoi(ebk) == 32) ? 32 : 16;
  const int swz = bk == 32 ? tc::kTmaSw128 : tc::kTmaSw64;
  for (int s = 0; s < npass; ++s) {
    PassShape ps;
    const int rc = pass_shape(n, hgt, wid, passes[s], &ps);
    if (rc != SCFLOW_OK) return rc;
    if (s == 1 && ps.ni != p.ni) return SCFLOW_E_BAD_SHAPE;
    p.ni = ps.ni; p.pass[s] = passes[s]; p.G[s] = ps.G; p.tiles_per_img[s] = ps.tiles_per_img; p.tiles[s] = ps.tiles;
    auto act_map = [&](CUtensorMap* mp, const float* base, int c) {
      const uint64_t dims[4] = {(uint64_t)c, (uint64_t)wid, (uint64_t)hgt, (uint64_t)n};
      const uint64_t strides[3] = {(uint64_t)c * 4, (uint64_t)wid * c * 4, (uint64_t)hgt * wid * c * 4};
      const uint32_t box[4] = {(uint32_t)bk, (uint32_t)(passes[s] == 0 ? wid : ps.G),
                               (uint32_t)(passes[s] == 0 ? ps.G : hgt), (uint32_t)ps.ni};
      return tc::encode_f32_map(mp, base, 4, dims, strides, box, swz);
    };
    // slot 1 reads the h that slot 0 wrote
    const float* hs = (s == 1 && h_out_tf32) ? h_out_tf32 : (s == 1 ? h_out : h_mma);
    if (!act_map(&maps.h[s], hs, CH) || !act_map(&maps.x[s], x, CX) || !act_map(&maps.rh[s], rh, CH))
      return SCFLOW_E_DRIVER;
    const uint64_t dims[2] = {BK, (uint64_t)TAPS * KB_ALL * ROWS};
    const uint64_t strides[1] = {BK * 4};
    const uint32_t box[2] = {(uint32_t)bk, (uint32_t)CH};
    if (!tc::encode_f32_map(&maps.w[s], wpack[s], 2, dims, strides, box, swz)) return SCFLOW_E_DRIVER;
    p.bias[s] = wpack[s] + (size_t)TAPS * KB_ALL * ROWS * BK;
  }
  if (npass == 1) {
    maps.h[1] = maps.h[0]; maps.x[1] = maps.x[0]; maps.rh[1] = maps.rh[0]; maps.w[1] = maps.w[0];
    p.bias[1] = p.bias[0]; p.pass[1] = p.pass[0]; p.G[1] = p.G[0]; p.tiles_per_img[1] = p.tiles_per_img[0];
    p.tiles[1] = 0;
  }
  p.h = h; p.h_out = h_out; p.h_out_tf32 = h_out_tf32; p.rh = rh; p.flags = flags;
  p.n = n; p.H = hgt; p.W = wid; p.npass = npass;
  p.need = p.tiles_per_img[0] * EPI_WARPS;
  {
    const char* e = getenv("SCFLOW_GRU_DBG_BOXES");
    p.dbg_max_boxes = e ? atoi(e) : 3;
    const char* m = getenv("SCFLOW_GRU_DBG_MMA");
    p.dbg_mma = m ? atoi(m) : 0;
    // stamps live in the 256 bytes after the flags (the workspace query reserves them)
    p.dbg_ts = (getenv("SCFLOW_GRU_DBG_TS") && npass == 2)
                   ? reinterpret_cast<long long*>(reinterpret_cast<char*>(flags) + (((size_t)n * 4 + 255) / 256) * 256)
                   : nullptr;
  }
  p.total_pix = (long long)n * hgt * wid;
  if (p.total_pix * CH >= (1ll << 31)) return SCFLOW_E_BAD_SHAPE;      // epilogue uses 32-bit offsets
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(convgru_fused_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Ring<32>::kSmem);
    cudaFuncSetAttribute(convgru_fused_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Ring<16>::kSmem);
    attr_set = true;
  }
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  const int items = p.tiles[0] + (npass == 2 ? p.tiles[1] : 0);
  if (npass == 2) {
    const int groups = (n + p.ni - 1) / p.ni;
    if (cudaMemsetAsync(flags, 0, (size_t)groups * sizeof(int), stream) != cudaSuccess) return check_launch();
  }
  const int grid = items < sms ? items : sms;
  if (bk == 32)
    convgru_fused_kernel<32><<<grid, THREADS, Ring<32>::kSmem, stream>>>(maps, p);
  else
    convgru_fused_kernel<16><<<grid, THREADS, Ring<16>::kSmem, stream>>>(maps, p);
  return check_launch();
}

}  // namespace gru
}  // namespace scflow

extern "C" size_t scflow_convgru_workspace_bytes(int n, int hgt, int wid) {
  (void)hgt; (void)wid;
  // one dependency counter per image (group) + 256 bytes of debug stamps
  return n > 0 ? (((size_t)n * sizeof(int) + 255) / 256) * 256 + 256 : 0;
}

extern "C" int scflow_convgru_fwd(float* h, float* h_tf32, const float* x, const float* wpack0,
                                  const float* wpack1, float* rh_scratch, void* workspace,
                                  size_t workspace_bytes, int n, int hgt, int wid, int ch, int cx,
                                  void* stream) {
  using namespace scflow;
  using namespace scflow::gru;
  if (!h || !x || !wpack0 || !wpack1 || !rh_scratch || !workspace) return SCFLOW_E_NULL;
  if (ch != CH || cx != CX) return SCFLOW_E_BAD_SHAPE;
  if (workspace_bytes < scflow_convgru_workspace_bytes(n, hgt, wid)) return SCFLOW_E_WORKSPACE;
  if (!aligned16(h) || !aligned16(x) || !aligned16(wpack0) || !aligned16(wpack1) || !aligned16(rh_scratch) ||
      !aligned16(h_tf32) || (reinterpret_cast<uintptr_t>(workspace) & 3u))
    return SCFLOW_E_MISALIGNED;
  const float* wp[2] = {wpack0, wpack1};
  const int passes[2] = {0, 1};
  return launch_fused(h, h_tf32, x, wp, h, h_tf32, rh_scratch, static_cast<int*>(workspace), n, hgt, wid, 2,
                      passes, (cudaStream_t)stream);
}
