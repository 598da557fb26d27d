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
#include "convgru_common.cuh"
#include <stdlib.h>

namespace scflow {
namespace gru {

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

