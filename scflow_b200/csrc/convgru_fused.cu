// ConvGRU (SeqConv), BOTH passes in one persistent launch: scflow_convgru_fwd.
// Replaces models/decoder/raft_decoder.py:245-252 (6 cuDNN convs + 6 activation kernels + 4 cat +
// 8 elementwise per iteration).  The per-pass mathematics, operand layouts and the two-stage
// accumulation are those of convgru.cu (read its header first); this file changes the schedule:
//
//   * One list of work items.  The (1,5) pass and the (5,1) pass are 512 tiles each at cfg2: as
//     two launches on 148 SMs that is 2 x 4 waves at 86 % fill, as one list of 1024 items it is 7
//     waves at 99 %.  A pass-1 item (columns of image group g) needs every pass-0 item of g:
//     pass-0 epilogue warps publish h with threadfence + proxy fence + a release add on flags[g];
//     the pass-1 TMA producer and epilogue warps acquire it before touching h.  Items are taken in
//     order (pass 0 first) by co-resident CTAs (grid <= SM count, 1 CTA / SM), so a wait only ever
//     points at work that is running or done: no deadlock.
//   * Sixteen epilogue warps (four per TMEM lane quarter, 32 channels each) and EARLY TMEM
//     RELEASE: the final epilogue pulls z and q out of TMEM into registers, hands the accumulators
//     back at once, and does the sigmoid / tanh / gate arithmetic and the stores while the tensor
//     core already runs the next item's stage 1.  Clock stamps of the previous schedule (8 warps,
//     TMEM held to the end): stage 1 29 us, r.h epilogue 4.2 us, stage 2 7.5 us, final epilogue
//     8 us, hand-over 2 us per item -- the final epilogue and the hand-over are now hidden.
//   * r.h leaves through shared memory: the epilogue warps write their 32 x 32 slices, swizzled, into
//     the activation slots of the first two stage-2 ring stages (idle between stage 1 and the r.h
//     loads) and ONE TMA tensor store per channel block carries them to L2 -- per-lane 16-byte global
//     stores at a 512-byte stride back-pressured the LSU for 4 us per item (epilogue 1: 5.0 -> 1.8 us).
//   * The MMA issue loop is written per (tap, k-block) with compile-time branches: the issuing
//     thread is on the critical path (two extra predicated branches per k-step cost 10 %).
//
// What was measured and NOT adopted (tools/kbench.py gru, tools/gru_stamps.py):
//   * a six-stage ring of 32 KB stages (16-channel k-blocks, SWIZZLE_64B): 26 % slower -- the cost
//     per k-block hand-shake dominates, not the refill latency;
//   * dropping every weight load changes nothing (203.8 -> 201.6 us): not L2 -> SM bandwidth;
//   * busy-polling (mbarrier.test_wait) instead of try_wait in the ring: no change;
//   * L2 prefetch (cp.async.bulk.prefetch.tensor) of the next item's activation tiles: 164.6 ->
//     168.3 us, slightly slower;
//   * software-pipelining items (z and r both read out in epilogue 1, q alternating between two
//     TMEM blocks, the next item's first k-blocks issued ahead of this item's stage 2): correct
//     but 6-30 % SLOWER -- with the TMA stream running, the r.h epilogue (strided 16-byte stores
//     + threadfence) takes 8-12 us instead of 5 and the gate epilogue ~20 us: the epilogue warps
//     become the bottleneck.  What remains exposed per item (41 us = 29 stage 1 + 5 r.h epilogue
//     + 6.5 stage 2) needs a coalesced epilogue store path first (no shared memory left for it
//     next to the 192 KB ring).
#include "convgru_common.cuh"
#include <stdlib.h>

namespace scflow {
namespace gru {

constexpr int FEPI_WARPS = 16;                   // four warps per TMEM lane quarter
constexpr int FTHREADS = 64 + 32 * FEPI_WARPS;   // TMA warp, MMA warp, epilogue warps
constexpr int FC = CH / 4;                       // channels per epilogue thread
constexpr size_t SMEMF_BYTES = 1024 + (size_t)STAGES * STAGE_BYTES + 2 * ROWS * 4 + 256;

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
  long long* stamps;     // optional (SCFLOW_GRU_DBG_TS): clock64 stamps of CTA 0, 8 per item
  int n, H, W, ni;
  int npass;             // 1 or 2
  int pass[2];           // conv axis of each slot: 0 = along x, 1 = along y
  int G[2], tiles_per_img[2], tiles[2];
  int need;              // flag value that releases a slot-1 item: tiles_per_img[0] * FEPI_WARPS
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

// Both epilogues of one item for one warp: 32 pixels (its TMEM lane quarter) x 32 channels.
// Hidden state through L2 (ld.cg: another SM may have written it during this launch).
__device__ __forceinline__ void epilogue_item(const FParams& p, const float* sbias, uint32_t taddr, int off,
                                              bool live, int cbase, int lane, uint64_t* s1_done,
                                              uint64_t* s2_done, uint64_t* rh_ready, uint64_t* tmem_free,
                                              uint32_t parity, const int* wait_flag, int* post_flag,
                                              long long* ts, uint8_t* rh_slot, const CUtensorMap* map_rh, int cx0,
                                              int cy0, int n_img, int m) {
  if (wait_flag != nullptr) {
    if (lane == 0) while (ld_acquire(wait_flag) < p.need) __nanosleep(64);
    __syncwarp();
  }
  const float4* hp = reinterpret_cast<const float4*>(p.h + (live ? off : 0) + cbase);
  {
    // ---- epilogue 1: r = sigmoid(r_pre + br); r.h (rounded to nearest TF32) -> scratch.  The
    // pixel's hidden-state slice is fetched before the wait, behind the stage-1 main loop.
    float h[FC];
#pragma unroll
    for (int k = 0; k < FC / 4; ++k) {
      const float4 t = live ? __ldcg(hp + k) : make_float4(0.f, 0.f, 0.f, 0.f);
      h[4 * k + 0] = t.x; h[4 * k + 1] = t.y; h[4 * k + 2] = t.z; h[4 * k + 3] = t.w;
    }
    tc::mbar_wait(s1_done, parity);
    tc::tc_fence_after();
    if (ts) ts[2] = clock64();
    float v[FC];
    tc::tmem_ld_32x32(taddr + CH + cbase, v);
    tc::tmem_ld_wait();
    // r.h of this warp's 32 pixels x 32 channels goes into the idle activation slot of a stage-2 ring
    // stage, as one 128-byte row per pixel in the 128-byte swizzle (chunk j of row m at j ^ (m & 7):
    // eight conflict-free STS.128), and leaves as ONE TMA tensor store per channel block: per-lane
    // 16-byte global stores at a 512-byte stride back-pressured the LSU for 4 us per item.
    {
      uint8_t* row = rh_slot + m * 128;
#pragma unroll
      for (int k = 0; k < FC / 4; ++k) {
        float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live) {
          o.x = rn_tf32(sigmoidf_(v[4 * k + 0] + sbias[CH + cbase + 4 * k + 0]) * h[4 * k + 0]);
          o.y = rn_tf32(sigmoidf_(v[4 * k + 1] + sbias[CH + cbase + 4 * k + 1]) * h[4 * k + 1]);
          o.z = rn_tf32(sigmoidf_(v[4 * k + 2] + sbias[CH + cbase + 4 * k + 2]) * h[4 * k + 2]);
          o.w = rn_tf32(sigmoidf_(v[4 * k + 3] + sbias[CH + cbase + 4 * k + 3]) * h[4 * k + 3]);
        }
        *reinterpret_cast<float4*>(row + ((k ^ (m & 7)) << 4)) = o;
      }
    }
    tc::fence_proxy_async();               // generic-proxy smem writes -> visible to the TMA store
    asm volatile("bar.sync %0, 128;" ::"r"(1 + cbase / FC) : "memory");     // the four warps of this channel block
    if (m == 0) {
      tc::tma_store_4d(map_rh, rh_slot, cbase, cx0, cy0, n_img);
      tc::bulk_commit();
      tc::bulk_wait_all<0>();              // written (L2) and the slot is free again
      tc::mbar_arrive(rh_ready);
    }
    if (ts) ts[3] = clock64();
  }

  // ---- epilogue 2: pull z and q out of TMEM, give the accumulators back, then gate
  tc::mbar_wait(s2_done, parity);
  tc::tc_fence_after();
  if (ts) ts[6] = clock64();
  float zv[FC], qv[FC];
  tc::tmem_ld_32x32(taddr + cbase, zv);
  tc::tmem_ld_32x32(taddr + 2 * CH + cbase, qv);
  tc::tmem_ld_wait();
  tc::tc_fence_before();
  __syncwarp();
  if (lane == 0) tc::mbar_arrive(tmem_free);     // the next item's MMAs may overwrite TMEM now
  if (ts) ts[7] = clock64();
  if (live) {
    float* op = p.h_out + off + cbase;
    float* opr = p.h_out_tf32 ? p.h_out_tf32 + off + cbase : nullptr;
#pragma unroll
    for (int k = 0; k < FC / 4; ++k) {
      const float4 t = __ldcg(hp + k);           // L2 hit (read a stage ago); frees 32 registers
      const float hv[4] = {t.x, t.y, t.z, t.w};
      float o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int ch = 4 * k + e;
        const float z = sigmoidf_(zv[ch] + sbias[cbase + ch]);
        const float q = tanhf_(qv[ch] + sbias[2 * CH + cbase + ch]);
        o[e] = (1.0f - z) * hv[e] + z * q;
      }
      *reinterpret_cast<float4*>(op + 4 * k) = make_float4(o[0], o[1], o[2], o[3]);
      if (opr)
        *reinterpret_cast<float4*>(opr + 4 * k) =
            make_float4(rn_tf32(o[0]), rn_tf32(o[1]), rn_tf32(o[2]), rn_tf32(o[3]));
    }
  }
  if (post_flag != nullptr) {
    __threadfence();                     // h, h_tf32 of this warp's pixels visible device-wide ...
    tc::fence_proxy_async_all();         // ... also to the other SMs' TMA reads
    __syncwarp();
    if (lane == 0) red_release_add(post_flag, 1);
  }
}

__global__ void __launch_bounds__(FTHREADS, 1)
convgru_fused_kernel(const __grid_constant__ GruMaps maps, const __grid_constant__ FParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* sbias = reinterpret_cast<float*>(smem + STAGES * STAGE_BYTES);          // [2][ROWS]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES + 2 * ROWS * 4);
  uint64_t* full = bars;                    // [STAGES]
  uint64_t* empty = bars + STAGES;          // [STAGES]
  uint64_t* s1_done = bars + 2 * STAGES;    // stage-1 accumulators complete
  uint64_t* s2_done = s1_done + 1;          // stage-2 accumulation complete
  uint64_t* rh_ready = s1_done + 2;         // r.h tile stored (one arrival per channel block)
  uint64_t* tmem_free = s1_done + 3;        // accumulators read out (16 epilogue warps)
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
      tc::mbar_init(rh_ready, CH / FC); tc::mbar_init(tmem_free, FEPI_WARPS);
      tc::fence_barrier_init();
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  for (int i = threadIdx.x; i < p.npass * ROWS; i += FTHREADS) sbias[i] = p.bias[i / ROWS][i % ROWS];
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0; uint32_t it_tile = 0;
      auto next_stage = [&]() { if (++stage == STAGES) { stage = 0; phase ^= 1; } };
      for (int w = blockIdx.x; w < total_items; w += gridDim.x, ++it_tile) {
        const Item item = decode_item(p, w);
        const int s = item.slot, pass = p.pass[s], G = p.G[s];
        if (s == 1) {                       // h of this image group comes from slot-0 items
          while (ld_acquire(p.flags + item.grp) < p.need) __nanosleep(64);
          tc::fence_proxy_async_all();
        }
        // tile origin (x, y) shifted by the tap along the convolution axis; TMA zero-fills
        // everything outside [0,W) x [0,H): that is the conv's zero padding
        const int bx = pass == 0 ? -2 : item.tl * G, by = pass == 0 ? item.tl * G : -2;
        const int dx = pass == 0 ? 1 : 0, dy = pass == 0 ? 0 : 1;
        // ---- stage 1: hx . [Wz|Wr]^T and x . Wq[:, x part]^T
        for (int t = 0; t < TAPS; ++t) {
          const int cx = bx + t * dx, cy = by + t * dy;
#pragma unroll
          for (int kb = 0; kb < KB_ALL; ++kb) {
            tc::mbar_wait(&empty[stage], phase ^ 1);
            uint8_t* a_s = smem + stage * STAGE_BYTES;
            uint8_t* b_s = a_s + A_BYTES;
            constexpr int kBox = CH * BK * 4;                          // one 128-row weight box
            const int nb = kb < KB_H ? 2 : 3;
            tc::mbar_expect_tx(&full[stage], A_BYTES + nb * kBox);
            if (kb < KB_H) tc::tma_load_4d(&maps.h[s], &full[stage], a_s, kb * BK, cx, cy, item.n);
            else tc::tma_load_4d(&maps.x[s], &full[stage], a_s, (kb - KB_H) * BK, cx, cy, item.n);
            const int wrow = (t * KB_ALL + kb) * ROWS;
            tc::tma_load_2d(&maps.w[s], &full[stage], b_s, 0, wrow);
            tc::tma_load_2d(&maps.w[s], &full[stage], b_s + kBox, 0, wrow + CH);
            if (kb >= KB_H) tc::tma_load_2d(&maps.w[s], &full[stage], b_s + 2 * kBox, 0, wrow + 2 * CH);
            next_stage();
          }
        }
        // ---- stage 2: (r.h) . Wq[:, h part]^T, TWO k-blocks per ring stage (2 x (16 KB pixels + 16 KB
        // weights) = one 64 KB stage): the ring's cost is per hand-shake, and these MMAs are small.
        // The weight boxes do not depend on r.h: those of the first STAGES stages are requested
        // BEFORE the wait, their activations after it.
        constexpr int kBox = CH * BK * 4;
        auto s2_coords = [&](int j, int& t, int& kb) { t = j / KB_H; kb = j - t * KB_H; };
        int pre_stage[STAGES];
#pragma unroll
        for (int i = 0; i < STAGES; ++i) {
          tc::mbar_wait(&empty[stage], phase ^ 1);
          tc::mbar_expect_tx(&full[stage], 2 * (A_BYTES + kBox));
          uint8_t* st = smem + stage * STAGE_BYTES;
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            int t, kb; s2_coords(2 * i + u, t, kb);
            tc::tma_load_2d(&maps.w[s], &full[stage], st + 2 * A_BYTES + u * kBox, 0, (t * KB_ALL + kb) * ROWS + 2 * CH);
          }
          pre_stage[i] = stage;
          next_stage();
        }
        tc::mbar_wait(rh_ready, it_tile & 1);                          // r.h of this tile is in L2
        if (p.stamps && blockIdx.x == 0 && it_tile < 4) p.stamps[it_tile * 8 + 4] = clock64();
#pragma unroll
        for (int i = 0; i < STAGES; ++i) {
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            int t, kb; s2_coords(2 * i + u, t, kb);
            tc::tma_load_4d(&maps.rh[s], &full[pre_stage[i]], smem + pre_stage[i] * STAGE_BYTES + u * A_BYTES, kb * BK,
                            bx + t * dx, by + t * dy, item.n);
          }
        }
        for (int i = STAGES; i < S2_ITERS / 2; ++i) {
          tc::mbar_wait(&empty[stage], phase ^ 1);
          tc::mbar_expect_tx(&full[stage], 2 * (A_BYTES + kBox));
          uint8_t* st = smem + stage * STAGE_BYTES;
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            int t, kb; s2_coords(2 * i + u, t, kb);
            tc::tma_load_4d(&maps.rh[s], &full[stage], st + u * A_BYTES, kb * BK, bx + t * dx, by + t * dy, item.n);
            tc::tma_load_2d(&maps.w[s], &full[stage], st + 2 * A_BYTES + u * kBox, 0, (t * KB_ALL + kb) * ROWS + 2 * CH);
          }
          next_stage();
        }
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t id256 = tc::make_idesc_tf32(BM, 256, 0, 0);   // K-major A and B
      constexpr uint32_t id128 = tc::make_idesc_tf32(BM, 128, 0, 0);
      int stage = 0; uint32_t phase = 0; uint32_t it_tile = 0;
      auto next_stage = [&]() { if (++stage == STAGES) { stage = 0; phase ^= 1; } };
      const uint32_t smem_a0 = tc::smem_u32(smem);
      for (int w = blockIdx.x; w < total_items; w += gridDim.x, ++it_tile) {
        tc::mbar_wait(tmem_free, (it_tile & 1) ^ 1);
        tc::tc_fence_after();
        long long* ts = (p.stamps && blockIdx.x == 0 && it_tile < 4) ? p.stamps + it_tile * 8 : nullptr;
        if (ts) ts[0] = clock64();
        for (int t = 0; t < TAPS; ++t) {
#pragma unroll
          for (int kb = 0; kb < KB_ALL; ++kb) {
            tc::mbar_wait(&full[stage], phase);
            tc::tc_fence_after();
            const uint32_t a_addr = smem_a0 + stage * STAGE_BYTES;
            const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
            for (int ks = 0; ks < BK / 8; ++ks) {
              // K-major SWIZZLE_128B: 8-row groups 1024 B apart; a k-step (8 tf32) is +32 B
              const uint64_t da = tc::make_smem_desc(a_addr + ks * 32, 16, 1024, tc::kLayoutSw128);
              const uint64_t dzr = tc::make_smem_desc(b_addr + ks * 32, 16, 1024, tc::kLayoutSw128);
              tc::mma_tf32_ss(tmem_base, da, dzr, id256, (t | kb | ks) != 0 ? 1u : 0u);
              if (kb >= KB_H) {
                const uint64_t dq = tc::make_smem_desc(b_addr + 2 * CH * BK * 4 + ks * 32, 16, 1024, tc::kLayoutSw128);
                tc::mma_tf32_ss(tmem_base + 2 * CH, da, dq, id128, (t != 0 || kb != KB_H || ks != 0) ? 1u : 0u);
              }
            }
            tc::tc_commit(&empty[stage]);
            next_stage();
          }
        }
        tc::tc_commit(s1_done);
        if (ts) ts[1] = clock64();
        for (int it = 0; it < S2_ITERS / 2; ++it) {                     // two k-blocks per stage
          tc::mbar_wait(&full[stage], phase);
          tc::tc_fence_after();
          const uint32_t a_addr = smem_a0 + stage * STAGE_BYTES;
          const uint32_t b_addr = a_addr + 2 * A_BYTES;
#pragma unroll
          for (int u = 0; u < 2; ++u) {
#pragma unroll
            for (int ks = 0; ks < BK / 8; ++ks) {
              const uint64_t da = tc::make_smem_desc(a_addr + u * A_BYTES + ks * 32, 16, 1024, tc::kLayoutSw128);
              const uint64_t dq = tc::make_smem_desc(b_addr + u * (CH * BK * 4) + ks * 32, 16, 1024, tc::kLayoutSw128);
              tc::mma_tf32_ss(tmem_base + 2 * CH, da, dq, id128, 1u);
            }
          }
          tc::tc_commit(&empty[stage]);
          next_stage();
        }
        tc::tc_commit(s2_done);
        if (ts) ts[5] = clock64();
      }
    }
  } else {
    // ---------------------------------------------------------------- epilogue (16 warps)
    const int quarter = warp & 3;                          // TMEM lanes this warp may read
    const int cbase = ((warp - 2) >> 2) * FC;              // its channel slice
    const int m = quarter * 32 + lane;                     // TMEM lane = pixel of the tile
    const bool fused = p.npass == 2;
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
      epilogue_item(p, sbias + s * ROWS, taddr, (int)(pix * CH), live, cbase, lane, s1_done, s2_done, rh_ready,
                    tmem_free, it_tile & 1, fused && s == 1 ? p.flags + item.grp : nullptr,
                    fused && s == 0 ? p.flags + item.grp : nullptr,
                    (p.stamps && blockIdx.x == 0 && it_tile < 4 && warp == 2 && lane == 0) ? p.stamps + it_tile * 8
                                                                                           : nullptr,
                    // the ring stands at stage (70 * item + 60) % 3 when stage 2 begins (60 + 10 stages per
                    // item); channel blocks 0, 1 use that stage's two activation slots, 2, 3 the next one's
                    smem + ((70u * it_tile + 60u + (uint32_t)(cbase / (2 * FC))) % STAGES) * STAGE_BYTES +
                        ((cbase / FC) & 1) * A_BYTES,
                    &maps.rh[s], p.pass[s] == 0 ? 0 : item.tl * G, p.pass[s] == 0 ? item.tl * G : 0, item.n, m);
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

static size_t flag_bytes(int n) { return (((size_t)n * sizeof(int) + 255) / 256) * 256; }

static int launch_fused(const float* h, const float* h_tf32, const float* x, const float* const* wpack,
                        float* h_out, float* h_out_tf32, float* rh, int* flags, int n, int hgt, int wid,
                        int npass, const int* passes, cudaStream_t stream) {
  GruMaps maps;
  FParams p;
  const float* h_mma = h_tf32 ? h_tf32 : h;
  for (int s = 0; s < npass; ++s) {
    PassShape ps;
    const int rc = pass_shape(n, hgt, wid, passes[s], &ps);
    if (rc != SCFLOW_OK) return rc;
    if (s == 1 && ps.ni != p.ni) return SCFLOW_E_BAD_SHAPE;
    p.ni = ps.ni; p.pass[s] = passes[s]; p.G[s] = ps.G; p.tiles_per_img[s] = ps.tiles_per_img; p.tiles[s] = ps.tiles;
    // activation maps: dims (c, x, y, n); box = 32 channels x the tile's pixel rectangle
    auto act_map = [&](CUtensorMap* mp, const float* base, int c) {
      const uint64_t dims[4] = {(uint64_t)c, (uint64_t)wid, (uint64_t)hgt, (uint64_t)n};
      const uint64_t strides[3] = {(uint64_t)c * 4, (uint64_t)wid * c * 4, (uint64_t)hgt * wid * c * 4};
      const uint32_t box[4] = {BK, (uint32_t)(passes[s] == 0 ? wid : ps.G), (uint32_t)(passes[s] == 0 ? ps.G : hgt),
                               (uint32_t)ps.ni};
      return tc::encode_f32_map(mp, base, 4, dims, strides, box, tc::kTmaSw128);
    };
    // slot 1 reads the h that slot 0 wrote
    const float* hs = (s == 1 && h_out_tf32) ? h_out_tf32 : (s == 1 ? h_out : h_mma);
    if (!act_map(&maps.h[s], hs, CH) || !act_map(&maps.x[s], x, CX) || !act_map(&maps.rh[s], rh, CH))
      return SCFLOW_E_DRIVER;
    const uint64_t dims[2] = {BK, (uint64_t)TAPS * KB_ALL * ROWS};
    const uint64_t strides[1] = {BK * 4};
    const uint32_t box[2] = {BK, (uint32_t)CH};
    if (!tc::encode_f32_map(&maps.w[s], wpack[s], 2, dims, strides, box, tc::kTmaSw128)) return SCFLOW_E_DRIVER;
    p.bias[s] = wpack[s] + (size_t)TAPS * KB_ALL * ROWS * BK;
  }
  if (npass == 1) {
    maps.h[1] = maps.h[0]; maps.x[1] = maps.x[0]; maps.rh[1] = maps.rh[0]; maps.w[1] = maps.w[0];
    p.bias[1] = p.bias[0]; p.pass[1] = p.pass[0]; p.G[1] = p.G[0]; p.tiles_per_img[1] = p.tiles_per_img[0];
    p.tiles[1] = 0;
  }
  p.h = h; p.h_out = h_out; p.h_out_tf32 = h_out_tf32; p.rh = rh; p.flags = flags;
  p.n = n; p.H = hgt; p.W = wid; p.npass = npass;
  p.need = p.tiles_per_img[0] * FEPI_WARPS;
  // clock stamps (tools/gru_stamps.py) live in the 256 bytes after the flags
  p.stamps = (getenv("SCFLOW_GRU_DBG_TS") && npass == 2)
                 ? reinterpret_cast<long long*>(reinterpret_cast<char*>(flags) + flag_bytes(n))
                 : nullptr;
  p.total_pix = (long long)n * hgt * wid;
  if (p.total_pix * CH >= (1ll << 31)) return SCFLOW_E_BAD_SHAPE;      // epilogue uses 32-bit offsets
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(convgru_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEMF_BYTES);
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
  convgru_fused_kernel<<<grid, FTHREADS, SMEMF_BYTES, stream>>>(maps, p);
  return check_launch();
}

}  // namespace gru
}  // namespace scflow

extern "C" size_t scflow_convgru_workspace_bytes(int n, int hgt, int wid) {
  (void)hgt; (void)wid;
  // one dependency counter per image (group) + 256 bytes of optional debug stamps
  return n > 0 ? scflow::gru::flag_bytes(n) + 256 : 0;
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
      !aligned16(h_tf32) || (reinterpret_cast<uintptr_t>(workspace) & 7u))
    return SCFLOW_E_MISALIGNED;
  const float* wp[2] = {wpack0, wpack1};
  const int passes[2] = {0, 1};
  return launch_fused(h, h_tf32, x, wp, h, h_tf32, rh_scratch, static_cast<int*>(workspace), n, hgt, wid, 2,
                      passes, (cudaStream_t)stream);
}
