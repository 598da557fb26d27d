// CorrelationPyramid forward: all-pairs correlation feat1^T . feat2 / sqrt(C) and its 2x2-mean
// pyramid.  Replaces models/decoder/raft_decoder.py:47-56 (bmm + full-volume div + 3 avg_pool2d).
//
// TF32 path (default): persistent warp-specialised tcgen05 GEMM.  Both operands are MN-major in
// the reference's NCHW layout (the contraction dim C is the slow axis), so TMA loads fp32 boxes
// straight from the feature maps into the UMMA MN-major SWIZZLE_128B_BASE32B layout -- no transpose or
// cast pre-pass -- and tcgen05.mma kind::tf32 accumulates 128 queries x 256 targets in TMEM.  A
// tile's 256 target columns are an 8-row x 32-column PATCH of the target plane, so the scale and
// all four pyramid levels are produced thread-locally in the epilogue (each thread owns one
// query row) and written once: algorithmic HBM traffic only, no re-read of the volume.
//
// FP32 path: CUDA-core SGEMM + one pooling pass.  Exact-arithmetic parity anchor for the TF32
// kernel (reference default is fp32 SGEMM, torch.backends.cuda.matmul.allow_tf32 = False).
#include "common.cuh"
#include "tc_common.cuh"
#include <cstdlib>

namespace scflow {

// =============================================================================== FP32 (SIMT)
constexpr int kSB = 64, kSK = 16;   // 64x64 output tile, 16-deep k slab, 256 threads x (4x4)

__global__ void __launch_bounds__(256)
corr_fp32_kernel(const float* __restrict__ f1, const float* __restrict__ f2, float* __restrict__ out,
                 int C, int P, float scale) {
  __shared__ __align__(16) float As[kSK][kSB];
  __shared__ __align__(16) float Bs[kSK][kSB];
  const int n = blockIdx.z;
  const int q0 = blockIdx.y * kSB, t0 = blockIdx.x * kSB;
  const float* A = f1 + (long long)n * C * P;
  const float* B = f2 + (long long)n * C * P;
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  const int lr = tid >> 4, lc = (tid & 15) * 4;   // load: row (k) and column quad
  float acc[4][4] = {};
  for (int k0 = 0; k0 < C; k0 += kSK) {
    *reinterpret_cast<float4*>(&As[lr][lc]) =
        *reinterpret_cast<const float4*>(A + (long long)(k0 + lr) * P + q0 + lc);
    *reinterpret_cast<float4*>(&Bs[lr][lc]) =
        *reinterpret_cast<const float4*>(B + (long long)(k0 + lr) * P + t0 + lc);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kSK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float* o = out + ((long long)n * P + q0 + ty * 4 + i) * P + t0 + tx * 4;
    *reinterpret_cast<float4*>(o) =
        make_float4(acc[i][0] / scale, acc[i][1] / scale, acc[i][2] / scale, acc[i][3] / scale);
  }
}

// 2x2 mean in ATen avg_pool2d order: ((a+b)+c)+d, then /4
__device__ __forceinline__ float pool4(float a, float b, float c, float d) {
  return (((a + b) + c) + d) * 0.25f;
}

// One thread per 8x8 block of one level-0 plane: writes its 4x4 / 2x2 / 1 cells of levels 1..3.
__global__ void __launch_bounds__(256)
pool_pyramid_kernel(const float* __restrict__ l0, float* __restrict__ l1, float* __restrict__ l2,
                    float* __restrict__ l3, long long planes, int h, int w, int num_levels) {
  const int bw = w / 8, bh = h / 8;
  const long long total = planes * bh * bw;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int bx = (int)(i % bw), by = (int)((i / bw) % bh);
    const long long pl = i / ((long long)bw * bh);
    const float* src = l0 + pl * h * w + (long long)(by * 8) * w + bx * 8;
    float v[8][8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const float4 a = *reinterpret_cast<const float4*>(src + r * w);
      const float4 b = *reinterpret_cast<const float4*>(src + r * w + 4);
      v[r][0] = a.x; v[r][1] = a.y; v[r][2] = a.z; v[r][3] = a.w;
      v[r][4] = b.x; v[r][5] = b.y; v[r][6] = b.z; v[r][7] = b.w;
    }
    float p1[4][4], p2[2][2];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c)
        p1[r][c] = pool4(v[2 * r][2 * c], v[2 * r][2 * c + 1], v[2 * r + 1][2 * c], v[2 * r + 1][2 * c + 1]);
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int c = 0; c < 2; ++c)
        p2[r][c] = pool4(p1[2 * r][2 * c], p1[2 * r][2 * c + 1], p1[2 * r + 1][2 * c], p1[2 * r + 1][2 * c + 1]);
    if (num_levels > 1) {
      float* d = l1 + pl * (h / 2) * (w / 2) + (long long)(by * 4) * (w / 2) + bx * 4;
#pragma unroll
      for (int r = 0; r < 4; ++r)
        *reinterpret_cast<float4*>(d + r * (w / 2)) = make_float4(p1[r][0], p1[r][1], p1[r][2], p1[r][3]);
    }
    if (num_levels > 2) {
      float* d = l2 + pl * (h / 4) * (w / 4) + (long long)(by * 2) * (w / 4) + bx * 2;
      *reinterpret_cast<float2*>(d) = make_float2(p2[0][0], p2[0][1]);
      *reinterpret_cast<float2*>(d + (w / 4)) = make_float2(p2[1][0], p2[1][1]);
    }
    if (num_levels > 3)
      l3[pl * (h / 8) * (w / 8) + (long long)by * (w / 8) + bx] = pool4(p2[0][0], p2[0][1], p2[1][0], p2[1][1]);
  }
}

static int corr_fp32(const float* f1, const float* f2, float* const* levels, int num_levels, int n,
                     int c, int h, int w, cudaStream_t st) {
  const int P = h * w;
  if ((P % kSB) != 0 || (c % kSK) != 0 || (h % 8) != 0 || (w % 8) != 0) return SCFLOW_E_BAD_SHAPE;
  dim3 grid(P / kSB, P / kSB, n);
  corr_fp32_kernel<<<grid, 256, 0, st>>>(f1, f2, levels[0], c, P, sqrtf((float)c));
  int rc = check_launch();
  if (rc != SCFLOW_OK || num_levels == 1) return rc;
  const long long planes = (long long)n * P;
  long long total = planes * (h / 8) * (w / 8);
  long long blocks = (total + 255) / 256;
  if (blocks > 148LL * 32) blocks = 148LL * 32;
  pool_pyramid_kernel<<<(unsigned)blocks, 256, 0, st>>>(levels[0], levels[1], num_levels > 2 ? levels[2] : nullptr,
                                                       num_levels > 3 ? levels[3] : nullptr, planes, h, w, num_levels);
  return check_launch();
}

// ============================================================================ TF32 (tcgen05)
namespace corr_tc {

constexpr int BM = 128;            // queries per tile (TMEM lanes)
constexpr int BN = 256;            // targets per tile: 8 rows x 32 columns of the target plane
constexpr int BK = 32;             // fp32 channels per pipeline stage (one 128-byte swizzle row)
constexpr int UK = 8;              // K of one tcgen05.mma kind::tf32
constexpr int STAGES = 3;
constexpr int A_BYTES = BM * BK * 4;             // 16 KB: [4 groups][32 k][32 q]
constexpr int B_BYTES = BN * BK * 4;             // 32 KB: [8 rows ][32 k][32 x]
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int EPI_GROUPS = 2;                    // one 4-warp epilogue group per accumulator stage
constexpr int EPI_BUF_BYTES = 32 * 128;          // 32 queries x one 32-float target row (swizzled)
constexpr int EPI_WARP_BYTES = 2 * EPI_BUF_BYTES;
constexpr int EPI_BYTES = EPI_GROUPS * 4 * EPI_WARP_BYTES;
constexpr int THREADS = 64 + EPI_GROUPS * 128;   // warp0 TMA, warp1 MMA + TMEM alloc, 8 epilogue warps
constexpr int TMEM_COLS = 512;                   // two 256-column accumulator stages
constexpr size_t SMEM_BYTES = 1024 + (size_t)STAGES * STAGE_BYTES + EPI_BYTES + 256;

struct Params {
  float* lvl[4];
  int n, c, h, w, num_levels;
  int tiles_m, tiles_y, tiles_x;
  float scale, inv_scale;
  int scale_is_pow2;
};

__device__ __forceinline__ void decode_tile(const Params& p, int tile, int tiles_per_m, int& pj, int& mt, int& n) {
  // target tile fastest: neighbours share the QUERY tile (query-tile-fastest measured the same)
  pj = tile % tiles_per_m;
  mt = (tile / tiles_per_m) % p.tiles_m;
  n = tile / (tiles_per_m * p.tiles_m);
}

template <bool POW2>
__global__ void __launch_bounds__(THREADS, 1)
corr_tf32_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 const __grid_constant__ CUtensorMap mapC, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* epi = smem + STAGES * STAGE_BYTES;                 // 1024-byte aligned (swizzle phase)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES + EPI_BYTES);
  uint64_t* full = bars;                 // [STAGES] TMA -> MMA
  uint64_t* empty = bars + STAGES;       // [STAGES] MMA -> TMA
  uint64_t* tfull = bars + 2 * STAGES;   // [2] MMA -> epilogue
  uint64_t* tempty = tfull + 2;          // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int P = p.h * p.w;
  const int KB = p.c / BK;
  const int tiles_per_m = p.tiles_y * p.tiles_x;
  const int total_tiles = p.n * p.tiles_m * tiles_per_m;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&mapA);
    tc::tma_prefetch_desc(&mapB);
    tc::tma_prefetch_desc(&mapC);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
      for (int i = 0; i < 2; ++i) { tc::mbar_init(&tfull[i], 1); tc::mbar_init(&tempty[i], 4); }
      tc::fence_barrier_init();
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        int pj, mt, n;
        decode_tile(p, tile, tiles_per_m, pj, mt, n);
        const int ty = pj / p.tiles_x, tx = pj % p.tiles_x;
        for (int kb = 0; kb < KB; ++kb) {
          tc::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* a_s = smem + stage * STAGE_BYTES;
          tc::mbar_expect_tx(&full[stage], STAGE_BYTES);
          tc::tma_load_4d(&mapA, &full[stage], a_s, 0, kb * BK, mt * (BM / 32), n);
          tc::tma_load_4d(&mapB, &full[stage], a_s + A_BYTES, tx * 32, kb * BK, ty * 8, n);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = tc::make_idesc_tf32(BM, BN, 1, 1);   // both operands MN-major
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        tc::mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < KB; ++kb) {
          tc::mbar_wait(&full[stage], phase);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / UK; ++kk) {
            // MN-major 32-bit operands: SWIZZLE_128B_BASE32B is the only layout tcgen05 accepts.
            // Atom = 4 k-rows x 128 B (SBO = 512 B between atoms along K); 32-element MN groups
            // are 4096 B apart (LBO); one MMA (K = 8) spans two atoms = 1024 B.  Pinned on
            // hardware with tools/umma_probe.cu.
            const uint64_t da = tc::make_smem_desc(a_addr + kk * 1024, 4096, 512, tc::kLayoutSw128Base32);
            const uint64_t db = tc::make_smem_desc(b_addr + kk * 1024, 4096, 512, tc::kLayoutSw128Base32);
            tc::mma_tf32_ss(d_tmem, da, db, idesc, (kb | kk) != 0 ? 1u : 0u);
          }
          tc::tc_commit(&empty[stage]);            // frees the smem slot when these MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        tc::tc_commit(&tfull[acc]);                // accumulator complete -> epilogue
        acc ^= 1; if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (2 groups x 4 warps)
    // Group g drains accumulator stage g (every other tile of this CTA), so one tile's epilogue may
    // take two MMA periods.  A warp owns 32 queries (its TMEM lane quarter); a lane owns one query.
    const int quarter = warp & 3;                  // TMEM lane quarter this warp may access
    const int group = (warp - 2) >> 2;
    uint8_t* stg = epi + ((group * 4 + quarter) * EPI_WARP_BYTES);
    const uint32_t stg_lane = tc::smem_u32(stg) + lane * 128;
    const int sw = lane & 7;                       // SWIZZLE_128B: 16-byte chunk index ^= row & 7
    uint32_t acc_phase = 0;
    const int w1 = p.w / 2, w2 = p.w / 4, w3 = p.w / 8;
    const long long P1 = (long long)(p.h / 2) * w1, P2 = (long long)(p.h / 4) * w2,
                    P3 = (long long)(p.h / 8) * w3;
    for (int tile = blockIdx.x + group * gridDim.x; tile < total_tiles; tile += EPI_GROUPS * gridDim.x) {
      int pj, mt, n;
      decode_tile(p, tile, tiles_per_m, pj, mt, n);
      const int ty = pj / p.tiles_x, tx = pj % p.tiles_x;
      const long long qbase = (long long)n * P + mt * BM + quarter * 32;   // first query of this warp
      tc::mbar_wait(&tfull[group], acc_phase);
      tc::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + group * BN;
      float vb[2][32], prev0[32], prev1[16], prev2[8];
      tc::tmem_ld_32x32(taddr, vb[0]);
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        float* v = vb[r & 1];
        tc::tmem_ld_wait_on(v);
        if (r < 7) tc::tmem_ld_32x32(taddr + (r + 1) * 32, vb[(r + 1) & 1]);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = POW2 ? v[j] * p.inv_scale : __fdiv_rn(v[j], p.scale);
        // level 0: the warp's 32 queries x 32 targets of row r go out as one TMA box; the store
        // issued two rows ago must have drained this staging buffer
        if (lane == 0) tc::bulk_wait_read<1>();
        __syncwarp();
        {
          const uint32_t base = stg_lane + (r & 1) * EPI_BUF_BYTES;
#pragma unroll
          for (int k = 0; k < 8; ++k)
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(base + ((k ^ sw) << 4)),
                         "f"(v[4 * k]), "f"(v[4 * k + 1]), "f"(v[4 * k + 2]), "f"(v[4 * k + 3])
                         : "memory");
        }
        tc::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tc::tma_store_3d(&mapC, stg + (r & 1) * EPI_BUF_BYTES, tx * 32, ty * 8 + r, (int)qbase);
          tc::bulk_commit();
        }
        if ((r & 1) == 0) {
#pragma unroll
          for (int j = 0; j < 32; ++j) prev0[j] = v[j];
        } else if (p.num_levels > 1) {
          float l1[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) l1[j] = pool4(prev0[2 * j], prev0[2 * j + 1], v[2 * j], v[2 * j + 1]);
          float* d1 = p.lvl[1] + (qbase + lane) * P1 + (long long)(ty * 4 + (r >> 1)) * w1 + tx * 16;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            *reinterpret_cast<float4*>(d1 + 4 * k) = make_float4(l1[4 * k], l1[4 * k + 1], l1[4 * k + 2], l1[4 * k + 3]);
          if ((r & 3) == 1) {
#pragma unroll
            for (int j = 0; j < 16; ++j) prev1[j] = l1[j];
          } else if (p.num_levels > 2) {
            float l2[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) l2[j] = pool4(prev1[2 * j], prev1[2 * j + 1], l1[2 * j], l1[2 * j + 1]);
            float* d2 = p.lvl[2] + (qbase + lane) * P2 + (long long)(ty * 2 + (r >> 2)) * w2 + tx * 8;
            *reinterpret_cast<float4*>(d2) = make_float4(l2[0], l2[1], l2[2], l2[3]);
            *reinterpret_cast<float4*>(d2 + 4) = make_float4(l2[4], l2[5], l2[6], l2[7]);
            if (r == 3) {
#pragma unroll
              for (int j = 0; j < 8; ++j) prev2[j] = l2[j];
            } else if (p.num_levels > 3) {
              float* d3 = p.lvl[3] + (qbase + lane) * P3 + (long long)ty * w3 + tx * 4;
              *reinterpret_cast<float4*>(d3) =
                  make_float4(pool4(prev2[0], prev2[1], l2[0], l2[1]), pool4(prev2[2], prev2[3], l2[2], l2[3]),
                              pool4(prev2[4], prev2[5], l2[4], l2[5]), pool4(prev2[6], prev2[7], l2[6], l2[7]));
            }
          }
        }
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&tempty[group]);
      acc_phase ^= 1;
    }
    if (lane == 0) tc::bulk_wait_all<0>();         // level-0 stores complete before the CTA exits
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace corr_tc

static int device_sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

static int corr_tf32(const float* f1, const float* f2, float* const* levels, int num_levels, int n,
                     int c, int h, int w, cudaStream_t st) {
  using namespace corr_tc;
  const int P = h * w;
  if ((c % BK) != 0 || (h % 8) != 0 || (w % 32) != 0 || (P % BM) != 0) return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(f1) || !aligned16(f2)) return SCFLOW_E_MISALIGNED;
  for (int l = 0; l < num_levels; ++l)
    if (!aligned16(levels[l])) return SCFLOW_E_MISALIGNED;
  CUtensorMap mapA, mapB, mapC;
  if ((long long)n * P > 0x7fffffffLL) return SCFLOW_E_BAD_SHAPE;
  {
    const uint64_t dims[4] = {32, (uint64_t)c, (uint64_t)(P / 32), (uint64_t)n};
    const uint64_t strides[3] = {(uint64_t)P * 4, 128, (uint64_t)c * P * 4};
    const uint32_t box[4] = {32, BK, BM / 32, 1};
    if (!tc::encode_f32_map(&mapA, f1, 4, dims, strides, box, tc::kTmaSw128Atom32)) return SCFLOW_E_DRIVER;
  }
  {
    const uint64_t dims[4] = {(uint64_t)w, (uint64_t)c, (uint64_t)h, (uint64_t)n};
    const uint64_t strides[3] = {(uint64_t)P * 4, (uint64_t)w * 4, (uint64_t)c * P * 4};
    const uint32_t box[4] = {32, BK, 8, 1};
    if (!tc::encode_f32_map(&mapB, f2, 4, dims, strides, box, tc::kTmaSw128Atom32)) return SCFLOW_E_DRIVER;
  }
  {
    // level 0 as (w, h, n*P) for the epilogue's stores: one box = 32 targets of one row x 32 queries
    const uint64_t dims[3] = {(uint64_t)w, (uint64_t)h, (uint64_t)n * P};
    const uint64_t strides[2] = {(uint64_t)w * 4, (uint64_t)P * 4};
    const uint32_t box[3] = {32, 1, 32};
    if (!tc::encode_f32_map(&mapC, levels[0], 3, dims, strides, box, tc::kTmaSw128)) return SCFLOW_E_DRIVER;
  }
  Params p;
  for (int l = 0; l < 4; ++l) p.lvl[l] = l < num_levels ? levels[l] : nullptr;
  p.n = n; p.c = c; p.h = h; p.w = w; p.num_levels = num_levels;
  p.tiles_m = P / BM; p.tiles_y = h / 8; p.tiles_x = w / 32;
  p.scale = sqrtf((float)c);
  p.inv_scale = 1.0f / p.scale;
  int e; const float fr = frexpf(p.scale, &e);
  p.scale_is_pow2 = (fr == 0.5f) ? 1 : 0;
  // function attributes are per device: set on every launch (cheap), no process-wide guard
  cudaFuncSetAttribute(corr_tf32_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
  cudaFuncSetAttribute(corr_tf32_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
  const long long total = (long long)n * p.tiles_m * p.tiles_y * p.tiles_x;
  const int grid = (int)(total < device_sm_count() ? total : device_sm_count());
  if (p.scale_is_pow2)
    corr_tf32_kernel<true><<<grid, THREADS, SMEM_BYTES, st>>>(mapA, mapB, mapC, p);
  else
    corr_tf32_kernel<false><<<grid, THREADS, SMEM_BYTES, st>>>(mapA, mapB, mapC, p);
  return check_launch();
}

}  // namespace scflow

extern "C" size_t scflow_corr_pyramid_workspace_bytes(int, int, int, int, int) { return 0; }

extern "C" int scflow_corr_pyramid_fwd(const float* feat1, const float* feat2, float* const* levels,
                                       int num_levels, int n, int c, int h, int w, int precision,
                                       void* workspace, size_t workspace_bytes, void* stream) {
  using namespace scflow;
  (void)workspace; (void)workspace_bytes;
  if (!feat1 || !feat2 || !levels) return SCFLOW_E_NULL;
  if (num_levels < 1 || num_levels > 4) return SCFLOW_E_UNSUPPORTED;
  for (int l = 0; l < num_levels; ++l)
    if (!levels[l]) return SCFLOW_E_NULL;
  if (n <= 0 || c <= 0 || h <= 0 || w <= 0) return SCFLOW_E_BAD_SHAPE;
  if (precision == SCFLOW_PREC_FP32) return corr_fp32(feat1, feat2, levels, num_levels, n, c, h, w, (cudaStream_t)stream);
  if (precision == SCFLOW_PREC_TF32) {
    // shapes outside the tcgen05 tiling (w % 32, h % 8, c % 32, h*w % 128) take the fp32 FMA kernel: a stricter
    // result than the one asked for, never a refusal (the same rule as scflow_corr_pyramid_bwd)
    const int rc = corr_tf32(feat1, feat2, levels, num_levels, n, c, h, w, (cudaStream_t)stream);
    return rc == SCFLOW_E_BAD_SHAPE ? corr_fp32(feat1, feat2, levels, num_levels, n, c, h, w, (cudaStream_t)stream) : rc;
  }
  return SCFLOW_E_UNSUPPORTED;
}
