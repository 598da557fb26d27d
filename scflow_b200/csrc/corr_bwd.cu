// Second half of the backward of CorrelationPyramid.forward (models/decoder/raft_decoder.py:47-56) on the tensor
// cores: scflow_corr_pyramid_bwd.  With G = d loss / d (f1^T f2) (n, P, P), folded from the level gradients by
// scflow_corr_fold_bwd (scale and 2x2-mean adjoints included),
//     d f1[c][q] = sum_t f2[c][t] G[q][t]          d f2[c][t] = sum_q f1[c][q] G[q][t]
// are two batched GEMMs per crop (2 x 2 P^2 C flops: as much as the forward).  Round 1 ran them as torch.bmm
// (cuBLAS); here they are ONE persistent warp-specialised tcgen05 launch, kind::tf32 (the gradient tolerance is
// 2e-3; TF32 is 5e-4), fp32 operands straight from the caller's tensors through TMA:
//   product 1 (d f1): A = f2 [c][t] and B = G [q][t], both K-major (the contraction index t is contiguous)
//   product 2 (d f2): A = f1 [c][q] K-major; B = G [q][t] with the contraction index q as the slow axis: MN-major,
//                     the SWIZZLE_128B_BASE32B layout of the forward kernel (csrc/corr_pyramid.cu)
// Tile = 128 channels x 256 positions, K = P in 32-float k-blocks through a 4-stage 48 KB ring; the accumulator is
// double-buffered in TMEM; the epilogue (8 warps) writes planar (n, c, P) rows, 128 contiguous bytes per lane.
#include "common.cuh"
#include "tc_common.cuh"
#include <string.h>

namespace scflow {
namespace cbw {

constexpr int BM = 128, BN = 256, BK = 32, UK = 8;
constexpr int STAGES = 4;
constexpr int A_BYTES = BM * BK * 4, B_BYTES = BN * BK * 4, STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int NEPI = 8;
constexpr int THREADS = 64 + 32 * NEPI;
constexpr int TMEM_COLS = 512;
constexpr size_t SMEM_BYTES = 1024 + (size_t)STAGES * STAGE_BYTES + 256;

struct Maps {
  CUtensorMap f1k, f2k;        // f[c][pos] as K-major A: box {32 pos, 128 c}
  CUtensorMap gk;              // G[q][t] as K-major B: box {32 t, 256 q}
  CUtensorMap gmn;             // G as MN-major B: dims {32, q, t / 32, n}: box {32 t, 32 q, 8 groups}
};
struct Params {
  float* d1; float* d2;        // (n, c, P) each, or null
  int n, c, P;
  int m_tiles, n_tiles;        // c / 128, P / 256
};

__global__ void __launch_bounds__(THREADS, 1)
corr_bwd_kernel(const __grid_constant__ Maps maps, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* acc_full = bars + 2 * STAGES;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int KB = p.P / BK;
  const int per_prod = p.m_tiles * p.n_tiles;
  const int prods = (p.d1 ? 1 : 0) + (p.d2 ? 1 : 0);
  const int total = p.n * prods * per_prod;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&maps.f1k); tc::tma_prefetch_desc(&maps.f2k);
    tc::tma_prefetch_desc(&maps.gk); tc::tma_prefetch_desc(&maps.gmn);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
      for (int i = 0; i < 2; ++i) { tc::mbar_init(&acc_full[i], 1); tc::mbar_init(&acc_empty[i], NEPI); }
      tc::fence_barrier_init();
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // tile t -> (crop, product, channel tile, position tile); product 0 = d f1 unless only d f2 is wanted
  auto decode = [&](int t, int& n, int& prod, int& mt, int& nt) {
    nt = t % p.n_tiles; t /= p.n_tiles;
    mt = t % p.m_tiles; t /= p.m_tiles;
    prod = t % prods; n = t / prods;
    if (!p.d1) prod = 1;
  };

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        int n, prod, mt, nt;
        decode(t, n, prod, mt, nt);
        for (int kb = 0; kb < KB; ++kb) {
          tc::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* st = smem + stage * STAGE_BYTES;
          tc::mbar_expect_tx(&full[stage], STAGE_BYTES);
          if (prod == 0) {      // d f1: A = f2[c][t-chunk], B = G[q-tile][t-chunk] (both K-major)
            tc::tma_load_3d(&maps.f2k, &full[stage], st, kb * BK, mt * BM, n);
            tc::tma_load_3d(&maps.gk, &full[stage], st + A_BYTES, kb * BK, nt * BN, n);
          } else {              // d f2: A = f1[c][q-chunk] (K-major), B = G[q-chunk][t-tile] (MN-major)
            tc::tma_load_3d(&maps.f1k, &full[stage], st, kb * BK, mt * BM, n);
            tc::tma_load_4d(&maps.gmn, &full[stage], st + A_BYTES, 0, kb * BK, nt * (BN / 32), n);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (tc::elect_one()) {
      constexpr uint32_t idesc_kk = tc::make_idesc_tf32(BM, BN, 0, 0);
      constexpr uint32_t idesc_kmn = tc::make_idesc_tf32(BM, BN, 0, 1);
      int stage = 0; uint32_t phase = 0, it = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
        int n, prod, mt, nt;
        decode(t, n, prod, mt, nt);
        const uint32_t buf = it & 1;
        tc::mbar_wait(&acc_empty[buf], ((it >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        const uint32_t d = tmem_base + buf * BN;
        for (int kb = 0; kb < KB; ++kb) {
          tc::mbar_wait(&full[stage], phase);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(smem + stage * STAGE_BYTES), b_addr = a_addr + A_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / UK; ++kk) {
            // K-major fp32, SWIZZLE_128B: a row is 32 floats = 128 bytes, a k-step (8 floats) is +32 bytes,
            // 8-row groups 1024 bytes apart.  MN-major: corr_pyramid.cu (atoms of 4 k-rows, LBO 4096, SBO 512).
            const uint64_t da = tc::make_smem_desc(a_addr + kk * 32, 16, 1024, tc::kLayoutSw128);
            if (prod == 0) {
              const uint64_t db = tc::make_smem_desc(b_addr + kk * 32, 16, 1024, tc::kLayoutSw128);
              tc::mma_tf32_ss(d, da, db, idesc_kk, (kb | kk) != 0 ? 1u : 0u);
            } else {
              const uint64_t db = tc::make_smem_desc(b_addr + kk * 1024, 4096, 512, tc::kLayoutSw128Base32);
              tc::mma_tf32_ss(d, da, db, idesc_kmn, (kb | kk) != 0 ? 1u : 0u);
            }
          }
          tc::tc_commit(&empty[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        tc::tc_commit(&acc_full[buf]);
      }
    }
    __syncwarp();
  } else {
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    uint32_t it = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
      int n, prod, mt, nt;
      decode(t, n, prod, mt, nt);
      const uint32_t buf = it & 1;
      tc::mbar_wait(&acc_full[buf], (it >> 1) & 1);
      tc::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + buf * BN;
      float* out = (prod == 0 ? p.d1 : p.d2) + ((size_t)n * p.c + mt * BM + quarter * 32 + lane) * p.P + nt * BN;
      for (int j = half; j < BN / 32; j += 2) {
        float v[32];
        tc::tmem_ld_32x32(taddr + j * 32, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 4; ++k)
          asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(out + j * 32 + 8 * k), "f"(v[8 * k]),
                       "f"(v[8 * k + 1]), "f"(v[8 * k + 2]), "f"(v[8 * k + 3]), "f"(v[8 * k + 4]), "f"(v[8 * k + 5]),
                       "f"(v[8 * k + 6]), "f"(v[8 * k + 7]) : "memory");
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&acc_empty[buf]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// SCFLOW_PREC_FP32 (exact-mode anchor, like the forward's corr_fp32) and shapes outside the tcgen05 tiling:
// fp32 FMA, 64 x 64 tile, 16-deep k-blocks, 4 x 4 outputs per thread.  D[c][j] = sum_k A[c][k] B(k, j) with
// B(k, j) = G[j][k] (TRANSB: d f1) or G[k][j] (d f2).
template <bool TRANSB>
__global__ void __launch_bounds__(256) corr_bwd_fp32_kernel(const float* __restrict__ A, const float* __restrict__ G,
                                                            float* __restrict__ D, int c, int P) {
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int n = blockIdx.z, c0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
  A += (size_t)n * c * P; G += (size_t)n * P * P; D += (size_t)n * c * P;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < P; k0 += 16) {
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      const int r = i >> 4, k = i & 15;                     // A[c0 + r][k0 + k]: k contiguous
      As[k][r] = (c0 + r < c && k0 + k < P) ? A[(size_t)(c0 + r) * P + k0 + k] : 0.f;
    }
    if (TRANSB) {
      for (int i = threadIdx.x; i < 64 * 16; i += 256) {
        const int j = i >> 4, k = i & 15;                   // G[j0 + j][k0 + k]: k contiguous
        Bs[k][j] = (j0 + j < P && k0 + k < P) ? G[(size_t)(j0 + j) * P + k0 + k] : 0.f;
      }
    } else {
      for (int i = threadIdx.x; i < 64 * 16; i += 256) {
        const int k = i >> 6, j = i & 63;                   // G[k0 + k][j0 + j]: j contiguous
        Bs[k][j] = (j0 + j < P && k0 + k < P) ? G[(size_t)(k0 + k) * P + j0 + j] : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[k][ty * 4 + i]; b[i] = Bs[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (c0 + ty * 4 + i < c && j0 + tx * 4 + j < P) D[(size_t)(c0 + ty * 4 + i) * P + j0 + tx * 4 + j] = acc[i][j];
}

}  // namespace cbw
}  // namespace scflow

using namespace scflow;

extern "C" int scflow_corr_pyramid_bwd(const float* grad_corr, const float* feat1, const float* feat2, float* grad_f1,
                                       float* grad_f2, int n, int c, int h, int w, int precision, void* stream) {
  using namespace cbw;
  if (!grad_corr || !feat1 || !feat2 || (!grad_f1 && !grad_f2)) return SCFLOW_E_NULL;
  if (precision != SCFLOW_PREC_FP32 && precision != SCFLOW_PREC_TF32) return SCFLOW_E_UNSUPPORTED;
  const long long P = (long long)h * w;
  if (n <= 0 || c <= 0 || P <= 0 || P > (1 << 20) || n > 65535) return SCFLOW_E_BAD_SHAPE;
  auto a32 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 31u) == 0; };
  const bool aligned = a32(grad_corr) && a32(feat1) && a32(feat2) && a32(grad_f1) && a32(grad_f2);
  if (precision == SCFLOW_PREC_FP32 || (c % BM) != 0 || (P % BN) != 0 || !aligned) {
    const dim3 grid((unsigned)((P + 63) / 64), (unsigned)((c + 63) / 64), (unsigned)n);
    if (grad_f1) corr_bwd_fp32_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(feat2, grad_corr, grad_f1, c, (int)P);
    if (grad_f2) corr_bwd_fp32_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(feat1, grad_corr, grad_f2, c, (int)P);
    return check_launch();
  }
  Maps maps;
  {
    const uint64_t dims[3] = {(uint64_t)P, (uint64_t)c, (uint64_t)n};
    const uint64_t strides[2] = {(uint64_t)P * 4, (uint64_t)c * P * 4};
    const uint32_t box[3] = {BK, BM, 1};
    if (!tc::encode_f32_map(&maps.f1k, feat1, 3, dims, strides, box, tc::kTmaSw128) ||
        !tc::encode_f32_map(&maps.f2k, feat2, 3, dims, strides, box, tc::kTmaSw128))
      return SCFLOW_E_DRIVER;
  }
  {
    const uint64_t dims[3] = {(uint64_t)P, (uint64_t)P, (uint64_t)n};
    const uint64_t strides[2] = {(uint64_t)P * 4, (uint64_t)P * P * 4};
    const uint32_t box[3] = {BK, BN, 1};
    if (!tc::encode_f32_map(&maps.gk, grad_corr, 3, dims, strides, box, tc::kTmaSw128)) return SCFLOW_E_DRIVER;
  }
  {
    const uint64_t dims[4] = {32, (uint64_t)P, (uint64_t)(P / 32), (uint64_t)n};
    const uint64_t strides[3] = {(uint64_t)P * 4, 128, (uint64_t)P * P * 4};
    const uint32_t box[4] = {32, BK, BN / 32, 1};
    if (!tc::encode_f32_map(&maps.gmn, grad_corr, 4, dims, strides, box, tc::kTmaSw128Atom32)) return SCFLOW_E_DRIVER;
  }
  Params p{grad_f1, grad_f2, n, c, (int)P, c / BM, (int)(P / BN)};
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  const int total = n * ((grad_f1 ? 1 : 0) + (grad_f2 ? 1 : 0)) * p.m_tiles * p.n_tiles;
  cudaFuncSetAttribute(corr_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
  corr_bwd_kernel<<<total < sms ? total : sms, THREADS, SMEM_BYTES, (cudaStream_t)stream>>>(maps, p);
  return check_launch();
}
