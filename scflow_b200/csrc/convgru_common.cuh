// Constants and scalar helpers shared by the ConvGRU kernels (convgru.cu, convgru_fused.cu).
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

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

}  // namespace gru
}  // namespace scflow
