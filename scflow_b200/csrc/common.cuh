// Shared helpers for the scflow_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "scflow_b200.h"

namespace scflow {

extern thread_local int g_last_cuda_error;

inline int check_launch() {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    g_last_cuda_error = (int)e;
    return SCFLOW_E_LAUNCH;
  }
  return SCFLOW_OK;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// 3x3 inverse through the adjugate, evaluated in fp64 and rounded once.  Replaces
// torch.inverse (fp32 LU) of models/utils/pose.py:36,38; both are within fp32 rounding of the
// exact inverse for the well-conditioned K and R the path sees.
__device__ __forceinline__ void inverse3x3(const float* __restrict__ m, float* __restrict__ o) {
  double a = m[0], b = m[1], c = m[2], d = m[3], e = m[4], f = m[5], g = m[6], h = m[7], i = m[8];
  double A = e * i - f * h, B = -(d * i - f * g), C = d * h - e * g;
  double det = a * A + b * B + c * C;
  double r = 1.0 / det;
  o[0] = (float)(A * r);
  o[1] = (float)(-(b * i - c * h) * r);
  o[2] = (float)((b * f - c * e) * r);
  o[3] = (float)(B * r);
  o[4] = (float)((a * i - c * g) * r);
  o[5] = (float)(-(a * f - c * d) * r);
  o[6] = (float)(C * r);
  o[7] = (float)(-(a * h - b * g) * r);
  o[8] = (float)((a * e - b * d) * r);
}

// SPLIT-BF16 operand format of the tensor-core convolutions (conv_tc.cu): v -> hi = rn_bf16(v), lo = rn_bf16(v - hi),
// two values per 32-bit word (low half = first value).  Packed conversion (cvt.rn.bf16x2.f32 = one F2FP in the
// ALU pipe; the scalar __float2bfloat16_rn is a quarter-rate F2F.BF16.F32).
__device__ __forceinline__ uint32_t cvt_bf16x2(float lo_half, float hi_half) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi_half), "f"(lo_half));
  return r;
}
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
  hi = cvt_bf16x2(a, b);
  lo = cvt_bf16x2(a - __uint_as_float(hi << 16), b - __uint_as_float(hi & 0xFFFF0000u));
}

}  // namespace scflow
