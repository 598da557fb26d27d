// scflow_pack_nhwc: concatenate / lay out fp32 channels-last and planar sources into a channel range of one
// channels-last buffer, optionally rounding to the nearest TF32 (the correlation kernel's operand rounding).
// Replaces the torch.cat + layout passes around models/decoder/raft_decoder.py:165 and the feature hand-over.
#include "common.cuh"

namespace scflow {
namespace pk {
// tcgen05 kind::tf32 ignores the low 13 mantissa bits (truncation, a biased error); operands we produce
// ourselves are rounded to nearest instead
__device__ __forceinline__ float rn_tf32(float v) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
  return __uint_as_float(u);
}
}  // namespace pk
}  // namespace scflow

namespace scflow {
namespace pk {
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
}  // namespace pk
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
    pk::pack4_nhwc_kernel<<<(unsigned)b4, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4*>(a), ca / 4, reinterpret_cast<float4*>(dst), cdst / 4, dst_off / 4,
        (unsigned)(total / 4), round_tf32);
    return check_launch();
  }
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  pk::pack_nhwc_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a, ca, b, cb, f, cf, dst, cdst,
                                                                          dst_off, npix, P, round_tf32);
  return check_launch();
}

