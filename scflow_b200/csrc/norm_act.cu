// Channels-last glue kernels of the refiner that are not convolutions:
//   nhwc_to_nchw_kernel     channels-last features -> planar operands of the correlation GEMM (RN to TF32)
//   group_norm4_act_kernel  GroupNorm(32, 128) + ReLU of the pose head, one launch per layer
// (The InstanceNorm / BatchNorm passes of the RAFT encoders live in conv_tc.cu: statistics in the convolution
// epilogue + scflow_norm_apply_split; eval BatchNorm is folded into the packed filters.)
#include "common.cuh"
#include <stdlib.h>


// ---------------------------------------------------------------------------------------------
// (n, P, C) channels-last -> (n, C, P) planar, optionally rounding to the nearest TF32: hands the
// channels-last encoder features to the correlation GEMM (whose operands are MN-major = planar)
// already rounded, instead of letting tcgen05 truncate them (a biased error that the 8 refinement
// iterations amplify: 6.4e-5 rad on the cfg1 fixture against 6e-6 rounded).
namespace scflow {
namespace na {
__global__ void __launch_bounds__(256)
nhwc_to_nchw_kernel(const float* __restrict__ in, float* __restrict__ out, int P, int C, int round_tf32) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z, p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const float* src = in + (size_t)n * P * C;
  float* dst = out + (size_t)n * C * P;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int p = p0 + ty + 8 * j, c = c0 + tx;
    float v = 0.0f;
    if (p < P && c < C) v = __ldg(src + (size_t)p * C + c);
    if (round_tf32) {
      uint32_t u;
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
      v = __uint_as_float(u);
    }
    tile[ty + 8 * j][tx] = v;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = c0 + ty + 8 * j, p = p0 + tx;
    if (p < P && c < C) dst[(size_t)c * P + p] = tile[tx][ty + 8 * j];
  }
}
}  // namespace na
}  // namespace scflow

extern "C" int scflow_nhwc_to_nchw(const float* in, float* out, int n, int P, int C, int round_tf32,
                                   void* stream) {
  using namespace scflow;
  if (!in || !out) return SCFLOW_E_NULL;
  if (n <= 0 || P <= 0 || C <= 0 || n > 65535 || (C + 31) / 32 > 65535) return SCFLOW_E_BAD_SHAPE;
  dim3 grid((P + 31) / 32, (C + 31) / 32, n);
  na::nhwc_to_nchw_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in, out, P, C, round_tf32);
  return check_launch();
}

// ---------------------------------------------------------------------------------------------
// GroupNorm (4 channels per group) + ReLU on a channels-last (n, P, C) tensor, one CTA per sample:
// the pose head's conv -> GN -> ReLU layers (models/head/pose_head.py:128-137) work on 16x16 ... 4x4
// maps, where torch's three-kernel native_group_norm + clamp_min_ chain costs 30-55 us per layer
// against a 128 KB tensor.  A thread owns one group (= one float4) of a pixel column; statistics are
// folded in fp64 through shared memory; the second pass re-reads the sample from L1/L2.
namespace scflow {
namespace na {
constexpr int kGNThreads = 512;

__global__ void __launch_bounds__(kGNThreads)
group_norm4_act_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                       const float* __restrict__ beta, float* __restrict__ out, uint2* __restrict__ out_split,
                       long long out_pitch, int out_lo, int P, int C, float eps, int relu) {
  __shared__ double red[kGNThreads][2];
  __shared__ float stat[256][2];                     // per group: mean, rstd
  const int G = C >> 2;                              // groups == float4 columns
  const int lanes = kGNThreads / G;
  const int g = threadIdx.x % G, pl = threadIdx.x / G;
  const float4* src = reinterpret_cast<const float4*>(x + (size_t)blockIdx.x * P * C) + g;
  float4* dst = out ? reinterpret_cast<float4*>(out + (size_t)blockIdx.x * P * C) + g : nullptr;
  // split planes of the next convolution's operand: pixel pitch out_pitch bf16, lo plane out_lo further
  uint2* dsp = out_split ? out_split + ((size_t)blockIdx.x * P * out_pitch) / 4 + g : nullptr;
  float s = 0.f, q = 0.f;
  if (pl < lanes) {
    int p = pl;
    for (; p + 3 * lanes < P; p += 4 * lanes) {
      float4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = __ldg(src + (size_t)(p + j * lanes) * G);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
        q = fmaf(v[j].x, v[j].x, q); q = fmaf(v[j].y, v[j].y, q);
        q = fmaf(v[j].z, v[j].z, q); q = fmaf(v[j].w, v[j].w, q);
      }
    }
    for (; p < P; p += lanes) {
      const float4 v = __ldg(src + (size_t)p * G);
      s += (v.x + v.y) + (v.z + v.w);
      q = fmaf(v.x, v.x, q); q = fmaf(v.y, v.y, q); q = fmaf(v.z, v.z, q); q = fmaf(v.w, v.w, q);
    }
  }
  red[threadIdx.x][0] = s; red[threadIdx.x][1] = q;
  __syncthreads();
  if (threadIdx.x < G) {
    double su = 0.0, sq = 0.0;
    for (int l = 0; l < lanes; ++l) { su += red[l * G + threadIdx.x][0]; sq += red[l * G + threadIdx.x][1]; }
    const double cnt = 4.0 * P;
    const double mean = su / cnt;
    const double var = fmax(sq / cnt - mean * mean, 0.0);          // biased, as F.group_norm
    stat[threadIdx.x][0] = (float)mean;
    stat[threadIdx.x][1] = (float)(1.0 / sqrt(var + (double)eps));
  }
  __syncthreads();
  if (pl >= lanes) return;
  const float mean = stat[g][0], rstd = stat[g][1];
  const float4 ga = __ldg(reinterpret_cast<const float4*>(gamma) + g);
  const float4 be = __ldg(reinterpret_cast<const float4*>(beta) + g);
  const float4 sc = make_float4(ga.x * rstd, ga.y * rstd, ga.z * rstd, ga.w * rstd);
  const float4 sh = make_float4(be.x - mean * sc.x, be.y - mean * sc.y, be.z - mean * sc.z, be.w - mean * sc.w);
  for (int p = pl; p < P; p += lanes) {
    const float4 v = src[(size_t)p * G];
    float4 y = make_float4(fmaf(v.x, sc.x, sh.x), fmaf(v.y, sc.y, sh.y), fmaf(v.z, sc.z, sh.z), fmaf(v.w, sc.w, sh.w));
    if (relu) { y.x = fmaxf(y.x, 0.f); y.y = fmaxf(y.y, 0.f); y.z = fmaxf(y.z, 0.f); y.w = fmaxf(y.w, 0.f); }
    if (dst) dst[(size_t)p * G] = y;
    if (dsp) {
      uint2 h, l;
      split2(y.x, y.y, h.x, l.x);
      split2(y.z, y.w, h.y, l.y);
      dsp[((size_t)p * out_pitch) / 4] = h;
      dsp[((size_t)p * out_pitch + out_lo) / 4] = l;
    }
  }
}
}  // namespace na
}  // namespace scflow

extern "C" int scflow_group_norm4_act_nhwc(const float* x, const float* gamma, const float* beta, float* out,
                                           void* out_split, long long out_pitch, int out_lo, int n, int P, int C,
                                           float eps, int relu, void* stream) {
  using namespace scflow;
  if (!x || !gamma || !beta || (!out && !out_split)) return SCFLOW_E_NULL;
  if (out_split && ((out_pitch % 4) || (out_lo % 4) || (reinterpret_cast<uintptr_t>(out_split) & 7u)))
    return SCFLOW_E_MISALIGNED;
  if (n <= 0 || P <= 0 || C < 4 || (C & 3) != 0 || (C >> 2) > 256 || na::kGNThreads % (C >> 2) != 0)
    return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(x) || !aligned16(out) || !aligned16(gamma) || !aligned16(beta)) return SCFLOW_E_MISALIGNED;
  na::group_norm4_act_kernel<<<n, na::kGNThreads, 0, (cudaStream_t)stream>>>(
      x, gamma, beta, out, static_cast<uint2*>(out_split), out_pitch, out_lo, P, C, eps, relu);
  return check_launch();
}
