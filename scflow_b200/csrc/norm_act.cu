// Channels-last normalisation glue of the RAFT encoders (models/encoder/raft_encoder.py:286-314,
// BasicBlock in models/backbone/resnet.py:12-92): InstanceNorm2d / eval BatchNorm2d + ReLU +
// residual add as TWO HBM-bound launches per layer instead of the reference's
// bias add_ + batch_norm(collect, transform) + relu_ + add + relu_ chain (each a full pass).
//
//   chan_stats_kernel   per (image, channel) sum and sum of squares over the pixels, fp32 partials
//                       folded into fp64 atomics: [n][C][2] doubles
//   norm_act_kernel     out = act( x*sc + sh  [+ res*rsc + rsh] ) with (sc, sh) either derived from
//                       the fp64 sums (instance norm: rstd, -mean*rstd), or a per-channel affine
//                       (eval batch norm folded with the convolution bias), or identity
//
// Layout (n, P, C) fp32, C % 4 == 0: a thread owns one float4 channel group and strides over the
// pixels of its chunk, so every warp access is a run of full 128-byte lines.
#include "common.cuh"
#include <stdlib.h>

namespace scflow {
namespace na {

constexpr int kThreads = 256;
constexpr int kChunk = 512;      // pixels per CTA

__global__ void __launch_bounds__(kThreads)
chan_stats_kernel(const float* __restrict__ x, double* __restrict__ sums, int P, int C) {
  const int groups = C >> 2;                 // float4 channel groups
  const int lanes = kThreads / groups;       // pixel lanes
  const int cg = threadIdx.x % groups, pl = threadIdx.x / groups;
  const int n = blockIdx.y;
  const int p0 = blockIdx.x * kChunk, p1 = min(p0 + kChunk, P);
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = s;
  if (pl < lanes) {
    const float4* src = reinterpret_cast<const float4*>(x + ((size_t)n * P) * C) + cg;
    auto acc = [&](const float4 v) {
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      q.x = fmaf(v.x, v.x, q.x); q.y = fmaf(v.y, v.y, q.y);
      q.z = fmaf(v.z, v.z, q.z); q.w = fmaf(v.w, v.w, q.w);
    };
    int p = p0 + pl;
    for (; p + 3 * lanes < p1; p += 4 * lanes) {      // 4 independent loads in flight
      float4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = __ldg(src + (size_t)(p + j * lanes) * groups);
#pragma unroll
      for (int j = 0; j < 4; ++j) acc(v[j]);
    }
    for (; p < p1; p += lanes) acc(__ldg(src + (size_t)p * groups));
  }
  // fold the pixel lanes of a channel group in fp64 through shared memory
  __shared__ double red[kThreads][8];
  double* mine = red[threadIdx.x];
  mine[0] = s.x; mine[1] = s.y; mine[2] = s.z; mine[3] = s.w;
  mine[4] = q.x; mine[5] = q.y; mine[6] = q.z; mine[7] = q.w;
  __syncthreads();
  for (int t = threadIdx.x; t < groups * 8; t += kThreads) {
    const int g = t >> 3, e = t & 7;
    double acc = 0.0;
    for (int l = 0; l < lanes; ++l) acc += red[l * groups + g][e];
    const int c = 4 * g + (e & 3);
    atomicAdd(&sums[((size_t)n * C + c) * 2 + (e >> 2)], acc);
  }
}

struct Side {
  const float* x;          // (n, P, C) or nullptr (side absent)
  const double* sums;      // instance-norm statistics [n][C][2], or nullptr
  const float* scale;      // per-channel affine [C] (with shift), or nullptr
  const float* shift;
};

__device__ __forceinline__ void side_coeffs(const Side& s, int n, int C, int c, int P, float eps,
                                            float* sc, float* sh) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (s.sums) {
      const double su = s.sums[((size_t)n * C + c + j) * 2], sq = s.sums[((size_t)n * C + c + j) * 2 + 1];
      const double mean = su / P;
      const double var = fmax(sq / P - mean * mean, 0.0);       // biased, as F.instance_norm
      const float rstd = 1.0f / sqrtf((float)var + eps);
      sc[j] = rstd;
      sh[j] = -(float)mean * rstd;
    } else if (s.scale) {
      sc[j] = s.scale[c + j];
      sh[j] = s.shift[c + j];
    } else {
      sc[j] = 1.0f;
      sh[j] = 0.0f;
    }
  }
}

__global__ void __launch_bounds__(kThreads)
norm_act_kernel(const Side a, const Side r, float* __restrict__ out, int P, int C, float eps, int relu) {
  const int groups = C >> 2;
  const int lanes = kThreads / groups;
  const int cg = threadIdx.x % groups, pl = threadIdx.x / groups;
  if (pl >= lanes) return;
  const int n = blockIdx.y;
  const int p0 = blockIdx.x * kChunk, p1 = min(p0 + kChunk, P);
  float sc[4], sh[4], rsc[4], rsh[4];
  side_coeffs(a, n, C, 4 * cg, P, eps, sc, sh);
  if (r.x) side_coeffs(r, n, C, 4 * cg, P, eps, rsc, rsh);
  const size_t base = ((size_t)n * P) * groups + cg;
  const float4* xa = reinterpret_cast<const float4*>(a.x) + base;
  const float4* xr = r.x ? reinterpret_cast<const float4*>(r.x) + base : nullptr;
  float4* o = reinterpret_cast<float4*>(out) + base;
  auto apply = [&](const float4 v, const float4 w) {
    float4 y;
    y.x = fmaf(v.x, sc[0], sh[0]); y.y = fmaf(v.y, sc[1], sh[1]);
    y.z = fmaf(v.z, sc[2], sh[2]); y.w = fmaf(v.w, sc[3], sh[3]);
    if (xr) {
      y.x += fmaf(w.x, rsc[0], rsh[0]); y.y += fmaf(w.y, rsc[1], rsh[1]);
      y.z += fmaf(w.z, rsc[2], rsh[2]); y.w += fmaf(w.w, rsc[3], rsh[3]);
    }
    if (relu) {
      y.x = fmaxf(y.x, 0.f); y.y = fmaxf(y.y, 0.f); y.z = fmaxf(y.z, 0.f); y.w = fmaxf(y.w, 0.f);
    }
    return y;
  };
  // four pixels per trip: 4 (8 with a residual) independent 16-byte loads in flight per thread
  int p = p0 + pl;
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  for (; p + 3 * lanes < p1; p += 4 * lanes) {
    float4 v[4], w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = xa[(size_t)(p + j * lanes) * groups];
#pragma unroll
    for (int j = 0; j < 4; ++j) w[j] = xr ? xr[(size_t)(p + j * lanes) * groups] : zero;
#pragma unroll
    for (int j = 0; j < 4; ++j) o[(size_t)(p + j * lanes) * groups] = apply(v[j], w[j]);
  }
  for (; p < p1; p += lanes)
    o[(size_t)p * groups] = apply(xa[(size_t)p * groups], xr ? xr[(size_t)p * groups] : zero);
}

}  // namespace na
}  // namespace scflow

static int check_shape(int n, int P, int C) {
  if (n <= 0 || P <= 0 || C < 4 || (C & 3) != 0 || C > 4 * scflow::na::kThreads || n > 65535)
    return SCFLOW_E_BAD_SHAPE;
  return SCFLOW_OK;
}

extern "C" int scflow_chan_stats_nhwc(const float* x, double* sums, int n, int P, int C, void* stream) {
  using namespace scflow;
  if (!x || !sums) return SCFLOW_E_NULL;
  if (int rc = check_shape(n, P, C)) return rc;
  if (!aligned16(x)) return SCFLOW_E_MISALIGNED;
  cudaStream_t st = (cudaStream_t)stream;
  if (cudaMemsetAsync(sums, 0, (size_t)n * C * 2 * sizeof(double), st) != cudaSuccess) return check_launch();
  dim3 grid((P + na::kChunk - 1) / na::kChunk, n);
  na::chan_stats_kernel<<<grid, na::kThreads, 0, st>>>(x, sums, P, C);
  return check_launch();
}

extern "C" int scflow_norm_act_nhwc(const float* x, const double* sums, const float* scale,
                                    const float* shift, const float* res, const double* res_sums,
                                    const float* res_scale, const float* res_shift, float* out,
                                    int n, int P, int C, float eps, int relu, void* stream) {
  using namespace scflow;
  if (!x || !out) return SCFLOW_E_NULL;
  if ((scale == nullptr) != (shift == nullptr) || (res_scale == nullptr) != (res_shift == nullptr))
    return SCFLOW_E_NULL;
  if (int rc = check_shape(n, P, C)) return rc;
  if (!aligned16(x) || !aligned16(out) || !aligned16(res)) return SCFLOW_E_MISALIGNED;
  na::Side a{x, sums, scale, shift};
  na::Side r{res, res_sums, res_scale, res_shift};
  dim3 grid((P + na::kChunk - 1) / na::kChunk, n);
  na::norm_act_kernel<<<grid, na::kThreads, 0, (cudaStream_t)stream>>>(a, r, out, P, C, eps, relu);
  return check_launch();
}

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
                       const float* __restrict__ beta, float* __restrict__ out, int P, int C, float eps,
                       int relu) {
  __shared__ double red[kGNThreads][2];
  __shared__ float stat[256][2];                     // per group: mean, rstd
  const int G = C >> 2;                              // groups == float4 columns
  const int lanes = kGNThreads / G;
  const int g = threadIdx.x % G, pl = threadIdx.x / G;
  const float4* src = reinterpret_cast<const float4*>(x + (size_t)blockIdx.x * P * C) + g;
  float4* dst = reinterpret_cast<float4*>(out + (size_t)blockIdx.x * P * C) + g;
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
    dst[(size_t)p * G] = y;
  }
}
}  // namespace na
}  // namespace scflow

extern "C" int scflow_group_norm4_act_nhwc(const float* x, const float* gamma, const float* beta, float* out,
                                           int n, int P, int C, float eps, int relu, void* stream) {
  using namespace scflow;
  if (!x || !gamma || !beta || !out) return SCFLOW_E_NULL;
  if (n <= 0 || P <= 0 || C < 4 || (C & 3) != 0 || (C >> 2) > 256 || na::kGNThreads % (C >> 2) != 0)
    return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(x) || !aligned16(out) || !aligned16(gamma) || !aligned16(beta)) return SCFLOW_E_MISALIGNED;
  na::group_norm4_act_kernel<<<n, na::kGNThreads, 0, (cudaStream_t)stream>>>(x, gamma, beta, out, P, C, eps, relu);
  return check_launch();
}

// ---------------------------------------------------------------------------------------------
// InstanceNorm + ReLU (+ residual) in ONE launch: a thread-block cluster owns an image, takes its
// statistics (first read: HBM), exchanges the per-CTA partial sums through distributed shared
// memory, and normalises its slice in place while that slice is still in L2 (second read: L2).
// HBM traffic per layer drops from 2 reads + 1 write to 1 read + 1 write.  One 1024-thread CTA
// per SM (the dynamic shared-memory request forces it), so the slices in flight -- 148 / cluster
// size images -- stay well inside the 126 MB L2.
//
// The statistics replay chan_stats_kernel exactly: the same 512-pixel chunks, the same 256-thread
// geometry inside a chunk, fp32 partials per thread in the same order, then fp64.  Only the
// order of the final fp64 additions differs (fixed here, atomic there).
#include <cooperative_groups.h>
namespace scflow {
namespace na {
namespace cg = cooperative_groups;

constexpr int kFThreads = 1024;
constexpr int kSub = kFThreads / kThreads;      // 4 chunk-sized sub-blocks per CTA
constexpr int kFusedSmem = 120 * 1024;          // > half an SM's shared memory: one CTA per SM

struct FusedParams {
  float* x;                // (n, P, C): normalised in place (out == x) or into out
  float* out;
  const float* res;        // optional residual (n, P, C), added as res*rsc + rsh (identity if null)
  const float* res_scale;
  const float* res_shift;
  double* sums_out;        // optional [n][C][2]: the statistics, for callers that want them
  int P, C, relu;
  float eps;
};

__global__ void __launch_bounds__(kFThreads, 1)
instnorm_act_cluster_kernel(const FusedParams p) {
  extern __shared__ __align__(16) uint8_t fsm[];
  double* red = reinterpret_cast<double*>(fsm);                    // [kFThreads][8]
  double* part = red + kFThreads * 8;                              // [2 * C] this CTA's partial sums
  float* coef = reinterpret_cast<float*>(part + 2 * 512);          // [2 * C] scale | shift
  cg::cluster_group cluster = cg::this_cluster();
  const int CL = (int)cluster.num_blocks(), rank = (int)cluster.block_rank();
  const int n = blockIdx.y, C = p.C, P = p.P;
  const int groups = C >> 2, lanes = kThreads / groups;
  const int sub = threadIdx.x / kThreads, vt = threadIdx.x % kThreads;   // sub-block, virtual thread
  const int cgi = vt % groups, pl = vt / groups;
  const int nchunks = (P + kChunk - 1) / kChunk;
  const float4* src = reinterpret_cast<const float4*>(p.x + ((size_t)n * P) * C) + cgi;

  // ---- pass 1: chunks rank*kSub + sub, + CL*kSub, ... ; per chunk the arithmetic of chan_stats_kernel.
  // The fp64 running sums live in shared memory (the registers go to 8 loads in flight per thread:
  // with one 1024-thread CTA per SM, HBM latency is covered by loads in flight, not by occupancy).
  double* mine = red + threadIdx.x * 8;
#pragma unroll
  for (int j = 0; j < 8; ++j) mine[j] = 0.0;
  if (pl < lanes) {
    for (int ch = rank * kSub + sub; ch < nchunks; ch += CL * kSub) {
      const int p0 = ch * kChunk, p1 = min(p0 + kChunk, P);
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = s;
      auto acc = [&](const float4 v) {
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        q.x = fmaf(v.x, v.x, q.x); q.y = fmaf(v.y, v.y, q.y);
        q.z = fmaf(v.z, v.z, q.z); q.w = fmaf(v.w, v.w, q.w);
      };
      int px = p0 + pl;
      for (; px + 7 * lanes < p1; px += 8 * lanes) {       // 8 independent loads in flight, summed in order
        float4 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __ldg(src + (size_t)(px + j * lanes) * groups);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc(v[j]);
      }
      for (; px < p1; px += lanes) acc(__ldg(src + (size_t)px * groups));
      mine[0] += s.x; mine[1] += s.y; mine[2] += s.z; mine[3] += s.w;
      mine[4] += q.x; mine[5] += q.y; mine[6] += q.z; mine[7] += q.w;
    }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < groups * 8; t += kFThreads) {
    const int g = t >> 3, e = t & 7;
    double a = 0.0;
    for (int sb = 0; sb < kSub; ++sb)
      for (int l = 0; l < lanes; ++l) a += red[(sb * kThreads + l * groups + g) * 8 + e];
    part[(4 * g + (e & 3)) * 2 + (e >> 2)] = a;            // [channel][sum | sum of squares]
  }
  cluster.sync();
  // ---- every CTA folds the cluster's partials (fixed order) into scale | shift of all channels
  for (int t = threadIdx.x; t < C; t += kFThreads) {
    double su = 0.0, sq = 0.0;
    for (int r = 0; r < CL; ++r) {
      const double* rp = cluster.map_shared_rank(part, r);
      su += rp[2 * t]; sq += rp[2 * t + 1];
    }
    const double mean = su / P;
    const double var = fmax(sq / P - mean * mean, 0.0);       // biased, as F.instance_norm
    const float rstd = 1.0f / sqrtf((float)var + p.eps);
    coef[t] = rstd;
    coef[C + t] = -(float)mean * rstd;
    if (p.sums_out && rank == 0) {
      p.sums_out[((size_t)n * C + t) * 2] = su;
      p.sums_out[((size_t)n * C + t) * 2 + 1] = sq;
    }
  }
  cluster.sync();      // remote partials stay alive until every CTA has read them; coef visible

  // ---- pass 2: the same chunks (still in L2), one pixel lane per thread
  if (pl >= lanes) return;
  float sc[4], sh[4], rsc[4], rsh[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    sc[j] = coef[4 * cgi + j]; sh[j] = coef[C + 4 * cgi + j];
    rsc[j] = p.res_scale ? p.res_scale[4 * cgi + j] : 1.0f;
    rsh[j] = p.res_shift ? p.res_shift[4 * cgi + j] : 0.0f;
  }
  const size_t base = ((size_t)n * P) * groups + cgi;
  const float4* xa = reinterpret_cast<const float4*>(p.x) + base;
  const float4* xr = p.res ? reinterpret_cast<const float4*>(p.res) + base : nullptr;
  float4* o = reinterpret_cast<float4*>(p.out) + base;
  const int relu = p.relu;
  auto apply = [&](const float4 v, const float4 w) {
    float4 y;
    y.x = fmaf(v.x, sc[0], sh[0]); y.y = fmaf(v.y, sc[1], sh[1]);
    y.z = fmaf(v.z, sc[2], sh[2]); y.w = fmaf(v.w, sc[3], sh[3]);
    if (xr) {
      y.x += fmaf(w.x, rsc[0], rsh[0]); y.y += fmaf(w.y, rsc[1], rsh[1]);
      y.z += fmaf(w.z, rsc[2], rsh[2]); y.w += fmaf(w.w, rsc[3], rsh[3]);
    }
    if (relu) {
      y.x = fmaxf(y.x, 0.f); y.y = fmaxf(y.y, 0.f); y.z = fmaxf(y.z, 0.f); y.w = fmaxf(y.w, 0.f);
    }
    return y;
  };
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int ch = rank * kSub + sub; ch < nchunks; ch += CL * kSub) {
    const int p0 = ch * kChunk, p1 = min(p0 + kChunk, P);
    int px = p0 + pl;
    for (; px + 3 * lanes < p1; px += 4 * lanes) {
      float4 v[4], w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = __ldcg(xa + (size_t)(px + j * lanes) * groups);
#pragma unroll
      for (int j = 0; j < 4; ++j) w[j] = xr ? __ldg(xr + (size_t)(px + j * lanes) * groups) : zero;
#pragma unroll
      for (int j = 0; j < 4; ++j) o[(size_t)(px + j * lanes) * groups] = apply(v[j], w[j]);
    }
    for (; px < p1; px += lanes)
      o[(size_t)px * groups] = apply(__ldcg(xa + (size_t)px * groups), xr ? __ldg(xr + (size_t)px * groups) : zero);
  }
}

}  // namespace na
}  // namespace scflow

extern "C" int scflow_instnorm_act_nhwc(float* x, const float* res, const float* res_scale,
                                        const float* res_shift, float* out, double* sums_out, int n,
                                        int P, int C, float eps, int relu, void* stream) {
  using namespace scflow;
  if (!x || !out) return SCFLOW_E_NULL;
  if ((res_scale == nullptr) != (res_shift == nullptr)) return SCFLOW_E_NULL;
  if (int rc = check_shape(n, P, C)) return rc;
  if (C > 512) return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(x) || !aligned16(out) || !aligned16(res)) return SCFLOW_E_MISALIGNED;
  na::FusedParams p{x, out, res, res_scale, res_shift, sums_out, P, C, relu, eps};
  // cluster size: enough chunks for every sub-block, at most 8 (portable)
  const int nchunks = (P + na::kChunk - 1) / na::kChunk;
  // cluster size: every sub-block gets at least one chunk; 4 for the large maps (measured on B200 at
  // 64 x 128x128x64: cluster 4 137 us, 8 143 us, 16 245 us; the statistics + normalise pair 148 us.
  // The second read of such a 268 MB layer does NOT stay in L2 whatever the cluster size; the
  // 64x64 and 32x32 layers gain 23-32 %.  Bounding the images in flight -- persistent clusters over
  // 10 / 14 / 18 images, i.e. 40-72 MB between a slice's two reads -- is slower still: 151-190 us)
  int cl = 8;
  while (cl > 1 && cl * na::kSub > nchunks) cl >>= 1;
  if (nchunks >= 32) cl = 4;
  const int smem = na::kFusedSmem;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(na::instnorm_act_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, na::kFusedSmem);
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cl, n, 1);
  cfg.blockDim = dim3(na::kFThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = cl; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  cfg.attrs = &attr; cfg.numAttrs = 1;
  if (cudaLaunchKernelEx(&cfg, na::instnorm_act_cluster_kernel, p) != cudaSuccess) return check_launch();
  return check_launch();
}
