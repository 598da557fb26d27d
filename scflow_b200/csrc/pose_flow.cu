// Pose compose + pose-induced flow + 1/scale down-sample (one launch per iteration), the dense
// 2D->3D lift, and the align_corners=True bilinear resampling glue.
// Replaces models/utils/pose.py:26-88,124-169 and models/decoder/scflow_decoder.py:189-198,
// 223-228 (~6k tiny launches and 64 host syncs per batch-64 forward in the reference).
//
// HBM-bound elementwise work: 4 B/pixel read (depth; the object-frame point is recomputed) or
// 13 B/pixel (stored points + mask), 8 B/pixel written.  One thread = 4 consecutive pixels
// (float4 load / 2x float4 store); every CTA recomputes its crop's 3x3 algebra (a few dozen
// flops) instead of synchronising on another kernel.
#include "common.cuh"

namespace scflow {

struct PoseSmem {
  float R[9], t[3], K[9], Kinv[9], R0inv[9], t0[3];
};

// get_rotation_matrix_from_ortho6d + get_pose_from_delta_pose (pose.py:124-169), fp32, one
// thread.  F.normalize: v / max(||v||, 1e-12).
__device__ void compose_pose(const float* dR6, const float* dt, const float* Rp, const float* tp,
                             float weight, float* R, float* t) {
  float x0 = dR6[0], x1 = dR6[1], x2 = dR6[2], y0 = dR6[3], y1 = dR6[4], y2 = dR6[5];
  float nx = fmaxf(sqrtf(x0 * x0 + x1 * x1 + x2 * x2), 1e-12f);
  x0 = x0 / nx; x1 = x1 / nx; x2 = x2 / nx;
  float z0 = x1 * y2 - x2 * y1, z1 = x2 * y0 - x0 * y2, z2 = x0 * y1 - x1 * y0;
  float nz = fmaxf(sqrtf(z0 * z0 + z1 * z1 + z2 * z2), 1e-12f);
  z0 = z0 / nz; z1 = z1 / nz; z2 = z2 / nz;
  float w0 = z1 * x2 - z2 * x1, w1 = z2 * x0 - z0 * x2, w2 = z0 * x1 - z1 * x0;  // y = z cross x
  // dR has columns (x, y, z):  dR[i][0] = x_i, dR[i][1] = y_i, dR[i][2] = z_i
  const float dR[9] = {x0, w0, z0, x1, w1, z1, x2, w2, z2};
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      R[i * 3 + j] = dR[i * 3 + 0] * Rp[0 * 3 + j] + dR[i * 3 + 1] * Rp[1 * 3 + j] +
                     dR[i * 3 + 2] * Rp[2 * 3 + j];
  const float vz = tp[2] / expf(dt[2]);
  t[0] = vz * (dt[0] / weight + tp[0] / tp[2]);
  t[1] = vz * (dt[1] / weight + tp[1] / tp[2]);
  t[2] = vz;
}

__device__ __forceinline__ void mat3_mul_vec(const float* M, float a, float b, float c, float& x,
                                             float& y, float& z) {
  x = M[0] * a + M[1] * b + M[2] * c;
  y = M[3] * a + M[4] * b + M[5] * c;
  z = M[6] * a + M[7] * b + M[8] * c;
}

// object-frame point of pixel (px,py) with depth d  (lift_2d_to_3d, pose.py:34-38)
__device__ __forceinline__ void lift_pixel(const PoseSmem& s, float px, float py, float d, float& ox,
                                           float& oy, float& oz) {
  float cx, cy, cz;
  mat3_mul_vec(s.Kinv, px * d, py * d, d, cx, cy, cz);
  mat3_mul_vec(s.R0inv, cx - s.t0[0], cy - s.t0[1], cz - s.t0[2], ox, oy, oz);
}

// flow of an object-frame point under the current pose (pose.py:82-84)
__device__ __forceinline__ void project_flow(const PoseSmem& s, float ox, float oy, float oz, float px,
                                             float py, float& fx, float& fy) {
  float rx, ry, rz, ux, uy, uz;
  mat3_mul_vec(s.R, ox, oy, oz, rx, ry, rz);
  mat3_mul_vec(s.K, rx + s.t[0], ry + s.t[1], rz + s.t[2], ux, uy, uz);
  fx = ux / uz - px;
  fy = uy / uz - py;
}

struct PoseFlowParams {
  const float *dR6, *dt, *R_prev, *t_prev, *K, *R0, *t0, *depth, *points3d;
  const unsigned char* mask;
  float *R_out, *t_out, *flow_full, *flow_lr;
  int n, H, W, scale;
  float invalid, weight;
};

template <bool kFromPoints>
__device__ __forceinline__ void pixel_flow(const PoseFlowParams& p, const PoseSmem& s, int n, int x,
                                           int y, float& fx, float& fy) {
  const long long hw = (long long)p.H * p.W;
  const long long o = (long long)y * p.W + x;
  fx = p.invalid; fy = p.invalid;
  if (kFromPoints) {
    if (p.mask[n * hw + o]) {
      const float* pt = p.points3d + (long long)n * 3 * hw + o;
      project_flow(s, pt[0], pt[hw], pt[2 * hw], (float)x, (float)y, fx, fy);
    }
  } else {
    const float d = __ldg(p.depth + n * hw + o);
    if (d > 0.0f) {
      float ox, oy, oz;
      lift_pixel(s, (float)x, (float)y, d, ox, oy, oz);
      project_flow(s, ox, oy, oz, (float)x, (float)y, fx, fy);
    }
  }
}

constexpr int kPFThreads = 256;

template <bool kFromPoints>
__global__ void __launch_bounds__(kPFThreads)
pose_flow_kernel(const PoseFlowParams p) {
  __shared__ PoseSmem s;
  const int n = blockIdx.y;
  if (threadIdx.x == 0) {
    if (p.dR6 != nullptr) {
      compose_pose(p.dR6 + n * 6, p.dt + n * 3, p.R_prev + n * 9, p.t_prev + n * 3, p.weight, s.R, s.t);
    } else {
      for (int i = 0; i < 9; ++i) s.R[i] = p.R_prev[n * 9 + i];
      for (int i = 0; i < 3; ++i) s.t[i] = p.t_prev[n * 3 + i];
    }
    for (int i = 0; i < 9; ++i) s.K[i] = p.K[n * 9 + i];
    if (!kFromPoints) {
      inverse3x3(p.K + n * 9, s.Kinv);
      inverse3x3(p.R0 + n * 9, s.R0inv);
      for (int i = 0; i < 3; ++i) s.t0[i] = p.t0[n * 3 + i];
    }
    if (blockIdx.x == 0) {
      if (p.R_out) for (int i = 0; i < 9; ++i) p.R_out[n * 9 + i] = s.R[i];
      if (p.t_out) for (int i = 0; i < 3; ++i) p.t_out[n * 3 + i] = s.t[i];
    }
  }
  __syncthreads();

  const long long hw = (long long)p.H * p.W;
  // ---- full-resolution flow: 4 pixels per thread along x
  if (p.flow_full != nullptr) {
    const int quads = (int)(hw / 4);
    for (int i = blockIdx.x * kPFThreads + threadIdx.x; i < quads; i += gridDim.x * kPFThreads) {
      const int pix = i * 4;
      const int y = pix / p.W, x = pix - y * p.W;
      float fx[4], fy[4];
      if (kFromPoints) {
#pragma unroll
        for (int k = 0; k < 4; ++k) pixel_flow<true>(p, s, n, x + k, y, fx[k], fy[k]);
      } else {
        const float4 d4 = __ldg(reinterpret_cast<const float4*>(p.depth + n * hw + pix));
        const float d[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          fx[k] = p.invalid; fy[k] = p.invalid;
          if (d[k] > 0.0f) {
            float ox, oy, oz;
            lift_pixel(s, (float)(x + k), (float)y, d[k], ox, oy, oz);
            project_flow(s, ox, oy, oz, (float)(x + k), (float)y, fx[k], fy[k]);
          }
        }
      }
      float* o = p.flow_full + (long long)n * 2 * hw + pix;
      *reinterpret_cast<float4*>(o) = make_float4(fx[0], fx[1], fx[2], fx[3]);
      *reinterpret_cast<float4*>(o + hw) = make_float4(fy[0], fy[1], fy[2], fy[3]);
    }
  }

  // ---- low-resolution flow for the next iteration:  1/scale * bilinear(align_corners=True)
  // ATen upsample_bilinear2d: src = dst * (in-1)/(out-1); i0 = (int)src; i1 = i0 + (i0 < in-1);
  // l1 = src - i0; l0 = 1 - l1; out = h0*(w0*v00 + w1*v01) + h1*(w0*v10 + w1*v11)
  if (p.flow_lr != nullptr) {
    const int h = p.H / p.scale, w = p.W / p.scale;
    const float sy = h > 1 ? (float)(p.H - 1) / (float)(h - 1) : 0.0f;
    const float sx = w > 1 ? (float)(p.W - 1) / (float)(w - 1) : 0.0f;
    const float mul = 1.0f / (float)p.scale;
    for (int i = blockIdx.x * kPFThreads + threadIdx.x; i < h * w; i += gridDim.x * kPFThreads) {
      const int j = i / w, k = i - j * w;
      const float srcy = sy * (float)j, srcx = sx * (float)k;
      const int y0 = (int)srcy, x0 = (int)srcx;
      const int y1 = y0 + (y0 < p.H - 1 ? 1 : 0), x1 = x0 + (x0 < p.W - 1 ? 1 : 0);
      const float ly1 = srcy - (float)y0, ly0 = 1.0f - ly1;
      const float lx1 = srcx - (float)x0, lx0 = 1.0f - lx1;
      float ax, ay, bx, by, cx, cy, dx, dy;
      pixel_flow<kFromPoints>(p, s, n, x0, y0, ax, ay);
      pixel_flow<kFromPoints>(p, s, n, x1, y0, bx, by);
      pixel_flow<kFromPoints>(p, s, n, x0, y1, cx, cy);
      pixel_flow<kFromPoints>(p, s, n, x1, y1, dx, dy);
      const float vx = ly0 * (lx0 * ax + lx1 * bx) + ly1 * (lx0 * cx + lx1 * dx);
      const float vy = ly0 * (lx0 * ay + lx1 * by) + ly1 * (lx0 * cy + lx1 * dy);
      p.flow_lr[((long long)n * 2 + 0) * h * w + i] = mul * vx;
      p.flow_lr[((long long)n * 2 + 1) * h * w + i] = mul * vy;
    }
  }
}

// ---- dense lift (cal_3d_2d_corr without the ragged nonzero())
__global__ void __launch_bounds__(kPFThreads)
lift_points_kernel(const float* depth, const float* K, const float* R0, const float* t0,
                   float* points3d, unsigned char* mask, int H, int W) {
  __shared__ PoseSmem s;
  const int n = blockIdx.y;
  if (threadIdx.x == 0) {
    inverse3x3(K + n * 9, s.Kinv);
    inverse3x3(R0 + n * 9, s.R0inv);
    for (int i = 0; i < 3; ++i) s.t0[i] = t0[n * 3 + i];
  }
  __syncthreads();
  const long long hw = (long long)H * W;
  for (long long i = (long long)blockIdx.x * kPFThreads + threadIdx.x; i < hw;
       i += (long long)gridDim.x * kPFThreads) {
    const int y = (int)(i / W), x = (int)(i - (long long)y * W);
    const float d = depth[n * hw + i];
    float ox = 0.f, oy = 0.f, oz = 0.f;
    if (d > 0.0f) lift_pixel(s, (float)x, (float)y, d, ox, oy, oz);
    float* o = points3d + (long long)n * 3 * hw + i;
    o[0] = ox; o[hw] = oy; o[2 * hw] = oz;
    if (mask) mask[n * hw + i] = d > 0.0f ? 1 : 0;
  }
}

// ---- align_corners=True bilinear resampling, out = mul * resize(in [+ add])
__global__ void __launch_bounds__(256)
resize_kernel(const float* __restrict__ in, const float* __restrict__ add, float* __restrict__ out,
              int planes, int ih, int iw, int oh, int ow, float mul) {
  const float sy = oh > 1 ? (float)(ih - 1) / (float)(oh - 1) : 0.0f;
  const float sx = ow > 1 ? (float)(iw - 1) / (float)(ow - 1) : 0.0f;
  const long long total = (long long)planes * oh * ow;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(i % ow);
    const int y = (int)((i / ow) % oh);
    const long long pl = i / ((long long)ow * oh);
    const float srcy = sy * (float)y, srcx = sx * (float)x;
    const int y0 = (int)srcy, x0 = (int)srcx;
    const int y1 = y0 + (y0 < ih - 1 ? 1 : 0), x1 = x0 + (x0 < iw - 1 ? 1 : 0);
    const float ly1 = srcy - (float)y0, ly0 = 1.0f - ly1;
    const float lx1 = srcx - (float)x0, lx0 = 1.0f - lx1;
    const float* a = in + pl * ih * iw;
    float v00 = a[y0 * iw + x0], v01 = a[y0 * iw + x1], v10 = a[y1 * iw + x0], v11 = a[y1 * iw + x1];
    if (add != nullptr) {
      const float* b = add + pl * ih * iw;
      v00 += b[y0 * iw + x0]; v01 += b[y0 * iw + x1]; v10 += b[y1 * iw + x0]; v11 += b[y1 * iw + x1];
    }
    out[i] = mul * (ly0 * (lx0 * v00 + lx1 * v01) + ly1 * (lx0 * v10 + lx1 * v11));
  }
}

// Same arithmetic, four consecutive outputs of a row per thread (one 16-byte store), 32-bit index
// math: the scalar kernel above spends its time in 64-bit div/mod (43 us for 8.4 M outputs).
__global__ void __launch_bounds__(256)
resize4_kernel(const float* __restrict__ in, const float* __restrict__ add, float* __restrict__ out,
               unsigned planes, int ih, int iw, int oh, int ow, float mul) {
  const float sy = oh > 1 ? (float)(ih - 1) / (float)(oh - 1) : 0.0f;
  const float sx = ow > 1 ? (float)(iw - 1) / (float)(ow - 1) : 0.0f;
  const unsigned qw = (unsigned)ow >> 2;                 // quads per row
  const unsigned total = planes * (unsigned)oh * qw;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned row = i / qw;                         // plane * oh + y
    const int x4 = (int)(i - row * qw) * 4;
    const unsigned pl = row / (unsigned)oh;
    const int y = (int)(row - pl * (unsigned)oh);
    const float srcy = sy * (float)y;
    const int y0 = (int)srcy;
    const int y1 = y0 + (y0 < ih - 1 ? 1 : 0);
    const float ly1 = srcy - (float)y0, ly0 = 1.0f - ly1;
    const float* a0 = in + (size_t)pl * ih * iw + y0 * iw;
    const float* a1 = in + (size_t)pl * ih * iw + y1 * iw;
    const float* b0 = add ? add + (size_t)pl * ih * iw + y0 * iw : nullptr;
    const float* b1 = add ? add + (size_t)pl * ih * iw + y1 * iw : nullptr;
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float srcx = sx * (float)(x4 + j);
      const int x0 = (int)srcx;
      const int x1 = x0 + (x0 < iw - 1 ? 1 : 0);
      const float lx1 = srcx - (float)x0, lx0 = 1.0f - lx1;
      float v00 = __ldg(a0 + x0), v01 = __ldg(a0 + x1), v10 = __ldg(a1 + x0), v11 = __ldg(a1 + x1);
      if (add != nullptr) {
        v00 += __ldg(b0 + x0); v01 += __ldg(b0 + x1); v10 += __ldg(b1 + x0); v11 += __ldg(b1 + x1);
      }
      o[j] = mul * (ly0 * (lx0 * v00 + lx1 * v01) + ly1 * (lx0 * v10 + lx1 * v11));
    }
    *reinterpret_cast<float4*>(out + (size_t)row * ow + x4) = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// Up-sampling by >= 4: the four outputs of a quad read at most three neighbouring source columns, so
// the 2 x 3 source values (+ the optional addend) are loaded once per quad instead of four corner
// loads per output (32 scalar loads per quad made the kernel LSU-bound: 18-20 us for 8.4 M outputs).
// Same arithmetic per output as resize4_kernel.
__global__ void __launch_bounds__(256)
resize4_up_kernel(const float* __restrict__ in, const float* __restrict__ add, float* __restrict__ out,
                  unsigned planes, int ih, int iw, int oh, int ow, float mul) {
  const float sy = oh > 1 ? (float)(ih - 1) / (float)(oh - 1) : 0.0f;
  const float sx = ow > 1 ? (float)(iw - 1) / (float)(ow - 1) : 0.0f;
  const unsigned qw = (unsigned)ow >> 2;                 // quads per row
  const unsigned total = planes * (unsigned)oh * qw;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned row = i / qw;                         // plane * oh + y
    const int x4 = (int)(i - row * qw) * 4;
    const unsigned pl = row / (unsigned)oh;
    const int y = (int)(row - pl * (unsigned)oh);
    const float srcy = sy * (float)y;
    const int y0 = (int)srcy;
    const int y1 = y0 + (y0 < ih - 1 ? 1 : 0);
    const float ly1 = srcy - (float)y0, ly0 = 1.0f - ly1;
    const size_t r0 = (size_t)pl * ih * iw + y0 * iw, r1 = (size_t)pl * ih * iw + y1 * iw;
    const int xb = (int)(sx * (float)x4);                // first output's left column
    float c0[3], c1[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int xc = min(xb + k, iw - 1);
      c0[k] = __ldg(in + r0 + xc);
      c1[k] = __ldg(in + r1 + xc);
      if (add != nullptr) { c0[k] += __ldg(add + r0 + xc); c1[k] += __ldg(add + r1 + xc); }
    }
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float srcx = sx * (float)(x4 + j);
      const int x0 = (int)srcx;
      const float lx1 = srcx - (float)x0, lx0 = 1.0f - lx1;
      const bool hi = x0 != xb;                          // x0 - xb is 0 or 1 (scale >= 4)
      const float v00 = hi ? c0[1] : c0[0], v01 = hi ? c0[2] : c0[1];
      const float v10 = hi ? c1[1] : c1[0], v11 = hi ? c1[2] : c1[1];
      o[j] = mul * (ly0 * (lx0 * v00 + lx1 * v01) + ly1 * (lx0 * v10 + lx1 * v11));
    }
    *reinterpret_cast<float4*>(out + (size_t)row * ow + x4) = make_float4(o[0], o[1], o[2], o[3]);
  }
}

}  // namespace scflow

extern "C" int scflow_pose_flow_fwd(const float* dR6, const float* dt, const float* R_prev,
                                    const float* t_prev, const float* K, const float* R0,
                                    const float* t0, const float* depth, const float* points3d,
                                    const unsigned char* mask, float* R_out, float* t_out,
                                    float* flow_full, float* flow_lr, int n, int H, int W,
                                    int scale, float invalid_num, float weight, void* stream) {
  using namespace scflow;
  if (!R_prev || !t_prev || !K) return SCFLOW_E_NULL;
  if ((dR6 == nullptr) != (dt == nullptr)) return SCFLOW_E_NULL;
  const bool from_points = points3d != nullptr;
  if (from_points ? (mask == nullptr) : (!depth || !R0 || !t0)) return SCFLOW_E_NULL;
  if (n <= 0 || H <= 0 || W <= 0 || (W % 4) != 0) return SCFLOW_E_BAD_SHAPE;
  if (flow_lr && (scale <= 0 || H % scale || W % scale)) return SCFLOW_E_BAD_SHAPE;
  if ((flow_full && !aligned16(flow_full)) || (!from_points && !aligned16(depth)))
    return SCFLOW_E_MISALIGNED;
  PoseFlowParams p{dR6, dt, R_prev, t_prev, K, R0, t0, depth, points3d, mask,
                   R_out, t_out, flow_full, flow_lr, n, H, W, scale, invalid_num, weight};
  long long quads = (long long)H * W / 4;
  // every CTA first composes the pose and inverts K and R0 in one thread (fp64), so few, fat CTAs
  // per crop: the whole grid (about 4 CTAs per SM at 64 crops) is resident in one wave
  int bx = (int)((quads + kPFThreads - 1) / kPFThreads);
  const int cap = n >= 64 ? 8 : n >= 16 ? 16 : 64;
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  dim3 grid(bx, n);
  if (from_points)
    pose_flow_kernel<true><<<grid, kPFThreads, 0, (cudaStream_t)stream>>>(p);
  else
    pose_flow_kernel<false><<<grid, kPFThreads, 0, (cudaStream_t)stream>>>(p);
  return check_launch();
}

extern "C" int scflow_lift_points(const float* depth, const float* K, const float* R0,
                                  const float* t0, float* points3d, unsigned char* mask, int n,
                                  int H, int W, void* stream) {
  using namespace scflow;
  if (!depth || !K || !R0 || !t0 || !points3d) return SCFLOW_E_NULL;
  if (n <= 0 || H <= 0 || W <= 0) return SCFLOW_E_BAD_SHAPE;
  long long hw = (long long)H * W;
  int bx = (int)((hw + kPFThreads - 1) / kPFThreads);
  if (bx > 64) bx = 64;
  lift_points_kernel<<<dim3(bx, n), kPFThreads, 0, (cudaStream_t)stream>>>(depth, K, R0, t0,
                                                                           points3d, mask, H, W);
  return check_launch();
}

static int launch_resize(const float* in, const float* add, float* out, long long planes, int ih,
                         int iw, int oh, int ow, float mul, void* stream) {
  long long total = planes * oh * ow;
  if ((ow & 3) == 0 && total < (1ll << 32) && scflow::aligned16(out)) {
    long long blocks4 = (total / 4 + 255) / 256;
    if (blocks4 > 148LL * 32) blocks4 = 148LL * 32;
    if (ow >= 4 * iw && iw >= 2)
      scflow::resize4_up_kernel<<<(unsigned)blocks4, 256, 0, (cudaStream_t)stream>>>(in, add, out, (unsigned)planes,
                                                                                   ih, iw, oh, ow, mul);
    else
      scflow::resize4_kernel<<<(unsigned)blocks4, 256, 0, (cudaStream_t)stream>>>(in, add, out, (unsigned)planes,
                                                                                ih, iw, oh, ow, mul);
    return scflow::check_launch();
  }
  long long blocks = (total + 255) / 256;
  if (blocks > 148LL * 16) blocks = 148LL * 16;
  scflow::resize_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(in, add, out, (int)planes,
                                                                           ih, iw, oh, ow, mul);
  return scflow::check_launch();
}

extern "C" int scflow_upsample_fwd(const float* in, const float* add, float* out, int n, int c,
                                   int h, int w, int scale, float mul, void* stream) {
  if (!in || !out) return SCFLOW_E_NULL;
  if (n <= 0 || c <= 0 || h <= 0 || w <= 0 || scale <= 0) return SCFLOW_E_BAD_SHAPE;
  return launch_resize(in, add, out, (long long)n * c, h, w, h * scale, w * scale, mul, stream);
}

extern "C" int scflow_downsample_fwd(const float* in, float* out, int n, int c, int H, int W,
                                     int scale, float mul, void* stream) {
  if (!in || !out) return SCFLOW_E_NULL;
  if (n <= 0 || c <= 0 || scale <= 0 || H % scale || W % scale) return SCFLOW_E_BAD_SHAPE;
  return launch_resize(in, nullptr, out, (long long)n * c, H, W, H / scale, W / scale, mul, stream);
}
