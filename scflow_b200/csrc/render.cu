// f4: the two pieces around the hot path that the reference leaves to third-party code or to numpy.
//
// 1. scflow_render_fwd -- what Renderer.forward (models/utils/rendering.py:185-248) asks of pytorch3d with the
//    shipped configuration (configs/refine_models/scflow_ycbv_real.py:261-274: HardPhongShader, one face per pixel,
//    blur radius 0, seperate_lights) and what the refiner reads back (models/refiner/base_refiner.py:106-112):
//    the Phong-shaded image, the view-space z-buffer and its mask, for a batch of (mesh label, R, t, K).
//    Three launches, no host synchronisation (the reference takes .item() of the batch's z range and joins
//    Python lists of meshes per call):
//      render_project_kernel   camera-frame vertices + their OpenCV image coordinates; per-image min z (light)
//      render_raster_kernel    one thread per (image, face): inside test at the pixel samples of its bounding box,
//                              perspective-correct z, 64-bit atomicMin of (z bits << 32 | face) -- the z-buffer
//      render_shade_kernel     one thread per pixel: barycentric interpolation of position / normal / colour,
//                              PointLights + Materials defaults of pytorch3d, hard blend over the background
//    Conventions (cameras_from_opencv_projection, rendering.py:17-47, + pytorch3d's pixel centres): pixel j samples
//    the OpenCV image coordinate (j + 0.5) (W - 1) / W.  Parity is against oracle/render_oracle.py (pytorch3d is
//    absent from this image: unpinned, see the oracle's header).
//
// 2. scflow_add_error_fwd -- ADD / ADD-S / 2-D projection error of ADD.eval_pose_error (metrics/add.py:354-394,
//    numpy on the host in the reference), pinned by tests/golden/add_eval.npz (outputs of the unmodified method).
//    ADD-S is the brute-force nearest neighbour the reference takes (V^2 distances per object): CTA = 128 ground-truth
//    vertices x all predicted vertices staged through shared memory.
#include "common.cuh"

namespace scflow {
namespace rd {

struct View {            // per image
  float R[9], t[3], fx, fy, cx, cy;
};

__device__ __forceinline__ View load_view(const float* R, const float* t, const float* K, int n) {
  View v;
#pragma unroll
  for (int i = 0; i < 9; ++i) v.R[i] = __ldg(R + n * 9 + i);
#pragma unroll
  for (int i = 0; i < 3; ++i) v.t[i] = __ldg(t + n * 3 + i);
  v.fx = __ldg(K + n * 9 + 0); v.fy = __ldg(K + n * 9 + 4); v.cx = __ldg(K + n * 9 + 2); v.cy = __ldg(K + n * 9 + 5);
  return v;
}

// float min through the integer atomics (valid for any sign)
__device__ __forceinline__ void atomic_min_float(float* addr, float v) {
  if (v >= 0) atomicMin(reinterpret_cast<int*>(addr), __float_as_int(v));
  else atomicMax(reinterpret_cast<unsigned*>(addr), __float_as_uint(v));
}

// cam (n, vmax, 4) = (u, v, z, 1/z) of every vertex of image n's mesh; zmin[n] = min camera z
__global__ void render_project_kernel(const float* __restrict__ verts, const int* __restrict__ voff,
                                      const long long* __restrict__ labels, const float* __restrict__ R,
                                      const float* __restrict__ t, const float* __restrict__ K,
                                      float4* __restrict__ cam, float* __restrict__ zmin, int vmax) {
  const int n = blockIdx.y;
  const int cls = (int)labels[n];
  const int v0 = voff[cls], nv = voff[cls + 1] - v0;
  const View vw = load_view(R, t, K, n);
  float zm = 3.0e38f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += gridDim.x * blockDim.x) {
    const float x = verts[3 * (v0 + i)], y = verts[3 * (v0 + i) + 1], z = verts[3 * (v0 + i) + 2];
    const float X = vw.R[0] * x + vw.R[1] * y + vw.R[2] * z + vw.t[0];
    const float Y = vw.R[3] * x + vw.R[4] * y + vw.R[5] * z + vw.t[1];
    const float Z = vw.R[6] * x + vw.R[7] * y + vw.R[8] * z + vw.t[2];
    const float iz = 1.0f / Z;
    cam[(size_t)n * vmax + i] = make_float4(vw.fx * X * iz + vw.cx, vw.fy * Y * iz + vw.cy, Z, iz);
    zm = fminf(zm, Z);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) zm = fminf(zm, __shfl_xor_sync(0xffffffffu, zm, o));
  if ((threadIdx.x & 31) == 0 && zm < 3.0e38f) atomic_min_float(zmin + n, zm);
}

struct Bary {
  float w0, w1, w2, z;
  bool inside;
};
// screen-space barycentrics of sample (sx, sy) in triangle (a, b, c) (float4 = u, v, z, 1/z), and the
// perspective-correct view z
__device__ __forceinline__ Bary bary_at(const float4& a, const float4& b, const float4& c, float inv_area, float sx,
                                        float sy) {
  Bary r;
  r.w0 = ((b.x - sx) * (c.y - sy) - (b.y - sy) * (c.x - sx)) * inv_area;
  r.w1 = ((c.x - sx) * (a.y - sy) - (c.y - sy) * (a.x - sx)) * inv_area;
  r.w2 = 1.0f - r.w0 - r.w1;
  r.inside = r.w0 >= 0.f && r.w1 >= 0.f && r.w2 >= 0.f;
  const float iz = r.w0 * a.w + r.w1 * b.w + r.w2 * c.w;
  r.z = iz > 0.f ? 1.0f / iz : 3.0e38f;
  return r;
}

// pixel j samples OpenCV coordinate (j + 0.5) * (W - 1) / W
__device__ __forceinline__ float sample_pos(int j, int W) { return ((float)j + 0.5f) * (float)(W - 1) / (float)W; }

__global__ void render_raster_kernel(const int* __restrict__ faces, const int* __restrict__ foff,
                                     const long long* __restrict__ labels, const float4* __restrict__ cam,
                                     unsigned long long* __restrict__ zkey, int vmax, int H, int W) {
  const int n = blockIdx.y;
  const int cls = (int)labels[n];
  const int f0 = foff[cls], nf = foff[cls + 1] - f0;
  const float4* cv = cam + (size_t)n * vmax;
  const float kx = (float)W / (float)(W - 1), ky = (float)H / (float)(H - 1);
  for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < nf; f += gridDim.x * blockDim.x) {
    const int ia = faces[3 * (f0 + f)], ib = faces[3 * (f0 + f) + 1], ic = faces[3 * (f0 + f) + 2];
    const float4 a = cv[ia], b = cv[ib], c = cv[ic];
    if (a.z <= 1e-6f || b.z <= 1e-6f || c.z <= 1e-6f) continue;
    const float area = (b.x - a.x) * (c.y - a.y) - (b.y - a.y) * (c.x - a.x);
    if (fabsf(area) < 1e-12f) continue;
    const float inv_area = 1.0f / area;
    // pixel indices whose samples can fall inside the bounding box (one pixel of slack, then exact tests)
    int j0 = (int)floorf(fminf(a.x, fminf(b.x, c.x)) * kx - 0.5f) - 1, j1 = (int)ceilf(fmaxf(a.x, fmaxf(b.x, c.x)) * kx - 0.5f) + 1;
    int i0 = (int)floorf(fminf(a.y, fminf(b.y, c.y)) * ky - 0.5f) - 1, i1 = (int)ceilf(fmaxf(a.y, fmaxf(b.y, c.y)) * ky - 0.5f) + 1;
    j0 = max(j0, 0); i0 = max(i0, 0); j1 = min(j1, W - 1); i1 = min(i1, H - 1);
    for (int i = i0; i <= i1; ++i) {
      const float sy = sample_pos(i, H);
      for (int j = j0; j <= j1; ++j) {
        const Bary r = bary_at(a, b, c, inv_area, sample_pos(j, W), sy);
        if (!r.inside || r.z >= 3.0e38f) continue;
        const unsigned long long key = ((unsigned long long)__float_as_uint(r.z) << 32) | (unsigned)f;
        atomicMin(zkey + ((size_t)n * H + i) * W + j, key);
      }
    }
  }
}

struct ShadeParams {
  const float *verts, *normals, *colors;
  const int *faces, *voff, *foff;
  const long long* labels;
  const float *R, *t, *K;
  const float4* cam;
  const float* zmin;
  const unsigned long long* zkey;
  float *image, *zbuf, *mask;
  int vmax, H, W;
  float bg[3], ambient, diffuse, specular, shininess;
};

__device__ __forceinline__ float3 nrm3(float3 v) {
  const float l = fmaxf(sqrtf(v.x * v.x + v.y * v.y + v.z * v.z), 1e-6f);
  return make_float3(v.x / l, v.y / l, v.z / l);
}

__global__ void render_shade_kernel(const ShadeParams p) {
  const int n = blockIdx.y;
  const int px = blockIdx.x * blockDim.x + threadIdx.x;
  if (px >= p.H * p.W) return;
  const int i = px / p.W, j = px - i * p.W;
  const size_t plane = (size_t)p.H * p.W;
  const unsigned long long key = p.zkey[(size_t)n * plane + px];
  float* img = p.image + (size_t)n * 3 * plane + px;
  if (key == ~0ull) {
    img[0] = p.bg[0]; img[plane] = p.bg[1]; img[2 * plane] = p.bg[2];
    p.zbuf[(size_t)n * plane + px] = -1.0f;
    if (p.mask) p.mask[(size_t)n * plane + px] = 0.0f;
    return;
  }
  const int cls = (int)p.labels[n];
  const int v0 = p.voff[cls], f = p.foff[cls] + (int)(key & 0xffffffffu);
  const int ia = p.faces[3 * f], ib = p.faces[3 * f + 1], ic = p.faces[3 * f + 2];
  const float4* cv = p.cam + (size_t)n * p.vmax;
  const float4 a = cv[ia], b = cv[ib], c = cv[ic];
  const float area = (b.x - a.x) * (c.y - a.y) - (b.y - a.y) * (c.x - a.x);
  const Bary r = bary_at(a, b, c, 1.0f / area, sample_pos(j, p.W), sample_pos(i, p.H));
  // perspective-correct barycentrics
  const float b0 = r.w0 * a.w * r.z, b1 = r.w1 * b.w * r.z, b2 = r.w2 * c.w * r.z;
  auto mix = [&](const float* base) {
    const float* A = base + 3 * (v0 + ia); const float* B = base + 3 * (v0 + ib); const float* C = base + 3 * (v0 + ic);
    return make_float3(b0 * A[0] + b1 * B[0] + b2 * C[0], b0 * A[1] + b1 * B[1] + b2 * C[1],
                       b0 * A[2] + b1 * B[2] + b2 * C[2]);
  };
  const float3 pos = mix(p.verts), tex = mix(p.colors);
  const float3 nr = nrm3(mix(p.normals));
  const View vw = load_view(p.R, p.t, p.K, n);
  // rendering.py:203-206: light at R @ (0, 0, max(min z - 400, 0)) in the object frame
  const float lz = fmaxf(p.zmin[n] - 400.0f, 0.0f);
  const float3 loc = make_float3(vw.R[2] * lz, vw.R[5] * lz, vw.R[8] * lz);
  const float3 ld = nrm3(make_float3(loc.x - pos.x, loc.y - pos.y, loc.z - pos.z));
  const float cosang = nr.x * ld.x + nr.y * ld.y + nr.z * ld.z;
  // camera centre in the object frame: -R^T t
  const float3 cc = make_float3(-(vw.R[0] * vw.t[0] + vw.R[3] * vw.t[1] + vw.R[6] * vw.t[2]),
                                -(vw.R[1] * vw.t[0] + vw.R[4] * vw.t[1] + vw.R[7] * vw.t[2]),
                                -(vw.R[2] * vw.t[0] + vw.R[5] * vw.t[1] + vw.R[8] * vw.t[2]));
  const float3 vd = nrm3(make_float3(cc.x - pos.x, cc.y - pos.y, cc.z - pos.z));
  const float3 rf = make_float3(-ld.x + 2.f * cosang * nr.x, -ld.y + 2.f * cosang * nr.y, -ld.z + 2.f * cosang * nr.z);
  const float sdot = fmaxf(vd.x * rf.x + vd.y * rf.y + vd.z * rf.z, 0.0f);
  const float spec = cosang > 0.f ? p.specular * powf(sdot, p.shininess) : 0.0f;
  const float lit = p.ambient + p.diffuse * fmaxf(cosang, 0.0f);
  img[0] = lit * tex.x + spec; img[plane] = lit * tex.y + spec; img[2 * plane] = lit * tex.z + spec;
  p.zbuf[(size_t)n * plane + px] = __uint_as_float((unsigned)(key >> 32));
  if (p.mask) p.mask[(size_t)n * plane + px] = 1.0f;
}

// ------------------------------------------------------------------------------------------- ADD / ADD-S
struct AddParams {
  const float *verts, *gt_r, *gt_t, *pr_r, *pr_t, *K;
  const int* voff;
  const long long* labels;
  const int* symmetric;          // per class
  double *e3, *e2;               // per object sums (divided by nv on the host side of the call: kernel 2)
  int n;
};

__device__ __forceinline__ void xform(const float* R, const float* t, float x, float y, float z, float& X, float& Y,
                                      float& Z) {
  X = R[0] * x + R[1] * y + R[2] * z + t[0];
  Y = R[3] * x + R[4] * y + R[5] * z + t[1];
  Z = R[6] * x + R[7] * y + R[8] * z + t[2];
}

constexpr int kAddThreads = 128;

// grid (vertex blocks, objects): thread = one model vertex under the ground-truth pose
__global__ void __launch_bounds__(kAddThreads)
add_error_kernel(const AddParams p) {
  __shared__ float sp[kAddThreads][3];
  __shared__ double red3[kAddThreads / 32], red2[kAddThreads / 32];
  const int n = blockIdx.y;
  const int cls = (int)p.labels[n];
  const int v0 = p.voff[cls], nv = p.voff[cls + 1] - v0;
  if ((int)(blockIdx.x * kAddThreads) >= nv) return;
  float Rg[9], tg[3], Rp[9], tp[3];
#pragma unroll
  for (int i = 0; i < 9; ++i) { Rg[i] = p.gt_r[n * 9 + i]; Rp[i] = p.pr_r[n * 9 + i]; }
#pragma unroll
  for (int i = 0; i < 3; ++i) { tg[i] = p.gt_t[n * 3 + i]; tp[i] = p.pr_t[n * 3 + i]; }
  const float fx = p.K[n * 9], fy = p.K[n * 9 + 4], cx = p.K[n * 9 + 2], cy = p.K[n * 9 + 5];
  const int i = blockIdx.x * kAddThreads + threadIdx.x;
  const bool live = i < nv;
  float gx = 0, gy = 0, gz = 1, qx = 0, qy = 0, qz = 1;
  if (live) {
    const float x = p.verts[3 * (v0 + i)], y = p.verts[3 * (v0 + i) + 1], z = p.verts[3 * (v0 + i) + 2];
    xform(Rg, tg, x, y, z, gx, gy, gz);
    xform(Rp, tp, x, y, z, qx, qy, qz);
  }
  // 2-D error always pairs the same vertex (add.py:392); project_3d_point divides by (z + 1e-8)
  float d2 = 0.f, d3;
  if (live) {
    const float gu = (fx * gx + cx * gz) / (gz + 1e-8f), gv = (fy * gy + cy * gz) / (gz + 1e-8f);
    const float pu = (fx * qx + cx * qz) / (qz + 1e-8f), pv = (fy * qy + cy * qz) / (qz + 1e-8f);
    d2 = sqrtf((gu - pu) * (gu - pu) + (gv - pv) * (gv - pv));
  }
  if (!p.symmetric[cls]) {
    d3 = live ? sqrtf((gx - qx) * (gx - qx) + (gy - qy) * (gy - qy) + (gz - qz) * (gz - qz)) : 0.f;
  } else {
    // ADD-S: nearest predicted vertex (argmin of the distance: the minimum itself is what enters the mean)
    float best = 3.0e38f;
    for (int j0 = 0; j0 < nv; j0 += kAddThreads) {
      __syncthreads();
      const int j = j0 + threadIdx.x;
      if (j < nv) {
        float X, Y, Z;
        xform(Rp, tp, p.verts[3 * (v0 + j)], p.verts[3 * (v0 + j) + 1], p.verts[3 * (v0 + j) + 2], X, Y, Z);
        sp[threadIdx.x][0] = X; sp[threadIdx.x][1] = Y; sp[threadIdx.x][2] = Z;
      }
      __syncthreads();
      const int cnt = min(kAddThreads, nv - j0);
      for (int k = 0; k < cnt; ++k) {
        const float dx = gx - sp[k][0], dy = gy - sp[k][1], dz = gz - sp[k][2];
        best = fminf(best, dx * dx + dy * dy + dz * dz);
      }
    }
    d3 = live ? sqrtf(best) : 0.f;
  }
  double s3 = d3, s2 = d2;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s3 += __shfl_xor_sync(0xffffffffu, s3, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
  if ((threadIdx.x & 31) == 0) { red3[threadIdx.x >> 5] = s3; red2[threadIdx.x >> 5] = s2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0;
    for (int w = 0; w < kAddThreads / 32; ++w) { a += red3[w]; b += red2[w]; }
    atomicAdd(p.e3 + n, a / nv);
    atomicAdd(p.e2 + n, b / nv);
  }
}

__global__ void fill_u64_kernel(unsigned long long* p, size_t n, unsigned long long v) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void fill_f32_kernel(float* p, int n, float v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

}  // namespace rd
}  // namespace scflow

using namespace scflow;

extern "C" size_t scflow_render_workspace_bytes(int n, int vmax, int h, int w) {
  if (n <= 0 || vmax <= 0 || h <= 0 || w <= 0) return 0;
  return (size_t)n * vmax * sizeof(float4) + (size_t)n * h * w * sizeof(unsigned long long) + 256 + (size_t)n * 4;
}

extern "C" int scflow_render_fwd(const float* verts, const float* normals, const float* colors, const int* faces,
                                 const int* voff, const int* foff, int vmax, int fmax, const long long* labels,
                                 const float* R, const float* t, const float* K, int n, int h, int w, float bg_r,
                                 float bg_g, float bg_b, float* image, float* zbuf, float* mask, void* workspace,
                                 void* stream) {
  if (!verts || !normals || !colors || !faces || !voff || !foff || !labels || !R || !t || !K || !image || !zbuf ||
      !workspace)
    return SCFLOW_E_NULL;
  if (n <= 0 || n > 65535 || vmax <= 0 || fmax <= 0 || h < 2 || w < 2) return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(workspace)) return SCFLOW_E_MISALIGNED;
  cudaStream_t st = (cudaStream_t)stream;
  float4* cam = static_cast<float4*>(workspace);
  size_t off = ((size_t)n * vmax * sizeof(float4) + 255) / 256 * 256;
  unsigned long long* zkey = reinterpret_cast<unsigned long long*>(static_cast<char*>(workspace) + off);
  float* zmin = reinterpret_cast<float*>(static_cast<char*>(workspace) + off + (size_t)n * h * w * 8);
  rd::fill_u64_kernel<<<148 * 8, 256, 0, st>>>(zkey, (size_t)n * h * w, ~0ull);
  rd::fill_f32_kernel<<<(n + 255) / 256, 256, 0, st>>>(zmin, n, 3.0e38f);
  rd::render_project_kernel<<<dim3((vmax + 255) / 256 < 64 ? (vmax + 255) / 256 : 64, n), 256, 0, st>>>(
      verts, voff, labels, R, t, K, cam, zmin, vmax);
  rd::render_raster_kernel<<<dim3((fmax + 127) / 128 < 256 ? (fmax + 127) / 128 : 256, n), 128, 0, st>>>(
      faces, foff, labels, cam, zkey, vmax, h, w);
  rd::ShadeParams p{verts, normals, colors, faces, voff, foff, labels, R, t, K, cam, zmin, zkey, image, zbuf, mask,
                    vmax, h, w, {bg_r, bg_g, bg_b}, 0.5f, 0.3f, 0.2f, 64.0f};   // PointLights / Materials defaults
  rd::render_shade_kernel<<<dim3((h * w + 255) / 256, n), 256, 0, st>>>(p);
  return check_launch();
}

extern "C" int scflow_add_error_fwd(const float* verts, const int* voff, int vmax, const long long* labels,
                                    const int* symmetric, const float* gt_r, const float* gt_t, const float* pred_r,
                                    const float* pred_t, const float* K, double* err3d, double* err2d, int n,
                                    void* stream) {
  if (!verts || !voff || !labels || !symmetric || !gt_r || !gt_t || !pred_r || !pred_t || !K || !err3d || !err2d)
    return SCFLOW_E_NULL;
  if (n <= 0 || n > 65535 || vmax <= 0) return SCFLOW_E_BAD_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(err3d, 0, sizeof(double) * n, st);
  cudaMemsetAsync(err2d, 0, sizeof(double) * n, st);
  rd::AddParams p{verts, gt_r, gt_t, pred_r, pred_t, K, voff, labels, symmetric, err3d, err2d, n};
  rd::add_error_kernel<<<dim3((vmax + rd::kAddThreads - 1) / rd::kAddThreads, n), rd::kAddThreads, 0, st>>>(p);
  return check_launch();
}
