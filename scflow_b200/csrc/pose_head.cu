// Fully connected tail of MultiClassPoseHead (models/head/pose_head.py:139-147, 201-211):
//   fc0 (fc_in -> 1024) + ReLU, fc1 (1024 -> 256) + ReLU, rotation_pred (256 -> 6*classes),
//   translation_pred (256 -> 3*classes), index_select of the class of label[0].
// The reference runs four cuBLAS fp32 SGEMMs with M = batch (64): skinny problems on which the library
// picks a 32x32 SIMT tile (61 us for fc0 alone), plus 4 elementwise and 2 gather launches.  Three launches:
//
//   linear_splitk_kernel   partial[s] = x[:, Ks] . W[:, Ks]^T   grid (N/16, splitk, M/64): 256 CTAs for fc0
//                          (64 x 2048 x 1024), so every SM streams its share of the 8 MB weight once; run a
//                          second time for fc1, reducing fc0's partials (+bias, ReLU) on the way in
//   pose_tail_kernel       one CTA per crop: reduces fc1's partials (+bias, ReLU), then only the 9 rows of
//                          the selected class of the two predictors, a warp per row
//
// Exact fp32, deterministic (no atomics).
#include "common.cuh"

namespace scflow {
namespace ph {

constexpr int BM = 64, BN = 16, KC = 128;

// x: (M, K) row-major, or -- when xsplits > 0 -- the previous layer's partial sums (xsplits, M, K), reduced
// on the way in with its bias and ReLU; w: (N, K) row-major; part: (splitk, M, N)
__global__ void __launch_bounds__(256)
linear_splitk_kernel(const float* __restrict__ x, const float* __restrict__ xbias, int xsplits,
                     const float* __restrict__ w, float* __restrict__ part, int M, int K, int N, int kper) {
  __shared__ float xs[KC][BM + 1];
  __shared__ __align__(16) float ws[KC][BN];
  const int n0 = blockIdx.x * BN, split = blockIdx.y, m0 = blockIdx.z * BM;
  const int kbeg = split * kper, kend = min(K, kbeg + kper);
  const int tm = threadIdx.x & 63, tq = threadIdx.x >> 6;       // row, quad of columns
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k0 = kbeg; k0 < kend; k0 += KC) {
    __syncthreads();
    // x chunk, transposed on the way in: 64 rows x 128 k = 2048 float4, 8 per thread
    for (int i = threadIdx.x; i < BM * KC / 4; i += 256) {
      const int r = i / (KC / 4), kq = i - r * (KC / 4);
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m0 + r < M && k0 + 4 * kq < kend) {
        v = __ldg(reinterpret_cast<const float4*>(x + (size_t)(m0 + r) * K + k0) + kq);
        if (xsplits > 0) {
          for (int sp = 1; sp < xsplits; ++sp) {
            const float4 u = __ldg(reinterpret_cast<const float4*>(x + ((size_t)sp * M + m0 + r) * K + k0) + kq);
            v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w;
          }
          const float4 b = __ldg(reinterpret_cast<const float4*>(xbias + k0) + kq);
          v.x = fmaxf(v.x + b.x, 0.f); v.y = fmaxf(v.y + b.y, 0.f);
          v.z = fmaxf(v.z + b.z, 0.f); v.w = fmaxf(v.w + b.w, 0.f);
        }
      }
      xs[4 * kq][r] = v.x; xs[4 * kq + 1][r] = v.y; xs[4 * kq + 2][r] = v.z; xs[4 * kq + 3][r] = v.w;
    }
    for (int i = threadIdx.x; i < BN * KC / 4; i += 256) {
      const int r = i / (KC / 4), kq = i - r * (KC / 4);
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (n0 + r < N && k0 + 4 * kq < kend)
        v = __ldg(reinterpret_cast<const float4*>(w + (size_t)(n0 + r) * K + k0) + kq);
      ws[4 * kq][r] = v.x; ws[4 * kq + 1][r] = v.y; ws[4 * kq + 2][r] = v.z; ws[4 * kq + 3][r] = v.w;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < KC; ++k) {
      const float xv = xs[k][tm];
      const float4 wv = *reinterpret_cast<const float4*>(&ws[k][4 * tq]);
      acc[0] = fmaf(xv, wv.x, acc[0]); acc[1] = fmaf(xv, wv.y, acc[1]);
      acc[2] = fmaf(xv, wv.z, acc[2]); acc[3] = fmaf(xv, wv.w, acc[3]);
    }
  }
  if (m0 + tm < M) {
    float* o = part + ((size_t)split * M + m0 + tm) * N + n0 + 4 * tq;
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (n0 + 4 * tq + j < N) o[j] = acc[j];
  }
}

struct TailParams {
  const float *part1, *b1, *wr, *br, *wt, *bt;
  const long long* label;
  float *dr, *dt;
  int M, splitk, n1, num_class;
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// one CTA (9 warps + 1 idle) per crop: a1 = ReLU(b1 + sum of fc1 partials), then the 9 predictor rows of
// the class of label[0] (pose_head.py:209-210: index_select(dim=1, index=label)[:, 0, :])
__global__ void __launch_bounds__(320)
pose_tail_kernel(const TailParams p) {
  extern __shared__ __align__(16) float a1[];
  const int m = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int k = threadIdx.x; k < p.n1; k += blockDim.x) {
    float s = __ldg(p.b1 + k);
    for (int sp = 0; sp < p.splitk; ++sp) s += __ldg(p.part1 + ((size_t)sp * p.M + m) * p.n1 + k);
    a1[k] = fmaxf(s, 0.0f);
  }
  __syncthreads();
  // An out-of-range class makes the reference's index_select raise (a device assert on CUDA).  No host
  // synchronisation here: the pose of such a batch comes back as NaN instead of a silently clamped class.
  long long cls = p.label[0];
  const bool bad_class = cls < 0 || cls >= p.num_class;
  if (bad_class) cls = 0;
  if (warp < 9) {
    const int o = warp;
    const bool rot = o < 6;
    const int row = rot ? (int)cls * 6 + o : (int)cls * 3 + (o - 6);
    const float* wrow = (rot ? p.wr : p.wt) + (size_t)row * p.n1;
    float s = 0.0f;
    for (int k = lane; k < p.n1; k += 32) s = fmaf(__ldg(wrow + k), a1[k], s);
    s = warp_sum(s);
    if (lane == 0) {
      if (bad_class) s = __int_as_float(0x7fc00000);
      if (rot) p.dr[m * 6 + o] = s + __ldg(p.br + row);
      else p.dt[m * 3 + (o - 6)] = s + __ldg(p.bt + row);
    }
  }
}

}  // namespace ph
}  // namespace scflow

static int pick_split(int K, int want) {          // split so that every CTA gets whole 128-wide k chunks
  int chunks = (K + scflow::ph::KC - 1) / scflow::ph::KC;
  return want < chunks ? want : chunks;
}

extern "C" size_t scflow_pose_fc_workspace_floats(int M, int n0, int n1, int splitk) {
  return (size_t)splitk * M * ((size_t)n0 + n1);
}

extern "C" int scflow_pose_fc_fwd(const float* x, const float* w0, const float* b0, const float* w1,
                                  const float* b1, const float* wr, const float* br, const float* wt,
                                  const float* bt, const long long* label, float* workspace, float* dr,
                                  float* dt, int M, int K, int n0, int n1, int num_class, int splitk,
                                  void* stream) {
  using namespace scflow;
  if (!x || !w0 || !b0 || !w1 || !b1 || !wr || !br || !wt || !bt || !label || !workspace || !dr || !dt)
    return SCFLOW_E_NULL;
  if (M <= 0 || K <= 0 || (K & 3) != 0 || n0 <= 0 || (n0 & 3) != 0 || n1 <= 0 || num_class <= 0 ||
      splitk <= 0 || splitk > 64 || n1 > 8192)
    return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(x) || !aligned16(w0) || !aligned16(w1) || !aligned16(b0) || !aligned16(workspace))
    return SCFLOW_E_MISALIGNED;
  cudaStream_t st = (cudaStream_t)stream;
  float* part0 = workspace;
  float* part1 = workspace + (size_t)splitk * M * n0;
  const int mb = (M + ph::BM - 1) / ph::BM;
  const int s0 = pick_split(K, splitk), s1 = pick_split(n0, splitk);
  auto kper = [](int K_, int s) { int k = (K_ + s - 1) / s; return (k + ph::KC - 1) / ph::KC * ph::KC; };
  ph::linear_splitk_kernel<<<dim3((n0 + ph::BN - 1) / ph::BN, s0, mb), 256, 0, st>>>(
      x, nullptr, 0, w0, part0, M, K, n0, kper(K, s0));
  if (int rc = check_launch()) return rc;
  ph::linear_splitk_kernel<<<dim3((n1 + ph::BN - 1) / ph::BN, s1, mb), 256, 0, st>>>(
      part0, b0, s0, w1, part1, M, n0, n1, kper(n0, s1));
  if (int rc = check_launch()) return rc;
  ph::TailParams p{part1, b1, wr, br, wt, bt, label, dr, dt, M, s1, n1, num_class};
  ph::pose_tail_kernel<<<M, 320, (size_t)n1 * sizeof(float), st>>>(p);
  return check_launch();
}

// The tail alone: a1 = ReLU(b1 + sum of `splitk` partial-sum blocks of fc1) and the 9 predictor rows of the class of
// label[0].  With splitk = 1 `part1` is simply fc1's pre-activation without bias (M, n1): what the tensor-core
// path of the pose head (fc0 as an hs x ws convolution, fc1 as a 1x1 convolution on scflow_conv_tc_fwd) hands over.
extern "C" int scflow_pose_tail_fwd(const float* part1, const float* b1, const float* wr, const float* br,
                                    const float* wt, const float* bt, const long long* label, float* dr, float* dt,
                                    int M, int splitk, int n1, int num_class, void* stream) {
  using namespace scflow;
  if (!part1 || !b1 || !wr || !br || !wt || !bt || !label || !dr || !dt) return SCFLOW_E_NULL;
  if (M <= 0 || n1 <= 0 || n1 > 8192 || num_class <= 0 || splitk <= 0 || splitk > 64) return SCFLOW_E_BAD_SHAPE;
  ph::TailParams p{part1, b1, wr, br, wt, bt, label, dr, dt, M, splitk, n1, num_class};
  ph::pose_tail_kernel<<<M, 320, (size_t)n1 * sizeof(float), (cudaStream_t)stream>>>(p);
  return check_launch();
}

// ---------------------------------------------------------------------------------------------
// scflow_pose_fc_fwd_static: the same three launches with the split-K partials in a file-scope device array
// instead of a caller-supplied workspace, for callers that cannot allocate device memory (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The array (SCFLOW_POSE_FC_STATIC_MAX_FLOATS floats = 4 MiB) is held by every
// GPU that loads the library until the process exits.  In host code `g_pose_fc_scratch` is only a stand-in
// for the symbol: its DEVICE address is looked up once per device, with that device current, through
// cudaGetSymbolAddress, and that pointer is what the kernels get.  Every call on a device shares the array:
// nothing in it has to start at zero (both partial-sum blocks are fully overwritten before they are read),
// but calls that overlap on the GPU (two streams, two threads) race on it -- serialise them.
namespace scflow {
namespace ph {
static __device__ __align__(256) float g_pose_fc_scratch[SCFLOW_POSE_FC_STATIC_MAX_FLOATS];

constexpr int kMaxDevices = 64;
static float* g_scratch_addr[kMaxDevices];      // device address of the array on device i, 0 = not looked up

static int scratch_address(float** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    g_last_cuda_error = (int)cudaGetLastError();
    return SCFLOW_E_LAUNCH;
  }
  float* p = __atomic_load_n(&g_scratch_addr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    void* sym = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&sym, g_pose_fc_scratch);     // the current device's copy
    if (e != cudaSuccess || sym == nullptr) {
      g_last_cuda_error = (int)e;
      cudaGetLastError();
      return SCFLOW_E_LAUNCH;
    }
    p = static_cast<float*>(sym);
    __atomic_store_n(&g_scratch_addr[dev], p, __ATOMIC_RELEASE);             // idempotent: same value from any thread
  }
  *out = p;
  return SCFLOW_OK;
}
}  // namespace ph
}  // namespace scflow

extern "C" int scflow_pose_fc_fwd_static(const float* x, const float* w0, const float* b0, const float* w1,
                                         const float* b1, const float* wr, const float* br, const float* wt,
                                         const float* bt, const long long* label, float* dr, float* dt, int M,
                                         int K, int n0, int n1, int num_class, int splitk, void* stream) {
  using namespace scflow;
  if (M <= 0 || n0 <= 0 || n1 <= 0 || splitk <= 0) return SCFLOW_E_BAD_SHAPE;
  // the partial sums must fit the static array: reject, never truncate
  if (scflow_pose_fc_workspace_floats(M, n0, n1, splitk) > (size_t)SCFLOW_POSE_FC_STATIC_MAX_FLOATS)
    return SCFLOW_E_STATIC_SCRATCH;
  float* scratch = nullptr;
  if (int rc = ph::scratch_address(&scratch)) return rc;
  return scflow_pose_fc_fwd(x, w0, b0, w1, b1, wr, br, wt, bt, label, scratch, dr, dt, M, K, n0, n1, num_class,
                            splitk, stream);
}
