// Probe: K-major tf32 operands, TMA SWIZZLE_128B, UMMA layout SWIZZLE_128B, N = 256 + 128.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include "../scflow_b200/csrc/tc_common.cuh"
using namespace scflow;
constexpr int M = 128, N = 384, K = 32;

__global__ void __launch_bounds__(128, 1)
probe(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
      uint32_t lbo, uint32_t sbo, uint32_t kstep, float* D) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                 // 128 rows x 128 B
  uint8_t* sB = smem + 16384;         // 384 rows x 128 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 49152);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { tc::mbar_init(&bar[0], 1); tc::mbar_init(&bar[1], 1); tc::fence_barrier_init(); }
  if (warp == 0) { __syncwarp(); tc::tmem_alloc(slot, 512); }
  tc::tc_fence_before(); __syncthreads(); tc::tc_fence_after();
  const uint32_t tmem = *slot;
  if (threadIdx.x == 0) {
    tc::mbar_expect_tx(&bar[0], 16384 + 49152);
    tc::tma_load_2d(&mapA, &bar[0], sA, 0, 0);
    for (int i = 0; i < 3; ++i) tc::tma_load_2d(&mapB, &bar[0], sB + i * 16384, 0, i * 128);
    tc::mbar_wait(&bar[0], 0);
    tc::tc_fence_after();
    const uint32_t id256 = tc::make_idesc_tf32(128, 256, 0, 0), id128 = tc::make_idesc_tf32(128, 128, 0, 0);
    for (int ks = 0; ks < 4; ++ks) {
      uint64_t da = tc::make_smem_desc(tc::smem_u32(sA) + ks * kstep, lbo, sbo, tc::kLayoutSw128);
      uint64_t db = tc::make_smem_desc(tc::smem_u32(sB) + ks * kstep, lbo, sbo, tc::kLayoutSw128);
      uint64_t dq = tc::make_smem_desc(tc::smem_u32(sB) + 32768 + ks * kstep, lbo, sbo, tc::kLayoutSw128);
      tc::mma_tf32_ss(tmem, da, db, id256, ks != 0);
      tc::mma_tf32_ss(tmem + 256, da, dq, id128, ks != 0);
    }
    tc::tc_commit(&bar[1]);
  }
  __syncthreads();
  tc::mbar_wait(&bar[1], 0);
  tc::tc_fence_after();
  for (int c = 0; c < 12; ++c) {
    float x[32];
    tc::tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + c * 32, x);
    tc::tmem_ld_wait();
    for (int j = 0; j < 32; ++j) D[(warp * 32 + lane) * N + c * 32 + j] = x[j];
  }
  tc::tc_fence_before(); __syncthreads();
  if (warp == 0) { tc::tc_fence_after(); tc::tmem_dealloc(tmem, 512); }
}

int main() {
  std::vector<float> A(M * K), B(N * K), ref(M * N);
  for (int m = 0; m < M; ++m) for (int k = 0; k < K; ++k) A[m * K + k] = float(((k * 7 + m * 3) % 5) - 2);
  for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) B[n * K + k] = float(((k * 5 + n) % 7) - 3);
  for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) { float s = 0; for (int k = 0; k < K; ++k) s += A[m * K + k] * B[n * K + k]; ref[m * N + n] = s; }
  float *dA, *dB, *dD;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, M * N * 4);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
  CUtensorMap mapA, mapB;
  { uint64_t d[2] = {K, M}, s[1] = {K * 4}; uint32_t b[2] = {32, 128};
    if (!tc::encode_f32_map(&mapA, dA, 2, d, s, b, tc::kTmaSw128)) { printf("encode A failed\n"); return 1; } }
  { uint64_t d[2] = {K, N}, s[1] = {K * 4}; uint32_t b[2] = {32, 128};
    if (!tc::encode_f32_map(&mapB, dB, 2, d, s, b, tc::kTmaSw128)) { printf("encode B failed\n"); return 1; } }
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000);
  std::vector<float> D(M * N);
  uint32_t cfg[][3] = {{16, 1024, 32}, {0, 1024, 32}, {1024, 16, 32}};
  for (auto& c : cfg) {
    cudaMemset(dD, 0xFF, M * N * 4);
    probe<<<1, 128, 70000>>>(mapA, mapB, c[0], c[1], c[2], dD);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("lbo=%u sbo=%u: CUDA error %s\n", c[0], c[1], cudaGetErrorString(e)); return 2; }
    cudaMemcpy(D.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
    double e1 = 0, e2 = 0;
    for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) { double d = std::fabs(D[m * N + n] - ref[m * N + n]); if (n < 256) e1 = std::fmax(e1, d); else e2 = std::fmax(e2, d); }
    printf("K-major lbo=%u sbo=%u kstep=%u : maxerr(N=256 part)=%g maxerr(N=128 part)=%g  D[5][7]=%g ref=%g\n", c[0], c[1], c[2], e1, e2, D[5 * N + 7], ref[5 * N + 7]);
  }
  return 0;
}
