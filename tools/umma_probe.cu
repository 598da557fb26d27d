// Stand-alone probe: one TMA load + 4 tcgen05.mma kind::tf32 with runtime descriptor variants.
// Dev aid for pinning down UMMA descriptor semantics on the GPU box; not part of the library.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include "../scflow_b200/csrc/tc_common.cuh"
using namespace scflow;

constexpr int M = 128, N = 256, K = 32;

struct Variant { uint32_t lbo, sbo, layout, kstride, idesc, version; };

__global__ void __launch_bounds__(128, 1)
probe(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
      Variant v, float* D, float* dumpA) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                 // 16 KB
  uint8_t* sB = smem + 16384;         // 32 KB
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 49152);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { tc::mbar_init(&bar[0], 1); tc::mbar_init(&bar[1], 1); tc::fence_barrier_init(); }
  if (warp == 0) { __syncwarp(); tc::tmem_alloc(slot, 512); }
  tc::tc_fence_before(); __syncthreads(); tc::tc_fence_after();
  const uint32_t tmem = *slot;
  if (threadIdx.x == 0) {
    tc::mbar_expect_tx(&bar[0], 49152);
    tc::tma_load_4d(&mapA, &bar[0], sA, 0, 0, 0, 0);
    tc::tma_load_4d(&mapB, &bar[0], sB, 0, 0, 0, 0);
    tc::mbar_wait(&bar[0], 0);
    tc::tc_fence_after();
    for (int kk = 0; kk < 4; ++kk) {
      uint64_t da = tc::make_smem_desc(tc::smem_u32(sA) + kk * v.kstride, v.lbo, v.sbo, v.layout);
      uint64_t db = tc::make_smem_desc(tc::smem_u32(sB) + kk * v.kstride, v.lbo, v.sbo, v.layout);
      if (!v.version) { da &= ~((uint64_t)3 << 46); db &= ~((uint64_t)3 << 46); }
      tc::mma_tf32_ss(tmem, da, db, v.idesc, kk != 0);
    }
    tc::tc_commit(&bar[1]);
  }
  __syncthreads();
  tc::mbar_wait(&bar[1], 0);
  tc::tc_fence_after();
  for (int i = threadIdx.x; i < 4096; i += 128) dumpA[i] = reinterpret_cast<float*>(sA)[i];
  for (int c = 0; c < 8; ++c) {
    float x[32];
    tc::tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + c * 32, x);
    tc::tmem_ld_wait();
    for (int j = 0; j < 32; ++j) D[(warp * 32 + lane) * N + c * 32 + j] = x[j];
  }
  tc::tc_fence_before(); __syncthreads();
  if (warp == 0) { tc::tc_fence_after(); tc::tmem_dealloc(tmem, 512); }
}

int main(int argc, char** argv) {
  const int sel = argc > 1 ? atoi(argv[1]) : -1;   // -1: smem dump only;  >=0: variant index;  100: K-major test

  std::vector<float> A(K * M), B(K * N), ref(M * N, 0.f);
  for (int k = 0; k < K; ++k) for (int m = 0; m < M; ++m) A[k * M + m] = float(((k * 7 + m * 3) % 5) - 2);
  for (int k = 0; k < K; ++k) for (int n = 0; n < N; ++n) B[k * N + n] = float(((k * 5 + n) % 7) - 3);
  for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) { float s = 0; for (int k = 0; k < K; ++k) s += A[k * M + m] * B[k * N + n]; ref[m * N + n] = s; }
  float *dA, *dB, *dD, *dDump;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, M * N * 4); cudaMalloc(&dDump, 16384);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
  const int tmasw = (sel >= 10 && sel < 100) ? tc::kTmaSw128Atom32 : tc::kTmaSw128;
  CUtensorMap mapA, mapB;
  { uint64_t d[4] = {32, K, M / 32, 1}, s[3] = {(uint64_t)M * 4, 128, (uint64_t)K * M * 4}; uint32_t b[4] = {32, 32, 4, 1};
    if (!tc::encode_f32_map(&mapA, dA, 4, d, s, b, tmasw)) { printf("encode A failed\n"); return 1; } }
  { uint64_t d[4] = {32, K, N / 32, 1}, s[3] = {(uint64_t)N * 4, 128, (uint64_t)K * N * 4}; uint32_t b[4] = {32, 32, 8, 1};
    if (!tc::encode_f32_map(&mapB, dB, 4, d, s, b, tmasw)) { printf("encode B failed\n"); return 1; } }
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 52000);
  std::vector<float> D(M * N), dump(4096);
  const uint32_t idesc_mn = tc::make_idesc_tf32(M, N, 1, 1);
  Variant vs[] = {
    {4096, 1024, 2, 1024, idesc_mn, 1}, {1024, 4096, 2, 1024, idesc_mn, 1},
    {4096, 512, 1, 1024, idesc_mn, 1}, {512, 4096, 1, 1024, idesc_mn, 1}, {4096, 1024, 1, 1024, idesc_mn, 1},
  };

  const int nv = sizeof(vs) / sizeof(vs[0]);
  for (int vi = 0; vi < nv; ++vi) {
    if (!(sel % 10 == vi || (sel == -1 && vi == 0))) continue;
    auto& v = vs[vi];
    cudaMemset(dD, 0xFF, M * N * 4);
    probe<<<1, 128, 52000>>>(mapA, mapB, v, dD, dDump);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("variant lbo=%u sbo=%u: CUDA error %s\n", v.lbo, v.sbo, cudaGetErrorString(e)); return 2; }
    cudaMemcpy(D.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(dump.data(), dDump, 16384, cudaMemcpyDeviceToHost);
    double maxerr = 0; int nz = 0, nan = 0;
    for (int i = 0; i < M * N; ++i) { if (std::isnan(D[i])) { ++nan; continue; } if (D[i] != 0) ++nz; maxerr = std::fmax(maxerr, std::fabs(D[i] - ref[i])); }
    printf("lbo=%u sbo=%u layout=%u ver=%u idesc=%08x : maxerr=%g nonzero=%d nan=%d  D[0][0..3]=%g %g %g %g ref=%g %g %g %g\n",
           v.lbo, v.sbo, v.layout, v.version, v.idesc, maxerr, nz, nan, D[0], D[1], D[2], D[3], ref[0], ref[1], ref[2], ref[3]);
  }
  // smem layout check of the A tile: element (m, k) expected at group g=m/32, row k, chunk ((m%32)/4) ^ (k%8)
  int bad = 0, zeros = 0;
  for (int m = 0; m < M; ++m) for (int k = 0; k < K; ++k) {
    int g = m / 32, p = m % 32;
    int idx = (tmasw == tc::kTmaSw128) ? g * 1024 + k * 32 + (((p / 4) ^ (k % 8)) * 4) + (p % 4)
                                       : g * 1024 + k * 32 + (((p / 8) ^ (k % 4)) * 8) + (p % 8);
    float got = dump[idx];
    if (got != A[k * M + m]) ++bad;
  }
  for (float x : dump) if (x == 0) ++zeros;
  printf("smem A layout mismatches: %d / %d   (zeros in dump: %d)\n", bad, M * K, zeros);
  printf("dump[0..7] = %g %g %g %g %g %g %g %g ; A[0][0..7] = %g %g %g %g %g %g %g %g\n", dump[0], dump[1], dump[2], dump[3], dump[4], dump[5], dump[6], dump[7], A[0], A[1], A[2], A[3], A[4], A[5], A[6], A[7]);
  return 0;
}
