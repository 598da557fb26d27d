// conv_tc: the convolution family of the refiner on the 5th-generation tensor cores, with fp32-grade
// results: scflow_conv_tc_fwd / scflow_conv_tc_pack_weights / scflow_split_bf16.
//
// Replaces every cuDNN convolution of the path: the RAFT encoders (models/encoder/raft_encoder.py:286-314,
// models/backbone/resnet.py:12-92), MotionEncoder (models/decoder/raft_decoder.py:152-166), the flow / mask
// heads (raft_decoder.py:256-294), the delta-flow / mask encoders (models/decoder/scflow_decoder.py:101-106)
// and the pose-head convolutions (models/head/pose_head.py:108-150).
//
// Arithmetic: SPLIT-BF16 ("bf16x3").  TF32 tensor-core convolutions put the 8-iteration refinement 3-6e-4 rad
// from the fp32 reference (budget 1e-4 rad); fp32 SIMT convolutions are 5x slower.  Here every operand v is
// carried as two bf16 numbers hi = rn(v), lo = rn(v - hi) (16 significand bits, |v - hi - lo| <= 2^-18 |v|),
// and a product is three tcgen05 kind::f16 MMAs into ONE fp32 TMEM accumulator:
//     a.b ~= a_lo.b_hi + a_hi.b_lo + a_hi.b_hi          (dropped: a_lo.b_lo <= 2^-18 |a.b|)
// i.e. ~40x more accurate than TF32 at 1.5x the tensor time of one TF32 MMA (bf16 runs at twice the TF32
// rate), and the same bytes as fp32 everywhere (2 + 2 per element) in HBM, L2 and shared memory.
//
// Layout: activations are channels-last split planes: pixel p holds [hi C | lo C] bf16 (the producer's
// epilogue writes both), weights are packed once to [tap][k-block][hi|lo][Cout_pad][BK] bf16.
//
// Kernel: persistent, warp-specialised implicit GEMM.  A tile is 128 output pixels (a TW x TH rectangle of
// one image, or whole small images) x up to 256 output channels.  The convolution is never unrolled in
// memory: for tap (ky,kx) the A operand of a k-block is ONE TMA box of the input at the tile's origin
// shifted by the tap -- out-of-bounds zero fill is the zero padding, elementStrides is the conv stride --
// landing in the K-major 128B-swizzled layout tcgen05 reads.  warp 0: TMA producer; warp 1: MMA issuer
// (one elected thread); warps 2..9: epilogue (tcgen05.ld -> bias / residual / ReLU -> split-bf16, fp32
// channels-last or planar fp32 stores, optional per-(image, channel) sum and sum of squares for
// InstanceNorm).  The fp32 accumulator is double-buffered in TMEM so the epilogue of tile i runs under the
// MMAs of tile i+1.
#include "common.cuh"
#include "tc_common.cuh"
#include <cuda_bf16.h>
#include <string.h>

namespace scflow {
namespace ctc {

constexpr int BM = 128;                        // output pixels per tile (TMEM lanes)
constexpr int NEPI = 8;                        // epilogue warps: two per TMEM lane quarter
constexpr int THREADS = 64 + 32 * NEPI;
constexpr int TMEM_COLS = 512;                 // two accumulator buffers of up to 256 columns
constexpr int SMEM_LIMIT = 227 * 1024;

struct Maps {
  CUtensorMap a_hi, a_lo, w;
  CUtensorMap b_hi, b_lo;                      // second input tensor (k-blocks kb_split ..): cat-free operands
};

struct Params {
  int n, OH, OW;
  int kh, kw, stride, pad_y, pad_x;
  int kblocks, ksteps_last;                    // k-blocks per tap; k-steps (16 channels) in the last one
  int kb_split;                                // k-blocks [kb_split, kblocks) read the second input (= kblocks: none)
  int TW, TH, NI, tiles_x, tiles_y;            // tile rectangle, images per tile, tiles per image
  int m_tiles, n_tiles, bn;                    // bn = output channels per tile (multiple of 16, <= 256)
  int cout_pad;                                // n_tiles * bn = packed weight rows per (tap, k-block, part)
  int cout_store;                              // channels written per pixel (multiple of 8, <= cout_pad)
  int stages, stage_bytes, a_bytes, b_bytes;   // ring geometry (a_bytes, b_bytes: ONE part)
  // epilogue
  const float* bias;                           // [cout_pad] or null
  int relu;
  const __nv_bfloat16* res_split;              // residual in split layout (hi plane), or null
  long long res_pitch; int res_lo;             // elements per pixel; offset of the lo plane
  const float* res_f32; long long resf_pitch;  // or fp32 channels-last residual
  __nv_bfloat16* out_split; long long out_pitch; int out_lo;
  float* out_f32; long long outf_pitch; int f32_tf32;   // fp32 channels-last, optionally RN to TF32
  float* out_planar; int planar_tf32;          // (N, cout_store, OH*OW) fp32, optionally RN to TF32
  double* stats;                               // [n][cout_store][2] sum, sum of squares (pre-activation)
  int fold;                                    // bn <= 128: A_hi x [B_hi;B_lo] as ONE N = 2 bn MMA (see the issuer)
  // ConvGRU epilogues (epi = SCFLOW_EPI_*): exact fp32 hidden state (n, OH, OW, gru_c), update gate z
  int epi, gru_c;
  float* gru_h;
  float* gru_z;
  // patch-resident kernel (conv_patch_kernel): input patch of a tile = PH x pitch pixel rows of 128 bytes per part
  int PH, pitch;                               // patch rows; patch row pitch in pixels (= box width)
  int patch_tx, patch_bytes;                   // bytes one part's TMA box writes; the same rounded up to 1024
  int a_stages, a_stage_bytes;                 // ring of patches: one entry = one 64-channel k-block, hi + lo
  int w_stages, w_stage_bytes, wb_bytes;       // ring of filter slabs: one entry = one (k-block, tap), hi + lo
  int w_resident;                              // the whole filter stays in shared memory (w_stages = all entries)
  int wide;                                    // epilogue pointers / pitches are 32-byte aligned: 256-bit accesses
  int terms;                                   // bf16 MMAs per product: 3 (fp32-grade, default), 2 (activations as bf16), 1 (bf16)
  const float* dot_w;                          // [cout_pad][SCFLOW_CONV_DOT_PITCH]: projections of the output pixel
  float* dot_out;                              // [pixels][dot_slots][dot_pitch]
  int dot_n, dot_pitch, dot_slots;
};

__device__ __forceinline__ uint32_t make_idesc_bf16(int m, int n) {
  // kind::f16: c_format F32 (1) at [4,6), a_format / b_format BF16 (1) at [7,10) / [10,13), K-major both
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ void mma_bf16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ void mma_bf16_ss_2cta(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                 uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// shared::cluster address of the object at this CTA-local shared address in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_rank(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}

// Arrive on a barrier of another CTA of the cluster without the release fence of mbar_arrive_remote (MEMBAR +
// ERRBAR per arrival, 15 % of the pair kernel's stall samples): the epilogue hands back TMEM, whose reads are
// complete (tcgen05.wait::ld) and ordered (tcgen05.fence::before_thread_sync) before this instruction issues.
__device__ __forceinline__ void mbar_arrive_remote_relaxed(uint32_t bar_addr) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}

__device__ __forceinline__ float rn_tf32(float v) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
  return __uint_as_float(u);
}

// sigmoid / tanh from one ex2.approx and one rcp.approx each (a few ulp of fp32: well below the 2^-17 of the
// split operands they feed)
__device__ __forceinline__ float sigmoidf_(float v) { return __fdividef(1.0f, 1.0f + __expf(-v)); }
__device__ __forceinline__ float tanhf_(float v) {
  const float e = __expf(-2.0f * fabsf(v));                 // in (0, 1]: no overflow
  return copysignf(__fdividef(1.0f - e, 1.0f + e), v);
}

__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

struct Tile {
  int n0, oy0, ox0, nt;
};
__device__ __forceinline__ Tile decode_tile(const Params& p, int t) {
  Tile r;
  r.nt = t % p.n_tiles;
  int m = t / p.n_tiles;
  const int per_img = p.tiles_x * p.tiles_y;
  const int g = m / per_img;
  m -= g * per_img;
  const int ty = m / p.tiles_x;
  r.n0 = g * p.NI;
  r.oy0 = ty * p.TH;
  r.ox0 = (m - ty * p.tiles_x) * p.TW;
  return r;
}

// 32 bytes per lane per instruction (LDG/STG.256, sm_100): a lane of the epilogue owns a PIXEL, so a warp-wide
// access touches 32 different 128-byte lines whatever its width, and the L1/LSU pipe (one line per cycle) bounds
// the heavier epilogues: the 16-byte form moved 16 bytes per line-cycle.  `wide`: pointers and pitches allow it.
__device__ __forceinline__ void ld8(const float* src, float* v, bool wide) {
  if (wide) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(src));
  } else {
    const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
}
// same, through the coherent path (the ConvGRU state is read and rewritten by the same launch)
__device__ __forceinline__ void ld8_rw(const float* src, float* v, bool wide) {
  if (wide) {
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(src) : "memory");
  } else {
    const float4 a = *reinterpret_cast<const float4*>(src), b = *(reinterpret_cast<const float4*>(src) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
}
__device__ __forceinline__ void st8(float* dst, const float* v, bool wide) {
  if (wide) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst), "f"(v[0]), "f"(v[1]), "f"(v[2]),
                 "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
  } else {
    reinterpret_cast<float4*>(dst)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(dst)[1] = make_float4(v[4], v[5], v[6], v[7]);
  }
}
// 16 values -> 16 bf16 hi at dst_hi and 16 bf16 lo at dst_lo (32 bytes each)
__device__ __forceinline__ void st_split16(__nv_bfloat16* dst_hi, __nv_bfloat16* dst_lo, const float* v, bool wide) {
  uint32_t h[8], l[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) split2(v[2 * e], v[2 * e + 1], h[e], l[e]);
  if (wide) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst_hi), "r"(h[0]), "r"(h[1]), "r"(h[2]),
                 "r"(h[3]), "r"(h[4]), "r"(h[5]), "r"(h[6]), "r"(h[7]) : "memory");
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst_lo), "r"(l[0]), "r"(l[1]), "r"(l[2]),
                 "r"(l[3]), "r"(l[4]), "r"(l[5]), "r"(l[6]), "r"(l[7]) : "memory");
  } else {
    reinterpret_cast<uint4*>(dst_hi)[0] = make_uint4(h[0], h[1], h[2], h[3]);
    reinterpret_cast<uint4*>(dst_hi)[1] = make_uint4(h[4], h[5], h[6], h[7]);
    reinterpret_cast<uint4*>(dst_lo)[0] = make_uint4(l[0], l[1], l[2], l[3]);
    reinterpret_cast<uint4*>(dst_lo)[1] = make_uint4(l[4], l[5], l[6], l[7]);
  }
}
// 16 bf16 hi + 16 bf16 lo added into 16 values
__device__ __forceinline__ void add_split16(const __nv_bfloat16* src_hi, const __nv_bfloat16* src_lo, float* v, bool wide) {
  uint32_t h[8], l[8];
  if (wide) {
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(h[0]), "=r"(h[1]), "=r"(h[2]), "=r"(h[3]), "=r"(h[4]), "=r"(h[5]), "=r"(h[6]), "=r"(h[7])
                 : "l"(src_hi));
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(l[0]), "=r"(l[1]), "=r"(l[2]), "=r"(l[3]), "=r"(l[4]), "=r"(l[5]), "=r"(l[6]), "=r"(l[7])
                 : "l"(src_lo));
  } else {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(src_hi)), b = __ldg(reinterpret_cast<const uint4*>(src_hi) + 1);
    const uint4 c = __ldg(reinterpret_cast<const uint4*>(src_lo)), d = __ldg(reinterpret_cast<const uint4*>(src_lo) + 1);
    h[0] = a.x; h[1] = a.y; h[2] = a.z; h[3] = a.w; h[4] = b.x; h[5] = b.y; h[6] = b.z; h[7] = b.w;
    l[0] = c.x; l[1] = c.y; l[2] = c.z; l[3] = c.w; l[4] = d.x; l[5] = d.y; l[6] = d.z; l[7] = d.w;
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    v[2 * e] += bf_lo(h[e]) + bf_lo(l[e]);
    v[2 * e + 1] += bf_hi(h[e]) + bf_hi(l[e]);
  }
}

constexpr int kEpiPlain = 0, kEpiGru = 1, kEpiDot = 2;   // epilogue a kernel instance is compiled with

// One accumulator tile (128 pixels x bn channels in TMEM at taddr) through the epilogue of this warp:
// chunks of 32 columns j = half, half + 2, ...; lane = pixel (n, oy, ox).  `live`: the pixel exists.
template <bool DOT>
__device__ __forceinline__ void epilogue_plain(const Params& p, const Tile& tl, uint32_t taddr, int half, int lane,
                                              int n, int oy, int ox, bool live, uint64_t* acc_full, uint32_t parity) {
  const int chunks = (p.bn + 31) >> 5;
  const long long plane = (long long)p.OH * p.OW;
  const long long pix = ((long long)n * p.OH + oy) * p.OW + ox;
  const bool wide = p.wide != 0;
  bool waited = false;
  float dacc[DOT ? SCFLOW_CONV_DOT_PITCH : 1];
#pragma unroll
  for (int k = 0; k < (DOT ? SCFLOW_CONV_DOT_PITCH : 1); ++k) dacc[k] = 0.f;
  for (int j = half; j < chunks; j += 2) {
    const int c0 = tl.nt * p.bn + j * 32;               // first output channel of the chunk
    if (!waited) {
      tc::mbar_wait(acc_full, parity);
      tc::tc_fence_after();
      waited = true;
    }
    float v[32];
    tc::tmem_ld_32x32(taddr + j * 32, v);
    if (p.fold) {
      float u[32];
      tc::tmem_ld_32x32(taddr + p.bn + j * 32, u);
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += u[i];
    } else {
      tc::tmem_ld_wait();
    }
    int nv = p.cout_store - c0;                          // channels of it that are stored
    nv = nv > 32 ? 32 : nv;
    if (tl.nt * p.bn + p.bn - c0 < nv) nv = tl.nt * p.bn + p.bn - c0;
    if (nv <= 0) continue;
    if (p.bias) {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < nv) v[i] += __ldg(p.bias + c0 + i);                      // bias is padded to cout_pad
    }
    if (p.stats) {
      // per-(image, channel) sum and sum of squares over the warp's 32 pixels (all of one image:
      // host guarantees OH*OW >= 32 with stats), butterfly so that lane i ends with channel i
      float s[32], q[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) { s[i] = live ? v[i] : 0.f; q[i] = s[i] * s[i]; }
#pragma unroll
      for (int w = 16; w >= 1; w >>= 1) {
#pragma unroll
        for (int i = 0; i < w; ++i) {
          const bool up = (lane & w) != 0;
          const float ss = up ? s[i] : s[i + w], sq = up ? q[i] : q[i + w];
          const float rs = __shfl_xor_sync(0xffffffffu, ss, w), rq = __shfl_xor_sync(0xffffffffu, sq, w);
          s[i] = (up ? s[i + w] : s[i]) + rs;
          q[i] = (up ? q[i + w] : q[i]) + rq;
        }
      }
      // lane i now holds channel bitrev-free index: after the halving steps element 0 of lane l is channel l
      const int nw = __shfl_sync(0xffffffffu, n, 0);
      if (lane < nv && nw < p.n) {
        // SCFLOW_STATS_REPLICAS accumulators per (image, channel): the CTAs of a wave work on one or two images, and
        // 512 same-address fp64 atomics per tile serialise in L2 (E1 at 256x256: 476 -> 37x us); the replica is
        // picked by the warp's TMEM quarter and the tile's parity, norm_apply adds them up
        const int rep = ((threadIdx.x >> 5) & 3) + 4 * (((tl.ox0 >> 3) ^ (tl.oy0 >> 4)) & 1);
        double* dst = p.stats + (((long long)nw * SCFLOW_STATS_REPLICAS + rep) * p.cout_store + c0 + lane) * 2;
        atomicAdd(dst, (double)s[0]);
        atomicAdd(dst + 1, (double)q[0]);
      }
    }
    if (!live) continue;
    if (p.res_split) {
      const __nv_bfloat16* rh = p.res_split + pix * p.res_pitch + c0;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        if (16 * k + 8 < nv) add_split16(rh + 16 * k, rh + p.res_lo + 16 * k, v + 16 * k, wide);
        else if (16 * k < nv) {                      // a last group of 8 channels
          const uint4 a = __ldg(reinterpret_cast<const uint4*>(rh + 16 * k));
          const uint4 b = __ldg(reinterpret_cast<const uint4*>(rh + p.res_lo + 16 * k));
          const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            v[16 * k + 2 * e] += bf_lo(aw[e]) + bf_lo(bw[e]);
            v[16 * k + 2 * e + 1] += bf_hi(aw[e]) + bf_hi(bw[e]);
          }
        }
      }
    } else if (p.res_f32) {
      const float* rp = p.res_f32 + pix * p.resf_pitch + c0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (8 * k < nv) {
          float r[8];
          ld8(rp + 8 * k, r, wide);
#pragma unroll
          for (int e = 0; e < 8; ++e) v[8 * k + e] += r[e];
        }
      }
    }
    if (p.relu) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
    }
    if (DOT) {
      // projections of the finished pixel onto dot_n vectors (the 1-2 channel predictor behind a head's last
      // 3x3 layer, raft_decoder.py:256-294): every lane reads the same weights (one L1 wavefront per load);
      // channels past cout carry zero weights and zero activations.  (fma.rn.f32x2 was measured: slower.)
      const int n4 = (p.dot_n + 3) >> 2;
      const float4* wr = reinterpret_cast<const float4*>(p.dot_w) + (size_t)c0 * (SCFLOW_CONV_DOT_PITCH / 4);
#pragma unroll
      for (int i = 0; i < 32; ++i) {
#pragma unroll
        for (int k4 = 0; k4 < SCFLOW_CONV_DOT_PITCH / 4; ++k4) {
          if (k4 < n4) {
            const float4 w4 = __ldg(wr + i * (SCFLOW_CONV_DOT_PITCH / 4) + k4);
            dacc[4 * k4] = fmaf(v[i], w4.x, dacc[4 * k4]); dacc[4 * k4 + 1] = fmaf(v[i], w4.y, dacc[4 * k4 + 1]);
            dacc[4 * k4 + 2] = fmaf(v[i], w4.z, dacc[4 * k4 + 2]); dacc[4 * k4 + 3] = fmaf(v[i], w4.w, dacc[4 * k4 + 3]);
          }
        }
      }
    }
    if (p.out_split) {
      __nv_bfloat16* oh = p.out_split + pix * p.out_pitch + c0;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        if (16 * k + 8 < nv) st_split16(oh + 16 * k, oh + p.out_lo + 16 * k, v + 16 * k, wide);
        else if (16 * k < nv) {                      // a last group of 8 channels: 16-byte stores
          uint4 h, l;
          split2(v[16 * k], v[16 * k + 1], h.x, l.x); split2(v[16 * k + 2], v[16 * k + 3], h.y, l.y);
          split2(v[16 * k + 4], v[16 * k + 5], h.z, l.z); split2(v[16 * k + 6], v[16 * k + 7], h.w, l.w);
          *reinterpret_cast<uint4*>(oh + 16 * k) = h;
          *reinterpret_cast<uint4*>(oh + p.out_lo + 16 * k) = l;
        }
      }
    }
    if (p.out_f32) {
      float* op = p.out_f32 + pix * p.outf_pitch + c0;
      if (p.f32_tf32) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = rn_tf32(v[i]);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (8 * k < nv) st8(op + 8 * k, v + 8 * k, wide);
    }
    if (p.out_planar) {
      float* op = p.out_planar + ((long long)n * p.cout_store + c0) * plane + (long long)oy * p.OW + ox;
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < nv) op[i * plane] = p.planar_tf32 ? rn_tf32(v[i]) : v[i];
    }
  }
  if (!waited) {                                       // a warp without a chunk of this tile still hands it back
    tc::mbar_wait(acc_full, parity);
    tc::tc_fence_after();
  }
  if (DOT && live) {
    // one slot per (output-channel tile, warp half): plain stores, summed in a fixed order by tap_sum_kernel --
    // no initialisation, and the same bits on every run (atomics would add the partial sums in arrival order)
    float4* dp = reinterpret_cast<float4*>(p.dot_out + (pix * p.dot_slots + tl.nt * 2 + half) * p.dot_pitch);
#pragma unroll
    for (int k4 = 0; k4 < SCFLOW_CONV_DOT_PITCH / 4; ++k4)
      if (4 * k4 < p.dot_pitch) dp[k4] = make_float4(dacc[4 * k4], dacc[4 * k4 + 1], dacc[4 * k4 + 2], dacc[4 * k4 + 3]);
  }
}

// The ConvGRU epilogues (SCFLOW_EPI_GRU_ZR / _GATE) of one accumulator tile.  The wait for the accumulator
// happens after the chunk's operands (update gate, state, x part of q: up to 384 bytes per lane) have been
// requested: they do not depend on the accumulator, and with eight epilogue warps per SM their latency was
// otherwise exposed once per chunk.  (With a folded accumulator the registers are needed for its second half
// first, and the operands are requested after it.)
__device__ __forceinline__ void epilogue_gru(const Params& p, const Tile& tl, uint32_t taddr, int half, int lane,
                                             int n, int oy, int ox, bool live, uint64_t* acc_full, uint32_t parity) {
  const int chunks = (p.bn + 31) >> 5;
  const long long pix = ((long long)n * p.OH + oy) * p.OW + ox;
  const bool wide = p.wide != 0, gate = p.epi == SCFLOW_EPI_GRU_GATE, fold = p.fold != 0;
  bool waited = false;
  for (int j = half; j < chunks; j += 2) {
    const int c0 = tl.nt * p.bn + j * 32;               // first output channel of the chunk
    const bool zhalf = !gate && c0 < p.gru_c;           // GRU_ZR: update-gate half | reset-gate half
    float ga[32], gb[32];
    auto load_operands = [&]() {
      if (gate) {
        const float* zp = p.gru_z + pix * p.gru_c + c0;
        const float* hp = p.gru_h + pix * p.gru_c + c0;
#pragma unroll
        for (int k = 0; k < 4; ++k) { ld8(zp + 8 * k, ga + 8 * k, wide); ld8_rw(hp + 8 * k, gb + 8 * k, wide); }
      } else {
        if (p.res_f32) {                             // iteration-invariant part of the pre-activation (see below)
          const float* rp = p.res_f32 + pix * p.resf_pitch + c0;
#pragma unroll
          for (int k = 0; k < 4; ++k) ld8(rp + 8 * k, ga + 8 * k, wide);
        }
        if (!zhalf) {
          const float* hp = p.gru_h + pix * p.gru_c + (c0 - p.gru_c);
#pragma unroll
          for (int k = 0; k < 4; ++k) ld8(hp + 8 * k, gb + 8 * k, wide);
        }
      }
    };
    if (live && !fold) load_operands();
    if (!waited) {
      tc::mbar_wait(acc_full, parity);
      tc::tc_fence_after();
      waited = true;
    }
    float v[32];
    tc::tmem_ld_32x32(taddr + j * 32, v);
    if (fold) {
      float u[32];
      tc::tmem_ld_32x32(taddr + p.bn + j * 32, u);
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += u[i];
    } else {
      tc::tmem_ld_wait();
    }
    if (!live) continue;
    if (fold) load_operands();
    if (p.bias) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += __ldg(p.bias + c0 + i);       // bias is padded to cout_pad
    }
    if (!gate) {
      // raft_decoder.py:247-250, first half: channels [0,C) z = sigmoid(.) -> gru_z (fp32);
      // channels [C,2C) r = sigmoid(.), r.h -> split planes (the operand of the q convolution).
      // res_f32 (2C channels): a part of the pre-activation computed elsewhere -- the context features are the same
      // in all iterations, their share of the convolution (and the bias) is computed once per forward
      if (p.res_f32) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += ga[i];
      }
      if (zhalf) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = sigmoidf_(v[i]);
        float* zp = p.gru_z + pix * p.gru_c + c0;
#pragma unroll
        for (int k = 0; k < 4; ++k) st8(zp + 8 * k, v + 8 * k, wide);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = sigmoidf_(v[i]) * gb[i];
        __nv_bfloat16* oh = p.out_split + pix * p.out_pitch + (c0 - p.gru_c);
        st_split16(oh, oh + p.out_lo, v, wide);
        st_split16(oh + 16, oh + p.out_lo + 16, v + 16, wide);
      }
    } else {
      // raft_decoder.py:251-252: q = tanh(conv(r.h) + [x part of the q convolution, res_f32] + b);
      // h = (1 - z) h + z q, in place in the exact fp32 state, and as split planes for the next operand
      float* hp = p.gru_h + pix * p.gru_c + c0;
      if (p.res_f32) {
        const float* qp = p.res_f32 + pix * p.resf_pitch + c0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float qx[8];
          ld8(qp + 8 * k, qx, wide);
#pragma unroll
          for (int e = 0; e < 8; ++e) v[8 * k + e] += qx[e];
        }
      }
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = (1.0f - ga[i]) * gb[i] + ga[i] * tanhf_(v[i]);
#pragma unroll
      for (int k = 0; k < 4; ++k) st8(hp + 8 * k, v + 8 * k, wide);
      __nv_bfloat16* oh = p.out_split + pix * p.out_pitch + c0;
      st_split16(oh, oh + p.out_lo, v, wide);
      st_split16(oh + 16, oh + p.out_lo + 16, v + 16, wide);
    }
  }
  if (!waited) {
    tc::mbar_wait(acc_full, parity);
    tc::tc_fence_after();
  }
}

template <int BK, int EPI>    // channels per k-block: 64 (128-byte swizzle) or 32 (64-byte swizzle); ConvGRU epilogues
__global__ void __launch_bounds__(THREADS, 1)
conv_tc_kernel(const __grid_constant__ Maps maps, const __grid_constant__ Params p) {
  constexpr uint32_t kLayout = BK == 64 ? 2u : 4u;          // SWIZZLE_128B : SWIZZLE_64B
  constexpr uint32_t kSbo = BK == 64 ? 1024u : 512u;        // 8 rows of the swizzle atom
  constexpr int MAX_STAGES = 8;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * p.stage_bytes);
  uint64_t* full = bars;                         // [stages]
  uint64_t* empty = bars + MAX_STAGES;           // [stages]
  uint64_t* acc_full = bars + 2 * MAX_STAGES;    // [2]
  uint64_t* acc_empty = acc_full + 2;            // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = p.m_tiles * p.n_tiles;
  const int taps = p.kh * p.kw;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&maps.a_hi); tc::tma_prefetch_desc(&maps.a_lo); tc::tma_prefetch_desc(&maps.w);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < p.stages; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
      for (int i = 0; i < 2; ++i) { tc::mbar_init(&acc_full[i], 1); tc::mbar_init(&acc_empty[i], NEPI); }
      tc::fence_barrier_init();
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x) {
        const Tile tl = decode_tile(p, t);
        const int wrow0 = tl.nt * p.bn;
        for (int tap = 0; tap < taps; ++tap) {
          const int ky = tap / p.kw, kx = tap - ky * p.kw;
          const int ix = tl.ox0 * p.stride + kx - p.pad_x, iy = tl.oy0 * p.stride + ky - p.pad_y;
          for (int kb = 0; kb < p.kblocks; ++kb) {
            tc::mbar_wait(&empty[stage], phase ^ 1);
            uint8_t* st = smem + (size_t)stage * p.stage_bytes;
            tc::mbar_expect_tx(&full[stage], (uint32_t)p.stage_bytes);
            const bool second = kb >= p.kb_split;
            const int ch = (second ? kb - p.kb_split : kb) * BK;
            tc::tma_load_4d(second ? &maps.b_hi : &maps.a_hi, &full[stage], st, ch, ix, iy, tl.n0);
            tc::tma_load_4d(second ? &maps.b_lo : &maps.a_lo, &full[stage], st + p.a_bytes, ch, ix, iy, tl.n0);
            const int wrow = ((tap * p.kblocks + kb) * 2) * p.cout_pad + wrow0;
            tc::tma_load_2d(&maps.w, &full[stage], st + 2 * p.a_bytes, 0, wrow);
            tc::tma_load_2d(&maps.w, &full[stage], st + 2 * p.a_bytes + p.b_bytes, 0, wrow + p.cout_pad);
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // ONE ELECTED lane (elect.sync, so the compiler knows exactly one thread runs the region and keeps
    // descriptors and addresses in uniform registers) walks the loops, issues the MMAs and commits.  Written as
    // `if (lane == 0)` the same loop compiled to ~55 SASS instructions per MMA (R2UR.BROADCAST + ELECT + BRA.U.ANY
    // sequences around every tcgen05.mma) and the issuing thread bound the narrow tiles (tensor pipe 24-33 %).
    if (tc::elect_one()) {
      const uint32_t idesc = make_idesc_bf16(BM, p.bn);
      const uint32_t idesc2 = make_idesc_bf16(BM, 2 * p.bn);
      const uint32_t smem0 = tc::smem_u32(smem);
      const bool fold = p.fold != 0;
      const int terms = p.terms;
      int stage = 0; uint32_t phase = 0; uint32_t it = 0;
      for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
        const uint32_t buf = it & 1;
        tc::mbar_wait(&acc_empty[buf], ((it >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        const uint32_t d = tmem_base + buf * 256u;
        uint32_t acc = 0;
        for (int tap = 0; tap < taps; ++tap) {
          for (int kb = 0; kb < p.kblocks; ++kb) {
            tc::mbar_wait(&full[stage], phase);
            tc::tc_fence_after();
            const uint32_t a_hi = smem0 + stage * p.stage_bytes;
            const uint32_t a_lo = a_hi + p.a_bytes;
            const uint32_t b_hi = a_lo + p.a_bytes;
            const uint32_t b_lo = b_hi + p.b_bytes;
            const int nks = kb == p.kblocks - 1 ? p.ksteps_last : BK / 16;
            // K-major swizzled: 8-row groups kSbo bytes apart; a k-step (16 bf16) is +32 bytes (+2 in the
            // descriptor's address field)
            const uint64_t dah = tc::make_smem_desc(a_hi, 16, kSbo, kLayout);
            const uint64_t dal = tc::make_smem_desc(a_lo, 16, kSbo, kLayout);
            const uint64_t dbh = tc::make_smem_desc(b_hi, 16, kSbo, kLayout);
            const uint64_t dbl = tc::make_smem_desc(b_lo, 16, kSbo, kLayout);
#pragma unroll
            for (int ks = 0; ks < BK / 16; ++ks) {
              if (ks < nks) {
                if (fold) {
                  // narrow tiles fold two of the three products into one instruction: [B_hi;B_lo] are
                  // adjacent in the stage = one 2 bn-row operand; columns [0,bn) get a_hi.b_hi, [bn,2bn)
                  // a_hi.b_lo, the epilogue adds the halves
                  mma_bf16_ss(d, dah + 2 * ks, dbh + 2 * ks, idesc2, acc | (uint32_t)ks);
                  if (terms == 3) mma_bf16_ss(d, dal + 2 * ks, dbh + 2 * ks, idesc, 1u);
                } else if (terms == 3) {
                  mma_bf16_ss(d, dal + 2 * ks, dbh + 2 * ks, idesc, acc | (uint32_t)ks);   // small terms first
                  mma_bf16_ss(d, dah + 2 * ks, dbl + 2 * ks, idesc, 1u);
                  mma_bf16_ss(d, dah + 2 * ks, dbh + 2 * ks, idesc, 1u);
                } else if (terms == 2) {                                                   // opt-in reduced modes
                  mma_bf16_ss(d, dah + 2 * ks, dbl + 2 * ks, idesc, acc | (uint32_t)ks);
                  mma_bf16_ss(d, dah + 2 * ks, dbh + 2 * ks, idesc, 1u);
                } else {
                  mma_bf16_ss(d, dah + 2 * ks, dbh + 2 * ks, idesc, acc | (uint32_t)ks);
                }
              }
            }
            tc::tc_commit(&empty[stage]);
            acc = 1u;
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
          }
        }
        tc::tc_commit(&acc_full[buf]);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ epilogue (8 warps)
    const int quarter = warp & 3;                           // TMEM lanes this warp may read
    const int half = (warp - 2) >> 2;                       // which 32-column chunks it takes
    const int m = quarter * 32 + lane;                      // TMEM lane = pixel of the tile
    const int per_tile_img = p.TW * p.TH;
    const int mi = m / per_tile_img, mr = m - mi * per_tile_img;
    const int my = mr / p.TW, mx = mr - my * p.TW;
    uint32_t it = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
      const Tile tl = decode_tile(p, t);
      const uint32_t buf = it & 1;
      const int n = tl.n0 + mi, oy = tl.oy0 + my, ox = tl.ox0 + mx;
      const bool live = n < p.n;
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + buf * 256u;
      if (EPI == kEpiGru) epilogue_gru(p, tl, taddr, half, lane, n, oy, ox, live, &acc_full[buf], (it >> 1) & 1);
      else epilogue_plain<EPI == kEpiDot>(p, tl, taddr, half, lane, n, oy, ox, live, &acc_full[buf], (it >> 1) & 1);
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&acc_empty[buf]);
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc::tc_fence_after();
    tc::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ================================================================================================
// Second generation for stride-1 filters: PATCH-RESIDENT operands.
//
// The kernel above pulls one shifted TMA box per (tap, k-block) and the whole filter slab per tile: a 3x3
// layer reads its input nine times and its filter once per 128 pixels through L2, and every layer shape
// of the refiner sat at the L2 -> SM throughput cap (10-14 TB/s of TMA traffic, tensor pipe 30-65 % busy;
// profiles/r02_conv_ncu.txt).  Here
//   * a tile is 16 rows x 8 pixels, and the (16 + kh - 1) x (8 + kw - 1) input PATCH of a 64-channel k-block
//     is loaded ONCE (one TMA box per part; out-of-bounds zero fill is the padding).  A pixel is one 128-byte
//     swizzled row, an output row of the tile is exactly one 8-row swizzle atom, so the A operand of tap
//     (ky, kx) is the same patch read through a descriptor whose start is advanced by (ky pitch + kx) rows
//     and whose stride between atoms is the patch pitch: no data moves between taps.  (Measured on B200: the
//     128-byte swizzle XOR comes from ABSOLUTE shared-memory address bits [7,10), so a start advanced by whole
//     rows and an atom stride that is not a multiple of 1024 bytes read exactly the rows TMA wrote, with the
//     descriptor's base-offset field left 0; setting it to (start >> 7) & 7 gives wrong products.);
//   * PAIR: a cluster of two CTAs computes two tiles as one M = 256 tcgen05.mma.cta_group::2, each CTA
//     loading only HALF of the filter rows -- filter traffic per pixel halves;
//   * small filters (<= 32 (k-block, tap) slabs that fit next to two patches) stay resident in shared
//     memory for the whole launch.
// Barrier waits are plain CTA-scope try_waits also in PAIR mode: what they guard is written by the async
// proxy (TMA, tcgen05.commit) and read by the async proxy; the cluster-scope acquire form costs an L1
// invalidate (CCTL.IVALL) per wait.
// warp 0: patch producer, warp 1: MMA issuer (one elected lane; PAIR: rank 0's), warps 2..9: epilogue,
// warp 10: filter producer.
constexpr int P_THREADS = THREADS + 32;
constexpr int MAX_A_STAGES = 4, MAX_W_STAGES = 32;
constexpr int P_BAR_BYTES = 1024;

template <bool PAIR, int EPI>
__global__ void __launch_bounds__(P_THREADS, 1)
conv_patch_kernel(const __grid_constant__ Maps maps, const __grid_constant__ Params p) {
  constexpr uint32_t kLayout = 2u;                          // SWIZZLE_128B
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* a_ring = smem;
  uint8_t* w_ring = smem + (size_t)p.a_stages * p.a_stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(w_ring + (size_t)p.w_stages * p.w_stage_bytes);
  uint64_t* a_full = bars;                                  // PAIR: rank 0's collect both CTAs' bytes
  uint64_t* a_empty = a_full + MAX_A_STAGES;
  uint64_t* w_full = a_empty + MAX_A_STAGES;
  uint64_t* w_empty = w_full + MAX_W_STAGES;
  uint64_t* acc_full = w_empty + MAX_W_STAGES;              // [2]
  uint64_t* acc_empty = acc_full + 2;                       // [2] PAIR: rank 0's, armed by both CTAs' epilogues
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = PAIR ? tc::cluster_ctarank() : 0u;
  const int unit0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int units_grid = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int m_units = PAIR ? (p.m_tiles + 1) >> 1 : p.m_tiles;
  const int total = m_units * p.n_tiles;
  const int taps = p.kh * p.kw;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&maps.a_hi); tc::tma_prefetch_desc(&maps.a_lo); tc::tma_prefetch_desc(&maps.w);
    if (p.kb_split < p.kblocks) { tc::tma_prefetch_desc(&maps.b_hi); tc::tma_prefetch_desc(&maps.b_lo); }
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < p.a_stages; ++i) { tc::mbar_init(&a_full[i], 1); tc::mbar_init(&a_empty[i], 1); }
      for (int i = 0; i < p.w_stages; ++i) { tc::mbar_init(&w_full[i], 1); tc::mbar_init(&w_empty[i], 1); }
      for (int i = 0; i < 2; ++i) {
        tc::mbar_init(&acc_full[i], 1);
        tc::mbar_init(&acc_empty[i], PAIR ? 2 * NEPI : NEPI);
      }
      tc::fence_barrier_init();
    }
    __syncwarp();
    if (PAIR) tc::tmem_alloc_2cta(tmem_slot, TMEM_COLS); else tc::tmem_alloc(tmem_slot, TMEM_COLS);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (PAIR) tc::cluster_sync();                  // the peer's barriers exist before anything signals them
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // this CTA's tile of unit u (PAIR: tile 2 m + rank of pair m; an odd tile count leaves the last one dead)
  auto tile_of = [&](int u, bool* alive) -> Tile {
    const int mu = u / p.n_tiles, nt = u - mu * p.n_tiles;
    int mt = PAIR ? 2 * mu + (int)rank : mu;
    *alive = mt < p.m_tiles;
    if (!*alive) mt = p.m_tiles - 1;
    return decode_tile(p, mt * p.n_tiles + nt);
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ patch producer
    if (lane == 0) {
      int as = 0; uint32_t aph = 0;
      for (int u = unit0; u < total; u += units_grid) {
        bool alive;
        const Tile tl = tile_of(u, &alive);
        const int ix = tl.ox0 - p.pad_x, iy = tl.oy0 - p.pad_y;
        for (int kb = 0; kb < p.kblocks; ++kb) {
          uint8_t* st = a_ring + (size_t)as * p.a_stage_bytes;
          const bool second = kb >= p.kb_split;
          const int ch = (second ? kb - p.kb_split : kb) * 64;
          const CUtensorMap* mh = second ? &maps.b_hi : &maps.a_hi;
          const CUtensorMap* ml = second ? &maps.b_lo : &maps.a_lo;
          if (PAIR) {
            tc::mbar_wait(&a_empty[as], aph ^ 1);
            const uint32_t fb = map_to_rank(tc::smem_u32(&a_full[as]), 0);
            if (rank == 0) tc::mbar_expect_tx(&a_full[as], 4u * (uint32_t)p.patch_tx);
            tc::tma_load_4d_2cta(mh, fb, st, ch, ix, iy, tl.n0);
            tc::tma_load_4d_2cta(ml, fb, st + p.patch_bytes, ch, ix, iy, tl.n0);
          } else {
            tc::mbar_wait(&a_empty[as], aph ^ 1);
            tc::mbar_expect_tx(&a_full[as], 2u * (uint32_t)p.patch_tx);
            tc::tma_load_4d(mh, &a_full[as], st, ch, ix, iy, tl.n0);
            tc::tma_load_4d(ml, &a_full[as], st + p.patch_bytes, ch, ix, iy, tl.n0);
          }
          if (++as == p.a_stages) { as = 0; aph ^= 1; }
        }
      }
    }
  } else if (warp == 10) {
    // ------------------------------------------------------------------ filter producer
    if (lane == 0) {
      int ws = 0; uint32_t wph = 0;
      const int rows = PAIR ? p.bn >> 1 : p.bn;               // filter rows this CTA holds per part
      for (int u = unit0; u < total; u += units_grid) {
        const int nt = u % p.n_tiles;
        const int wrow0 = nt * p.bn + (int)rank * rows;
        for (int kb = 0; kb < p.kblocks; ++kb) {
          for (int tap = 0; tap < taps; ++tap) {
            uint8_t* st = w_ring + (size_t)ws * p.w_stage_bytes;
            const int wrow = ((tap * p.kblocks + kb) * 2) * p.cout_pad + wrow0;
            if (PAIR) {
              if (!p.w_resident) tc::mbar_wait(&w_empty[ws], wph ^ 1);
              const uint32_t fb = map_to_rank(tc::smem_u32(&w_full[ws]), 0);
              if (rank == 0) tc::mbar_expect_tx(&w_full[ws], 4u * (uint32_t)p.wb_bytes);
              tc::tma_load_2d_2cta(&maps.w, fb, st, 0, wrow);
              tc::tma_load_2d_2cta(&maps.w, fb, st + p.wb_bytes, 0, wrow + p.cout_pad);
            } else {
              if (!p.w_resident) tc::mbar_wait(&w_empty[ws], wph ^ 1);
              tc::mbar_expect_tx(&w_full[ws], 2u * (uint32_t)p.wb_bytes);
              tc::tma_load_2d(&maps.w, &w_full[ws], st, 0, wrow);
              tc::tma_load_2d(&maps.w, &w_full[ws], st + p.wb_bytes, 0, wrow + p.cout_pad);
            }
            if (++ws == p.w_stages) { ws = 0; wph ^= 1; }
          }
        }
        if (p.w_resident) break;                              // every slab is in place for the whole launch
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one elected lane; see above)
    if (rank == 0 && tc::elect_one()) {
      const uint32_t idesc = make_idesc_bf16(PAIR ? 2 * BM : BM, p.bn);
      const uint32_t idesc2 = make_idesc_bf16(BM, 2 * p.bn);
      const uint32_t a0 = tc::smem_u32(a_ring), w0 = tc::smem_u32(w_ring);
      const uint32_t sbo = (uint32_t)p.pitch * 128u;          // one output row of the tile = one swizzle atom
      const bool fold = p.fold != 0, resident = p.w_resident != 0;
      const int terms = p.terms;
      const uint64_t lo_off = (uint64_t)(p.wb_bytes >> 4);    // B_lo behind B_hi, in descriptor address units
      int as = 0, ws = 0; uint32_t aph = 0, wph = 0, it = 0;
      for (int u = unit0; u < total; u += units_grid, ++it) {
        const uint32_t buf = it & 1;
        tc::mbar_wait(&acc_empty[buf], ((it >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        const uint32_t d = tmem_base + buf * 256u;
        uint32_t acc = 0;
        // a resident filter is in place after the first tile: nothing to wait for inside a k-block then (the
        // wait + fence per tap is most of the issue cost of the 64-filter encoder layers, whose 12 MMAs per tap
        // are 32 cycles each)
        const bool wait_w = !(resident && it > 0);
        for (int kb = 0; kb < p.kblocks; ++kb) {
          tc::mbar_wait(&a_full[as], aph);
          tc::tc_fence_after();
          const uint32_t a_hi = a0 + as * p.a_stage_bytes;
          const uint64_t dah0 = tc::make_smem_desc(a_hi, 16, sbo, kLayout);
          const uint64_t dal0 = tc::make_smem_desc(a_hi + p.patch_bytes, 16, sbo, kLayout);
          const int nks = kb == p.kblocks - 1 ? p.ksteps_last : 4;
          int ky = 0, kx = 0;
          for (int tap = 0; tap < taps; ++tap) {
            if (wait_w) {
              tc::mbar_wait(&w_full[ws], wph);                // a resident slab: parity 0 of a barrier that completes once
              tc::tc_fence_after();
            }
            const uint64_t sh = (uint64_t)((ky * p.pitch + kx) * 8);            // rows of 128 bytes, >> 4
            const uint64_t dah = dah0 + sh, dal = dal0 + sh;
            const uint64_t dbh = tc::make_smem_desc(w0 + ws * p.w_stage_bytes, 16, 1024, kLayout), dbl = dbh + lo_off;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              if (ks < nks) {
                if (PAIR) {
                  if (terms == 3) {
                    mma_bf16_ss_2cta(d, dal + 2 * ks, dbh + 2 * ks, idesc, acc | (uint32_t)ks);
                    mma_bf16_ss_2cta(d, dah + 2 * ks, dbl + 2 * ks, idesc, 1u);
                    mma_bf16_ss_2cta(d, dah + 2 * ks, dbh + 2 * ks, idesc, 1u);
                  } else if (terms == 2) {                                                 // opt-in reduced modes
                    mma_bf16_ss_2cta(d, dah + 2 * ks, dbl + 2 * ks, idesc, acc | (uint32_t)ks);
                    mma_bf16_ss_2cta(d, dah + 2 * ks, dbh + 2 * ks, idesc, 1u);
                  } else {
                    mma_bf16_ss_2cta(d, dah + 2 * ks, dbh + 2 * ks, idesc, acc | (uint32_t)ks);
                  }
                } else if (fold) {
                  mma_bf16_ss(d, dah + 2 * ks, dbh + 2 * ks, idesc2, acc | (uint32_t)ks);   // [B_hi;B_lo]: 2 bn rows
                  if (terms == 3) mma_bf16_ss(d, dal + 2 * ks, dbh + 2 * ks, idesc, 1u);
                } else if (terms == 3) {
                  mma_bf16_ss(d, dal + 2 * ks, dbh + 2 * ks, idesc, acc | (uint32_t)ks);
                  mma_bf16_ss(d, dah + 2 * ks, dbl + 2 * ks, idesc, 1u);
                  mma_bf16_ss(d, dah + 2 * ks, dbh + 2 * ks, idesc, 1u);
                } else if (terms == 2) {
                  mma_bf16_ss(d, dah + 2 * ks, dbl + 2 * ks, idesc, acc | (uint32_t)ks);
                  mma_bf16_ss(d, dah + 2 * ks, dbh + 2 * ks, idesc, 1u);
                } else {
                  mma_bf16_ss(d, dah + 2 * ks, dbh + 2 * ks, idesc, acc | (uint32_t)ks);
                }
              }
            }
            if (!resident) {
              if (PAIR) tc::tc_commit_2cta(&w_empty[ws]); else tc::tc_commit(&w_empty[ws]);
            }
            acc = 1u;
            if (++kx == p.kw) { kx = 0; ++ky; }
            if (++ws == p.w_stages) { ws = 0; if (!resident) wph ^= 1; }
          }
          if (PAIR) tc::tc_commit_2cta(&a_empty[as]); else tc::tc_commit(&a_empty[as]);
          if (++as == p.a_stages) { as = 0; aph ^= 1; }
        }
        if (PAIR) tc::tc_commit_2cta(&acc_full[buf]); else tc::tc_commit(&acc_full[buf]);
      }
    }
    __syncwarp();
  } else if (warp >= 2 && warp < 2 + NEPI) {
    // ------------------------------------------------------------------ epilogue (8 warps, own TMEM)
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;
    const int m = quarter * 32 + lane;
    const int my = m >> 3, mx = m & 7;                       // TW = 8
    const uint32_t empty0 = PAIR ? map_to_rank(tc::smem_u32(&acc_empty[0]), 0) : 0u;
    uint32_t it = 0;
    for (int u = unit0; u < total; u += units_grid, ++it) {
      bool alive;
      const Tile tl = tile_of(u, &alive);
      const uint32_t buf = it & 1;
      const int n = alive ? tl.n0 : p.n, oy = tl.oy0 + my, ox = tl.ox0 + mx;
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + buf * 256u;
      if (EPI == kEpiGru) epilogue_gru(p, tl, taddr, half, lane, n, oy, ox, alive, &acc_full[buf], (it >> 1) & 1);
      else epilogue_plain<EPI == kEpiDot>(p, tl, taddr, half, lane, n, oy, ox, alive, &acc_full[buf], (it >> 1) & 1);
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (PAIR) mbar_arrive_remote_relaxed(empty0 + buf * 8u); else tc::mbar_arrive(&acc_empty[buf]);
      }
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (PAIR) tc::cluster_sync();                  // nobody leaves while the pair's MMAs may touch its memory
  if (warp == 1) {
    tc::tc_fence_after();
    if (PAIR) tc::tmem_dealloc_2cta(tmem_base, TMEM_COLS); else tc::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// fp32 (Cout,Cin,KH,KW) [* scale[co]] -> packed bf16 rows [tap][kb][part][cout_pad][bk]
__global__ void pack_weights_kernel(const float* __restrict__ w, const float* __restrict__ scale,
                                    __nv_bfloat16* __restrict__ out, int cout, int cin, int taps, int kblocks,
                                    int cout_pad, int bk, long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int kk = (int)(i % bk);
    long long r = i / bk;
    const int co = (int)(r % cout_pad); r /= cout_pad;
    const int part = (int)(r & 1); r >>= 1;
    const int kb = (int)(r % kblocks);
    const int tap = (int)(r / kblocks);
    const int ci = kb * bk + kk;
    float v = 0.f;
    if (co < cout && ci < cin) {
      v = w[((long long)co * cin + ci) * taps + tap];
      if (scale) v *= scale[co];
    }
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    out[i] = part == 0 ? hi : __float2bfloat16_rn(v - __bfloat162float(hi));
  }
}

// fp32 channels-last (pixel pitch in_pitch) -> split planes; 8 channels per thread
__global__ void split_bf16_kernel(const float* __restrict__ in, long long in_pitch, __nv_bfloat16* __restrict__ out,
                                  long long out_pitch, int out_lo, int c8, int c_valid, long long pixels) {
  const long long total = pixels * c8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long pix = i / c8;
    const int c = (int)(i - pix * c8) * 8;
    float v[8];
    if (c + 8 <= c_valid) {
      const float4 a = *reinterpret_cast<const float4*>(in + pix * in_pitch + c);
      const float4 b = *reinterpret_cast<const float4*>(in + pix * in_pitch + c + 4);
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = c + e < c_valid ? in[pix * in_pitch + c + e] : 0.f;
    }
    uint4 h, l;
    split2(v[0], v[1], h.x, l.x);
    split2(v[2], v[3], h.y, l.y);
    split2(v[4], v[5], h.z, l.z);
    split2(v[6], v[7], h.w, l.w);
    *reinterpret_cast<uint4*>(out + pix * out_pitch + c) = h;
    *reinterpret_cast<uint4*>(out + pix * out_pitch + out_lo + c) = l;
  }
}

// InstanceNorm (+ residual) + ReLU of a raw fp32 convolution output, from the statistics the convolution's
// epilogue accumulated, written as split planes for the next convolution (resnet.py:64-92):
//   y = act( (x - mean) * rstd  +  [ res (split planes)  |  (rx - rmean) * rrstd (raw + its own statistics) ] )
// One pass: 4 B read (+4 B residual) and 4 B written per element.  Thread = 8 channels of a pixel.
struct ApplyParams {
  const float* x; const double* sums;
  const __nv_bfloat16* res_split; long long res_pitch; int res_lo;
  const float* res_raw; const double* res_sums;
  __nv_bfloat16* out; long long out_pitch; int out_lo;
  float* out_f32;
  int P, C, relu; float eps;
};
constexpr int kApplyThreads = 256;
constexpr int kApplyChunk = 256;     // pixels per CTA
constexpr int kApplyMaxC = 256;

// scale / shift of InstanceNorm for the channels of image n, once per CTA: thread c adds the replicas up
__device__ __forceinline__ void in_coeffs(const double* sums, int n, int C, int P, float eps, float* sc, float* sh) {
  for (int c = threadIdx.x; c < C; c += kApplyThreads) {
    double su = 0.0, sq = 0.0;
#pragma unroll
    for (int r = 0; r < SCFLOW_STATS_REPLICAS; ++r) {
      const double2 v = *reinterpret_cast<const double2*>(sums + (((size_t)n * SCFLOW_STATS_REPLICAS + r) * C + c) * 2);
      su += v.x; sq += v.y;
    }
    const double mean = su / P;
    const double var = fmax(sq / P - mean * mean, 0.0);       // biased, as F.instance_norm
    const float rstd = 1.0f / sqrtf((float)var + eps);
    sc[c] = rstd;
    sh[c] = -(float)mean * rstd;
  }
}

__global__ void __launch_bounds__(kApplyThreads, 2)
norm_apply_split_kernel(const ApplyParams p) {
  const int groups = p.C >> 3;
  const int lanes = kApplyThreads / groups;
  const int cg = threadIdx.x % groups, pl = threadIdx.x / groups;
  const int n = blockIdx.y;
  __shared__ float coef[4][kApplyMaxC];
  in_coeffs(p.sums, n, p.C, p.P, p.eps, coef[0], coef[1]);
  if (p.res_raw) in_coeffs(p.res_sums, n, p.C, p.P, p.eps, coef[2], coef[3]);
  __syncthreads();
  if (pl >= lanes) return;
  const int p0 = blockIdx.x * kApplyChunk, p1 = min(p0 + kApplyChunk, p.P);
  const int c = 8 * cg;
  float sc[8], sh[8], rsc[8], rsh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = coef[0][c + j]; sh[j] = coef[1][c + j];
    if (p.res_raw) { rsc[j] = coef[2][c + j]; rsh[j] = coef[3][c + j]; }
  }
  // two pixels per thread and iteration, every load issued before the first use (two CTAs per SM): with one
  // pixel a thread had 32-64 bytes in flight and the pass sat at 0.65 of the HBM peak
  constexpr int U = 2;
  for (int px0 = p0 + pl; px0 < p1; px0 += U * lanes) {
    float v[U][8], r[U][8];
    uint32_t hw[U][4], lw[U][4];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int px = px0 + u * lanes;
      if (px < p1) {
        const size_t pix = (size_t)n * p.P + px;
        const float4 a = __ldcs(reinterpret_cast<const float4*>(p.x + pix * p.C + c));
        const float4 b = __ldcs(reinterpret_cast<const float4*>(p.x + pix * p.C + c + 4));
        v[u][0] = a.x; v[u][1] = a.y; v[u][2] = a.z; v[u][3] = a.w; v[u][4] = b.x; v[u][5] = b.y; v[u][6] = b.z; v[u][7] = b.w;
        if (p.res_raw) {
          const float4 ra = __ldcs(reinterpret_cast<const float4*>(p.res_raw + pix * p.C + c));
          const float4 rb = __ldcs(reinterpret_cast<const float4*>(p.res_raw + pix * p.C + c + 4));
          r[u][0] = ra.x; r[u][1] = ra.y; r[u][2] = ra.z; r[u][3] = ra.w; r[u][4] = rb.x; r[u][5] = rb.y; r[u][6] = rb.z; r[u][7] = rb.w;
        } else if (p.res_split) {
          const uint4 h = __ldg(reinterpret_cast<const uint4*>(p.res_split + pix * p.res_pitch + c));
          const uint4 l = __ldg(reinterpret_cast<const uint4*>(p.res_split + pix * p.res_pitch + p.res_lo + c));
          hw[u][0] = h.x; hw[u][1] = h.y; hw[u][2] = h.z; hw[u][3] = h.w;
          lw[u][0] = l.x; lw[u][1] = l.y; lw[u][2] = l.z; lw[u][3] = l.w;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int px = px0 + u * lanes;
      if (px >= p1) break;
      const size_t pix = (size_t)n * p.P + px;
#pragma unroll
      for (int j = 0; j < 8; ++j) v[u][j] = fmaf(v[u][j], sc[j], sh[j]);
      if (p.res_raw) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[u][j] += fmaf(r[u][j], rsc[j], rsh[j]);
      } else if (p.res_split) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          v[u][2 * e] += bf_lo(hw[u][e]) + bf_lo(lw[u][e]);
          v[u][2 * e + 1] += bf_hi(hw[u][e]) + bf_hi(lw[u][e]);
        }
      }
      if (p.relu) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[u][j] = fmaxf(v[u][j], 0.f);
      }
      uint4 h, l;
      split2(v[u][0], v[u][1], h.x, l.x);
      split2(v[u][2], v[u][3], h.y, l.y);
      split2(v[u][4], v[u][5], h.z, l.z);
      split2(v[u][6], v[u][7], h.w, l.w);
      *reinterpret_cast<uint4*>(p.out + pix * p.out_pitch + c) = h;
      *reinterpret_cast<uint4*>(p.out + pix * p.out_pitch + p.out_lo + c) = l;
      if (p.out_f32) {
        *reinterpret_cast<float4*>(p.out_f32 + pix * p.C + c) = make_float4(v[u][0], v[u][1], v[u][2], v[u][3]);
        *reinterpret_cast<float4*>(p.out_f32 + pix * p.C + c + 4) = make_float4(v[u][4], v[u][5], v[u][6], v[u][7]);
      }
    }
  }
}

// planar fp32 (n, cf, P) -> channels [0, cf) of a split-plane slice (element stores: cf is 1-2 channels,
// the flow pair the reference concatenates behind the motion features, raft_decoder.py:165)
__global__ void split_planar_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long out_pitch,
                                    int out_lo, int cf, long long P, long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long pix = i % P;
    const long long r = i / P;
    const int c = (int)(r % cf);
    const long long n = r / cf;
    const float v = in[i];
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    __nv_bfloat16* o = out + (n * P + pix) * out_pitch + c;
    o[0] = hi;
    o[out_lo] = __float2bfloat16_rn(v - __bfloat162float(hi));
  }
}

// planar fp32 (n, c <= 2, h, w) -> split planes (n, h, w, [16 | 16]) holding, per pixel, its kw <= 8 x-neighbours
// x - kw/2 .. x + kw/2 of every channel (channel index dx * c + ci, zero outside the image and in the unused slots):
// a kh x kw convolution on 1-2 channels becomes a kh x 1 convolution on 16 channels, i.e. kh taps of one 16-wide
// k-step on the tensor cores, instead of kh * kw * c FMAs per output on the FP32 pipe.
__global__ void expand_taps_kernel(const float* __restrict__ in, uint4* __restrict__ out, int c, int kw, int H, int W,
                                   long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(i % W);
    const long long r = i / W;
    const int y = (int)(r % H);
    const long long n = r / H;
    float v[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = 0.f;
    for (int dx = 0; dx < kw; ++dx) {
      const int xx = x + dx - kw / 2;
      if (xx < 0 || xx >= W) continue;
      for (int ci = 0; ci < c; ++ci) v[dx * c + ci] = __ldg(in + ((n * c + ci) * H + y) * W + xx);
    }
    uint32_t hi[8], lo[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) split2(v[2 * k], v[2 * k + 1], hi[k], lo[k]);
    uint4* o = out + i * 4;                                   // 64 bytes per pixel: 16 hi, 16 lo
    o[0] = make_uint4(hi[0], hi[1], hi[2], hi[3]); o[1] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
    o[2] = make_uint4(lo[0], lo[1], lo[2], lo[3]); o[3] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
  }
}

static bool encode_bf16_map(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                            const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* estr, int bk) {
  tc::EncodeTiledFn fn = tc::get_encode_tiled();
  if (!fn) return false;
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t bdim[5], es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bdim[i] = box[i]; es[i] = estr ? estr[i] : 1; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr,
                  bdim, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  bk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

static int cout_pad_of(int cout, int* bn, int* n_tiles) {
  int c16 = (cout + 15) / 16 * 16;
  if (c16 <= 256) { *bn = c16; *n_tiles = 1; return c16; }
  *n_tiles = (c16 + 255) / 256;
  // equal tiles, multiples of 16
  int b = ((c16 + *n_tiles - 1) / *n_tiles + 15) / 16 * 16;
  *bn = b;
  return b * *n_tiles;
}

// Geometry of the patch-resident kernel for this layer, or false when the layer is outside it (stride 2,
// 32-channel k-blocks, maps that 16 x 8 tiles do not cover, rings that do not fit): the caller then takes the
// per-tap kernel.  algo: SCFLOW_CONV_AUTO | _PATCH | _PATCH_PAIR.
static bool plan_patch(const scflow_conv_tc_desc* d, Params& p, int algo, bool* pair, size_t* smem_bytes) {
  if (d->stride != 1 || d->bk != 64 || (p.OW % 8) || (p.OH % 16) || d->kw > 9 || d->kh * d->kw > MAX_W_STAGES)
    return false;
  const int taps = d->kh * d->kw;
  const int budget = SMEM_LIMIT - 1024 - P_BAR_BYTES;
  const int tiles = p.n * (p.OW / 8) * (p.OH / 16);
  // PAIR halves the filter traffic per pixel; it needs an even split of the filter rows into 8-row atoms
  bool want_pair = algo == SCFLOW_CONV_PATCH_PAIR || (algo == SCFLOW_CONV_AUTO && tiles >= 2);
  if (want_pair && (p.bn % 16)) {
    if (algo == SCFLOW_CONV_PATCH_PAIR) return false;
    want_pair = false;
  }
  p.PH = 16 + d->kh - 1;
  p.pitch = 8 + d->kw - 1;
  p.patch_tx = p.PH * p.pitch * 128;
  p.patch_bytes = (p.patch_tx + 1023) / 1024 * 1024;
  p.a_stage_bytes = 2 * p.patch_bytes;
  for (int attempt = 0; attempt < 2; ++attempt) {
    const bool pr = want_pair && attempt == 0;
    p.wb_bytes = (pr ? p.bn / 2 : p.bn) * 128;
    p.w_stage_bytes = 2 * p.wb_bytes;
    const int entries = taps * p.kblocks;
    bool ok = false;
    if (entries <= MAX_W_STAGES && p.n_tiles == 1 && entries * p.w_stage_bytes + 2 * p.a_stage_bytes <= budget) {
      p.w_resident = 1;
      p.w_stages = entries;
      p.a_stages = (budget - entries * p.w_stage_bytes) / p.a_stage_bytes;
      ok = true;
    } else {
      p.w_resident = 0;
      p.a_stages = 2;
      p.w_stages = (budget - 2 * p.a_stage_bytes) / p.w_stage_bytes;
      if (p.w_stages >= 6 && budget - 3 * p.a_stage_bytes >= 4 * p.w_stage_bytes) {
        p.a_stages = 3;
        p.w_stages = (budget - 3 * p.a_stage_bytes) / p.w_stage_bytes;
      }
      ok = p.w_stages >= 2;
    }
    if (p.a_stages > MAX_A_STAGES) p.a_stages = MAX_A_STAGES;
    if (p.w_stages > MAX_W_STAGES) p.w_stages = MAX_W_STAGES;
    if (ok) {
      *pair = pr;
      p.TW = 8; p.TH = 16; p.NI = 1;
      p.tiles_x = p.OW / 8; p.tiles_y = p.OH / 16;
      p.m_tiles = tiles;
      p.fold = (!pr && p.bn <= 128 && p.terms != 1) ? 1 : 0;
      *smem_bytes = 1024 + (size_t)p.a_stages * p.a_stage_bytes + (size_t)p.w_stages * p.w_stage_bytes + P_BAR_BYTES;
      return true;
    }
    if (algo == SCFLOW_CONV_PATCH_PAIR || !want_pair) break;
  }
  return false;
}

}  // namespace ctc
}  // namespace scflow

using namespace scflow;

extern "C" size_t scflow_conv_tc_packed_bytes(int cout, int cin, int kh, int kw, int bk) {
  if (cout <= 0 || cin <= 0 || kh <= 0 || kw <= 0 || (bk != 64 && bk != 32)) return 0;
  int bn, nt;
  const int cp = ctc::cout_pad_of(cout, &bn, &nt);
  const int kblocks = (cin + bk - 1) / bk;
  return (size_t)kh * kw * kblocks * 2 * cp * bk * sizeof(__nv_bfloat16);
}

extern "C" int scflow_conv_tc_pack_weights(const float* w, const float* scale, void* packed, int cout, int cin,
                                           int kh, int kw, int bk, void* stream) {
  if (!w || !packed) return SCFLOW_E_NULL;
  const size_t bytes = scflow_conv_tc_packed_bytes(cout, cin, kh, kw, bk);
  if (bytes == 0) return SCFLOW_E_BAD_SHAPE;
  int bn, nt;
  const int cp = ctc::cout_pad_of(cout, &bn, &nt);
  const int kblocks = (cin + bk - 1) / bk;
  const long long total = (long long)(bytes / sizeof(__nv_bfloat16));
  const int blocks = (int)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
  ctc::pack_weights_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(
      w, scale, static_cast<__nv_bfloat16*>(packed), cout, cin, kh * kw, kblocks, cp, bk, total);
  return check_launch();
}

extern "C" int scflow_split_bf16(const float* in, long long in_pitch, void* out, long long out_pitch, int out_lo,
                                 int channels, long long pixels, void* stream) {
  if (!in || !out) return SCFLOW_E_NULL;
  if (channels <= 0 || pixels < 0 || (out_pitch % 8) || (out_lo % 8)) return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(out)) return SCFLOW_E_MISALIGNED;
  if (pixels == 0) return SCFLOW_OK;
  const int c8 = (channels + 7) / 8;
  // the vector path reads float4s: needs an aligned base and pitch; otherwise every group is the tail
  const bool vec = aligned16(in) && (in_pitch % 4) == 0;
  const long long total = pixels * c8;
  const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  ctc::split_bf16_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(
      in, in_pitch, static_cast<__nv_bfloat16*>(out), out_pitch, out_lo, c8, vec ? channels : 0, pixels);
  return check_launch();
}

namespace scflow {
namespace ctc {
// A CTA finishes a band of kTapRows image rows: the slot sums of the band and its halo rows go to shared memory
// (coalesced: the pitch floats of a pixel are contiguous), then one thread per (row, x, channel) adds the taps.
constexpr int kTapRows = 4;
__global__ void __launch_bounds__(256) tap_sum_kernel(const float* __restrict__ dot, int slots, int pitch,
                                                      const float* __restrict__ bias, float* __restrict__ out,
                                                      int h, int w, int ks, int cout) {
  extern __shared__ float tsm[];                       // [kTapRows + 2 pad][w][pitch]
  const int pad = ks >> 1, rows = kTapRows + 2 * pad;
  const int bands = (h + kTapRows - 1) / kTapRows;
  const int n = blockIdx.x / bands, y0 = (blockIdx.x % bands) * kTapRows - pad;
  const int p4 = pitch >> 2, per_row4 = w * p4;         // pitch % 4 == 0: 16-byte loads, four slots in flight each
  float4* tsm4 = reinterpret_cast<float4*>(tsm);
#pragma unroll 4
  for (int i = threadIdx.x; i < rows * per_row4; i += blockDim.x) {
    const int r = i / per_row4, j = i - r * per_row4, y = y0 + r;
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    if (y >= 0 && y < h) {
      const int px = j / p4, k4 = j - px * p4;
      const float4* d = reinterpret_cast<const float4*>(dot + (((long long)n * h + y) * w + px) * slots * pitch) + k4;
      for (int sl = 0; sl < slots; ++sl) {
        const float4 a = __ldg(d + sl * p4);
        t.x += a.x; t.y += a.y; t.z += a.z; t.w += a.w;
      }
    }
    tsm4[i] = t;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kTapRows * w * cout; i += blockDim.x) {
    const int x = i % w, r = (i / w) % kTapRows, co = i / (w * kTapRows), y = y0 + pad + r;
    if (y >= h) continue;
    float s = bias ? __ldg(bias + co) : 0.f;
    for (int ky = 0; ky < ks; ++ky)
      for (int kx = 0; kx < ks; ++kx) {
        const int xx = x + kx - pad;
        if (xx >= 0 && xx < w) s += tsm[((r + ky) * w + xx) * pitch + (ky * ks + kx) * cout + co];
      }
    out[(((long long)n * cout + co) * h + y) * w + x] = s;
  }
}
}  // namespace ctc
}  // namespace scflow

extern "C" int scflow_tap_sum_fwd(const float* dot, int slots, int pitch, const float* bias, float* out, int n, int h,
                                  int w, int ks, int cout, void* stream) {
  using namespace scflow::ctc;
  if (!dot || !out) return SCFLOW_E_NULL;
  if (n <= 0 || h <= 0 || w <= 0 || cout <= 0 || slots <= 0 || (ks != 1 && ks != 3) || ks * ks * cout > pitch ||
      (pitch % 4))
    return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(dot)) return SCFLOW_E_MISALIGNED;
  const size_t smem = (size_t)(kTapRows + 2 * (ks >> 1)) * w * pitch * sizeof(float);
  if (smem > 200 * 1024) return SCFLOW_E_BAD_SHAPE;
  cudaFuncSetAttribute(tap_sum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int bands = (h + kTapRows - 1) / kTapRows;
  tap_sum_kernel<<<(unsigned)(n * bands), 256, smem, (cudaStream_t)stream>>>(dot, slots, pitch, bias, out, h, w, ks, cout);
  return check_launch();
}

extern "C" int scflow_expand_taps_fwd(const float* in, void* out, int n, int c, int kw, int h, int w, void* stream) {
  if (!in || !out) return SCFLOW_E_NULL;
  if (n <= 0 || h <= 0 || w <= 0 || c <= 0 || kw <= 0 || !(kw & 1) || c * kw > 16) return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(out)) return SCFLOW_E_MISALIGNED;
  const long long total = (long long)n * h * w;
  const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  ctc::expand_taps_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(in, static_cast<uint4*>(out), c, kw, h, w, total);
  return check_launch();
}

extern "C" int scflow_split_planar(const float* in, void* out, long long out_pitch, int out_lo, int n, int cf,
                                   long long pixels, void* stream) {
  if (!in || !out) return SCFLOW_E_NULL;
  if (n <= 0 || cf <= 0 || pixels <= 0) return SCFLOW_E_BAD_SHAPE;
  const long long total = (long long)n * cf * pixels;
  const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  ctc::split_planar_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(in, static_cast<__nv_bfloat16*>(out), out_pitch,
                                                                     out_lo, cf, pixels, total);
  return check_launch();
}

extern "C" int scflow_norm_apply_split(const float* x, const double* sums, const void* res_split, long long res_pitch,
                                       int res_lo, const float* res_raw, const double* res_sums, void* out,
                                       long long out_pitch, int out_lo, float* out_f32, int n, int P, int C,
                                       float eps, int relu, void* stream) {
  using namespace scflow::ctc;
  if (!x || !sums || !out) return SCFLOW_E_NULL;
  if ((res_raw == nullptr) != (res_sums == nullptr)) return SCFLOW_E_NULL;
  if (n <= 0 || P <= 0 || C < 8 || (C & 7) || C > kApplyMaxC || n > 65535) return SCFLOW_E_BAD_SHAPE;
  if (!aligned16(x) || !aligned16(out) || !aligned16(sums) || !aligned16(res_sums) || !aligned16(res_split) || !aligned16(res_raw) || !aligned16(out_f32) ||
      (out_pitch % 8) || (out_lo % 8) || (res_split && ((res_pitch % 8) || (res_lo % 8))))
    return SCFLOW_E_MISALIGNED;
  ApplyParams p{x, sums, static_cast<const __nv_bfloat16*>(res_split), res_pitch, res_lo, res_raw, res_sums,
                static_cast<__nv_bfloat16*>(out), out_pitch, out_lo, out_f32, P, C, relu, eps};
  dim3 grid((P + kApplyChunk - 1) / kApplyChunk, n);
  norm_apply_split_kernel<<<grid, kApplyThreads, 0, (cudaStream_t)stream>>>(p);
  return check_launch();
}

extern "C" int scflow_conv_tc_fwd(const scflow_conv_tc_desc* d, void* stream) {
  using namespace scflow::ctc;
  if (!d || !d->x || !d->wpack) return SCFLOW_E_NULL;
  if (d->bk != 64 && d->bk != 32) return SCFLOW_E_BAD_SHAPE;
  if (d->n <= 0 || d->h <= 0 || d->w <= 0 || d->cin <= 0 || d->cout <= 0 || d->kh <= 0 || d->kw <= 0 ||
      d->stride < 1 || d->stride > 2 || d->kh * d->kw > 64)
    return SCFLOW_E_BAD_SHAPE;
  Params p;
  memset(&p, 0, sizeof(p));
  p.n = d->n;
  p.OH = (d->h + 2 * d->pad_h - d->kh) / d->stride + 1;
  p.OW = (d->w + 2 * d->pad_w - d->kw) / d->stride + 1;
  if (p.OH <= 0 || p.OW <= 0) return SCFLOW_E_BAD_SHAPE;
  p.kh = d->kh; p.kw = d->kw; p.stride = d->stride; p.pad_y = d->pad_h; p.pad_x = d->pad_w;
  if (d->mma_terms < 0 || d->mma_terms > 3) return SCFLOW_E_UNSUPPORTED;
  p.terms = d->mma_terms == 0 ? 3 : d->mma_terms;
  p.cout_pad = cout_pad_of(d->cout, &p.bn, &p.n_tiles);
  if (d->bn > 0) {                             // caller-chosen tile width: must tile the pack
    if ((d->bn % 16) || d->bn > 256 || (p.cout_pad % d->bn)) return SCFLOW_E_BAD_SHAPE;
    p.bn = d->bn;
    p.n_tiles = p.cout_pad / d->bn;
  }
  p.cout_store = (d->cout + 7) / 8 * 8;
  const int bk = d->bk;
  // a second input tensor (d->x2, cin2 channels) continues the filter's input channels after the first one's:
  // what the reference builds with torch.cat before a convolution is read from where it lies
  const int cin2 = d->x2 ? d->cin2 : 0;
  if (cin2 < 0 || (d->x2 && (cin2 == 0 || (d->cin % bk) != 0))) return SCFLOW_E_BAD_SHAPE;
  const int cin_all = d->cin + cin2;
  p.kblocks = (cin_all + bk - 1) / bk;
  p.kb_split = d->x2 ? d->cin / bk : p.kblocks;
  p.ksteps_last = ((cin_all - (p.kblocks - 1) * bk) + 15) / 16;
  const int algo = d->algo & 0xff;
  if (algo > SCFLOW_CONV_PATCH_PAIR) return SCFLOW_E_BAD_SHAPE;
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  bool pair = false, patch = false;
  size_t smem_bytes = 0;
  // Geometry for tile width q.bn; returns the number of waves of the persistent grid, or a negative error.
  // AUTO: 1x1 filters have no taps to share a patch between and measured 5-9 % faster on the per-tap kernel
  // (its 128-pixel row tiles store wider runs); everything else stride-1 goes to the patch kernel, on CTA
  // pairs where the filter rows split evenly (measured faster on every 3x3 / 1x5 / 5x1 shape of the refiner)
  auto plan = [&](Params& q, bool* q_pair, bool* q_patch, size_t* q_smem) -> int {
    *q_pair = false;
    *q_patch = algo != SCFLOW_CONV_TAP_BOXES && !(algo == SCFLOW_CONV_AUTO && d->kh * d->kw == 1) &&
               plan_patch(d, q, algo, q_pair, q_smem);
    if (!*q_patch && (algo == SCFLOW_CONV_PATCH || algo == SCFLOW_CONV_PATCH_PAIR)) return SCFLOW_E_BAD_SHAPE;
    if (!*q_patch) {
      // tile rectangle: 128 output pixels
      if (q.OW >= BM) {
        if (q.OW % BM) return SCFLOW_E_BAD_SHAPE;
        q.TW = BM; q.TH = 1; q.NI = 1;
      } else {
        if (BM % q.OW) return SCFLOW_E_BAD_SHAPE;
        q.TW = q.OW;
        const int rows = BM / q.OW;
        if (q.OH >= rows) {
          if (q.OH % rows) return SCFLOW_E_BAD_SHAPE;
          q.TH = rows; q.NI = 1;
        } else {
          if (rows % q.OH) return SCFLOW_E_BAD_SHAPE;
          q.TH = q.OH; q.NI = rows / q.OH;
        }
      }
      q.tiles_x = q.OW / q.TW; q.tiles_y = q.OH / q.TH;
      const int groups = (q.n + q.NI - 1) / q.NI;
      q.m_tiles = groups * q.tiles_x * q.tiles_y;
      q.fold = (q.bn <= 128 && q.terms != 1) ? 1 : 0;
      q.a_bytes = BM * bk * 2;
      q.b_bytes = q.bn * bk * 2;
      q.stage_bytes = 2 * (q.a_bytes + q.b_bytes);
      q.stages = (SMEM_LIMIT - 2048) / q.stage_bytes;
      if (q.stages > 8) q.stages = 8;
      if (q.stages < 2) return SCFLOW_E_BAD_SHAPE;
      *q_smem = 1024 + (size_t)q.stages * q.stage_bytes + 256;
    }
    const int units = (*q_pair ? (q.m_tiles + 1) / 2 : q.m_tiles) * q.n_tiles;
    const int slots = *q_pair ? sms / 2 : sms;
    return (units + slots - 1) / slots;
  };
  {
    const int waves = plan(p, &pair, &patch, &smem_bytes);
    if (waves < 0) return waves;
    // Wave quantisation: the grid is persistent (one CTA per SM), so a launch costs whole waves of tiles.  Half
    // as wide a tile doubles the units (and the activation reads) at about half the cost each; taken when it
    // fills the last wave enough to win under cost ~ waves x (bn + 32) (the 32: the activation side of a tile).
    // 64 crops of 32 x 32 and 256 filters: 4 waves of 256 -> 7 waves of 128 (measured 131 -> 123 us, 86 -> 77 us).
    const int half = p.bn / 2;
    if (d->bn == 0 && p.n_tiles == 1 && half >= 64 && (half % (d->epi != SCFLOW_EPI_PLAIN ? 32 : 16)) == 0) {
      Params q = p;
      q.bn = half; q.n_tiles = 2;
      bool q_pair = false, q_patch = false;
      size_t q_smem = 0;
      const int w2 = plan(q, &q_pair, &q_patch, &q_smem);
      if (w2 > 0 && (long long)w2 * (half + 32) < (long long)waves * (p.bn + 32)) {
        p = q; pair = q_pair; patch = q_patch; smem_bytes = q_smem;
      }
    }
  }
  // epilogue
  p.bias = d->bias; p.relu = d->relu;
  p.res_split = static_cast<const __nv_bfloat16*>(d->res_split); p.res_pitch = d->res_pitch; p.res_lo = d->res_lo;
  p.res_f32 = d->res_f32; p.resf_pitch = d->resf_pitch;
  p.out_split = static_cast<__nv_bfloat16*>(d->out_split); p.out_pitch = d->out_pitch; p.out_lo = d->out_lo;
  p.out_f32 = d->out_f32; p.outf_pitch = d->outf_pitch; p.f32_tf32 = d->f32_tf32;
  p.out_planar = d->out_planar; p.planar_tf32 = d->planar_tf32;
  p.stats = d->stats;
  p.epi = d->epi; p.gru_c = d->gru_c; p.gru_h = d->gru_h; p.gru_z = d->gru_z;
  if (p.epi != SCFLOW_EPI_PLAIN) {
    // the gate epilogues work on whole 32-channel chunks of a 32-aligned hidden width
    if (p.epi != SCFLOW_EPI_GRU_ZR && p.epi != SCFLOW_EPI_GRU_GATE) return SCFLOW_E_UNSUPPORTED;
    if (!p.gru_h || !p.gru_z || !p.out_split || p.gru_c <= 0 || (p.gru_c % 32) || (p.bn % 32) || p.stats ||
        d->cout != (p.epi == SCFLOW_EPI_GRU_ZR ? 2 : 1) * p.gru_c)
      return SCFLOW_E_BAD_SHAPE;
    if (!aligned16(p.gru_h) || !aligned16(p.gru_z)) return SCFLOW_E_MISALIGNED;
  }
  p.dot_w = d->dot_w; p.dot_out = d->dot_out; p.dot_n = d->dot_n; p.dot_pitch = d->dot_pitch; p.dot_slots = d->dot_slots;
  if (p.dot_out) {
    if (!p.dot_w) return SCFLOW_E_NULL;
    // the slot count is part of the buffer's layout: the caller fixes the tile width (bn) and with it n_tiles
    if (p.epi != SCFLOW_EPI_PLAIN || p.dot_n < 1 || p.dot_n > SCFLOW_CONV_DOT_PITCH || p.dot_pitch < p.dot_n ||
        (p.dot_pitch % 4) || p.dot_pitch > SCFLOW_CONV_DOT_PITCH || d->bn == 0 || p.dot_slots != 2 * p.n_tiles)
      return SCFLOW_E_BAD_SHAPE;
    if (!aligned16(p.dot_w) || !aligned16(p.dot_out)) return SCFLOW_E_MISALIGNED;
  }
  if (!p.out_split && !p.out_f32 && !p.out_planar && !p.dot_out) return SCFLOW_E_NULL;
  {
    auto a32 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 31u) == 0; };
    p.wide = a32(p.out_split) && (p.out_pitch % 16) == 0 && (p.out_lo % 16) == 0 && a32(p.out_f32) &&
             (p.outf_pitch % 8) == 0 && a32(p.res_split) && (p.res_pitch % 16) == 0 && (p.res_lo % 16) == 0 &&
             a32(p.res_f32) && (p.resf_pitch % 8) == 0 && a32(p.gru_h) && a32(p.gru_z) && (p.gru_c % 8) == 0;
  }
  if (p.stats && p.OH * p.OW < 32) return SCFLOW_E_BAD_SHAPE;
  if ((p.out_split && (!aligned16(p.out_split) || (p.out_pitch % 8) || (p.out_lo % 8))) ||
      (p.out_f32 && (!aligned16(p.out_f32) || (p.outf_pitch % 4))) ||
      (p.res_split && (!aligned16(p.res_split) || (p.res_pitch % 8) || (p.res_lo % 8))) ||
      (p.res_f32 && (!aligned16(p.res_f32) || (p.resf_pitch % 4))) || !aligned16(d->x) || !aligned16(d->wpack) ||
      (d->x_lo_off % 8) || (d->x_stride_w % 8) || (d->x_stride_h % 8) || (d->x_stride_n % 8) ||
      (d->x2 && (!aligned16(d->x2) || (d->x2_lo_off % 8) || (d->x2_stride_w % 8) || (d->x2_stride_h % 8) ||
                 (d->x2_stride_n % 8))))
    return SCFLOW_E_MISALIGNED;

  Maps maps;
  {
    const uint64_t dims[4] = {(uint64_t)d->cin, (uint64_t)d->w, (uint64_t)d->h, (uint64_t)d->n};
    const uint64_t strides[3] = {(uint64_t)d->x_stride_w * 2, (uint64_t)d->x_stride_h * 2,
                                 (uint64_t)d->x_stride_n * 2};
    const uint32_t box_tap[4] = {(uint32_t)bk, (uint32_t)((p.TW - 1) * p.stride + 1),
                                 (uint32_t)((p.TH - 1) * p.stride + 1), (uint32_t)p.NI};
    const uint32_t box_patch[4] = {64u, (uint32_t)p.pitch, (uint32_t)p.PH, 1u};
    const uint32_t es[4] = {1, (uint32_t)p.stride, (uint32_t)p.stride, 1};
    const uint32_t* box = patch ? box_patch : box_tap;
    const __nv_bfloat16* xb = static_cast<const __nv_bfloat16*>(d->x);
    if (!encode_bf16_map(&maps.a_hi, xb, 4, dims, strides, box, es, bk) ||
        !encode_bf16_map(&maps.a_lo, xb + d->x_lo_off, 4, dims, strides, box, es, bk))
      return SCFLOW_E_DRIVER;
    maps.b_hi = maps.a_hi; maps.b_lo = maps.a_lo;
    if (d->x2) {
      const uint64_t dims2[4] = {(uint64_t)cin2, (uint64_t)d->w, (uint64_t)d->h, (uint64_t)d->n};
      const uint64_t strides2[3] = {(uint64_t)d->x2_stride_w * 2, (uint64_t)d->x2_stride_h * 2,
                                    (uint64_t)d->x2_stride_n * 2};
      const __nv_bfloat16* xb2 = static_cast<const __nv_bfloat16*>(d->x2);
      if (!encode_bf16_map(&maps.b_hi, xb2, 4, dims2, strides2, box, es, bk) ||
          !encode_bf16_map(&maps.b_lo, xb2 + d->x2_lo_off, 4, dims2, strides2, box, es, bk))
        return SCFLOW_E_DRIVER;
    }
    const uint64_t wd[2] = {(uint64_t)bk, (uint64_t)d->kh * d->kw * p.kblocks * 2 * p.cout_pad};
    const uint64_t ws[1] = {(uint64_t)bk * 2};
    const uint32_t wb[2] = {(uint32_t)bk, (uint32_t)(pair ? p.bn / 2 : p.bn)};
    if (!encode_bf16_map(&maps.w, d->wpack, 2, wd, ws, wb, nullptr, bk)) return SCFLOW_E_DRIVER;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (patch) {
    const int units = (pair ? (p.m_tiles + 1) / 2 : p.m_tiles) * p.n_tiles;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.blockDim = dim3(P_THREADS);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = pair ? 2 : 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    const bool gru = p.epi != SCFLOW_EPI_PLAIN;
    auto launch = [&](auto kern) -> int {
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
      if (pair) {
        int clusters = 0;
        cfg.gridDim = dim3(2 * (sms / 2));
        if (cudaOccupancyMaxActiveClusters(&clusters, kern, &cfg) != cudaSuccess || clusters <= 0) {
          cudaGetLastError();
          return SCFLOW_E_LAUNCH;
        }
        if (clusters > sms / 2) clusters = sms / 2;
        cfg.gridDim = dim3(2 * (units < clusters ? units : clusters));
      } else {
        cfg.gridDim = dim3(units < sms ? units : sms);
      }
      cudaLaunchKernelEx(&cfg, kern, maps, p);
      return check_launch();
    };
    if (pair) return gru ? launch(conv_patch_kernel<true, kEpiGru>)
                         : p.dot_out ? launch(conv_patch_kernel<true, kEpiDot>) : launch(conv_patch_kernel<true, kEpiPlain>);
    return gru ? launch(conv_patch_kernel<false, kEpiGru>)
               : p.dot_out ? launch(conv_patch_kernel<false, kEpiDot>) : launch(conv_patch_kernel<false, kEpiPlain>);
  }
  const int total = p.m_tiles * p.n_tiles;
  const int grid = total < sms ? total : sms;
  auto launch1 = [&](auto kern) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    kern<<<grid, THREADS, smem_bytes, st>>>(maps, p);
  };
  const bool gru = p.epi != SCFLOW_EPI_PLAIN;
  if (bk == 64) {
    if (gru) launch1(conv_tc_kernel<64, kEpiGru>);
    else if (p.dot_out) launch1(conv_tc_kernel<64, kEpiDot>);
    else launch1(conv_tc_kernel<64, kEpiPlain>);
  } else {
    if (gru) launch1(conv_tc_kernel<32, kEpiGru>);
    else if (p.dot_out) launch1(conv_tc_kernel<32, kEpiDot>);
    else launch1(conv_tc_kernel<32, kEpiPlain>);
  }
  return check_launch();
}
