// ConvGRU (SeqConv) pass as a fused tcgen05 implicit GEMM -- placeholder entry points until the
// kernel lands: they reject the call (no fallback), the Python ConvGRU module then uses cuDNN.
#include "common.cuh"

extern "C" size_t scflow_convgru_packed_floats(int ch, int cx) {
  return (size_t)3 * ch * (ch + cx) * 5 + 3 * (size_t)ch;
}

extern "C" int scflow_convgru_pack_weights(const float*, const float*, const float*, const float*,
                                           const float*, const float*, float*, int, int, int, void*) {
  return SCFLOW_E_UNIMPLEMENTED;
}

extern "C" int scflow_convgru_pass_fwd(const float*, const float*, const float*, float*, int, int,
                                       int, int, int, int, void*) {
  return SCFLOW_E_UNIMPLEMENTED;
}
