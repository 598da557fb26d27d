// ABI bookkeeping: version, error strings.
#include "common.cuh"

namespace scflow {
thread_local int g_last_cuda_error = 0;
}

extern "C" int scflow_abi_version(void) { return SCFLOW_ABI_VERSION; }

extern "C" int scflow_last_cuda_error(void) { return scflow::g_last_cuda_error; }

extern "C" const char* scflow_strerror(int rc) {
  switch (rc) {
    case SCFLOW_OK: return "ok";
    case SCFLOW_E_BAD_SHAPE: return "shape not supported by the sm_100a kernel";
    case SCFLOW_E_UNSUPPORTED: return "mode/flag not implemented (no fallback by design)";
    case SCFLOW_E_MISALIGNED: return "pointer is not 16-byte aligned";
    case SCFLOW_E_NULL: return "required pointer is NULL";
    case SCFLOW_E_LAUNCH: return "CUDA launch failed (see scflow_last_cuda_error)";
    case SCFLOW_E_WORKSPACE: return "workspace too small";
    case SCFLOW_E_DRIVER: return "CUDA driver entry point unavailable";
    case SCFLOW_E_UNIMPLEMENTED: return "kernel not implemented yet";
    case SCFLOW_E_STATIC_SCRATCH: return "problem exceeds the static device scratch of the no-allocation entry point";
    default: return "unknown scflow error";
  }
}
