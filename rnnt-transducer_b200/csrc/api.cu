// Version / status strings / shape helpers of the C ABI (include/rnnt_b200.h).
#include "common.cuh"

extern "C" int rnnt_b200_version(void) { return RNNT_B200_VERSION; }

extern "C" const char* rnnt_b200_status_string(int status) {
  switch (status) {
    case RNNT_B200_OK: return "ok";
    case RNNT_B200_E_NULL: return "required pointer is NULL";
    case RNNT_B200_E_SHAPE: return "shape out of the supported range";
    case RNNT_B200_E_BLANK: return "blank index not in [0,V)";
    case RNNT_B200_E_ALIGN: return "pointer alignment";
    case RNNT_B200_E_WORKSPACE: return "workspace too small";
    case RNNT_B200_E_DRIVER: return "cuTensorMapEncodeTiled unavailable or failed";
    case RNNT_B200_E_ARCH: return "device is not sm_100";
    default: break;
  }
  if (status > 0) return cudaGetErrorString((cudaError_t)status);
  return "unknown status";
}

extern "C" size_t rnnt_b200_plane_elems(int B, int T, int U1) {
  if (B <= 0 || T <= 0 || U1 <= 0) return 0;
  return (size_t)B * (size_t)(T + U1 - 1) * (size_t)U1;
}
