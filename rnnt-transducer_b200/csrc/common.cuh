// Shared helpers for the rnnt_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/rnnt_b200.h"

#define RNNT_TILE_M 128            // lattice cells per GEMM tile
#define RNNT_LOG2E 1.4426950408889634f
#define RNNT_LN2 0.6931471805599453f
#define RNNT_NEG_BIG (-1.0e30f)    // bias of padded vocabulary columns (log2 domain)

// Timing-experiment knobs (environment variables, clock64 traces) exist only in the -DRNNT_B200_TUNE build
// (`make tune` -> librnnt_b200_tune.so, loaded through RNNT_B200_LIB by tools/). The release library reads no
// environment variable, allocates nothing and never synchronises inside an entry point.
#ifdef RNNT_B200_TUNE
#include <stdlib.h>
static inline int rnnt_tune_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}
#else
#define rnnt_tune_int(name, dflt) (dflt)
#endif

#define RNNT_CUDA_RETURN_IF_ERROR()                        \
  do {                                                     \
    cudaError_t e__ = cudaGetLastError();                  \
    if (e__ != cudaSuccess) return (int)e__;               \
  } while (0)

// RNNT_B200_E_ARCH unless the current device is sm_100 (the kernels are compiled for sm_100a only).
static inline int rnnt_check_arch() {
  static int cached[64];  // 0 = not looked at yet, 1 = sm_100, -1 = something else (benign race: same value)
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return RNNT_B200_E_ARCH;
  if (dev >= 0 && dev < 64 && cached[dev]) return cached[dev] > 0 ? RNNT_B200_OK : RNNT_B200_E_ARCH;
  int major = 0, minor = 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev) != cudaSuccess)
    return RNNT_B200_E_ARCH;
  const int ok = (major == 10 && minor == 0) ? 1 : -1;
  if (dev >= 0 && dev < 64) cached[dev] = ok;
  return ok > 0 ? RNNT_B200_OK : RNNT_B200_E_ARCH;
}

// Row pitch (elements) of the tile-packed bf16 hidden-activation stash. H = 128, 256, 512 (the shapes the
// own dW2 GEMM covers, which sums db2 itself): H, rows of whole 128-byte lines. Other H: H + 32 with the
// ones column at column H (its library GEMM against dS gives d b2) and padding up to the pitch.
#ifndef RNNT_H_PAD
#define RNNT_H_PAD 32
#endif
__host__ __device__ constexpr int rnnt_hp(int H) {
  return (H == 128 || H == 256 || H == 512) ? H : H + RNNT_H_PAD;
}

// Anti-diagonal-major plane index of cell (t,u) of utterance b (see rnnt_b200.h).
__host__ __device__ __forceinline__ size_t plane_index(int b, int t, int u, int T, int U1) {
  return ((size_t)b * (size_t)(T + U1 - 1) + (size_t)(t + u)) * (size_t)U1 + (size_t)u;
}

// 16-byte read of data that is touched once: read-only path, no L1 allocation.
__device__ __forceinline__ uint4 ld_stream_u4(const uint4* p) {
  uint4 v;
  asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

__device__ __forceinline__ int clamp_int(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// log(exp(a)+exp(b)): the running value and the difference a-b are float64 (alpha/beta reach
// ~1e3..1e4 nats, where fp32 would lose the 1e-4 gradient tolerance -- SURVEY.md F7); the
// increment log1p(exp(-|a-b|)) lies in [0, ln 2] and is evaluated in fp32 (absolute error
// ~1e-7 per wavefront step), which takes the float64 transcendentals off the serial chain.
__device__ __forceinline__ double log_add_exp(double a, double b) {
  const double m = fmax(a, b);
  const float d = (float)(-fabs(a - b));
  // d is NaN only when both are -inf (inf-inf); then m = -inf is the answer.
  if (!(d == d)) return m;
  return m + (double)log1pf(__expf(d));
}
