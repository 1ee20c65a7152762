// alpha/beta anti-diagonal wavefront + occupation coefficients.
//
// Replaces the lattice stage of warprnnt_pytorch.RNNTLoss as called at
// /root/reference rnnt/model.py:70. One CTA per
// (utterance, direction): blockIdx.y = 0 runs alpha from (0,0), = 1 runs beta from
// (T_b-1,U_b), concurrently on different SMs. Thread u owns lattice column u; a
// wavefront step d handles the cells t+u = d. The planes are stored anti-diagonal
// major (rnnt_b200.h), so each step reads one contiguous row of lp_blank/lp_label
// (coalesced, requested RING-1 steps ahead with cp.async) and writes one contiguous row of alpha/beta.
// Neighbour exchange goes through a double-buffered shared-memory diagonal.
// Accumulators are float64: with V=4232 the per-utterance log-likelihood is ~1e3 nats
// and fp32 alpha/beta miss the 1e-4 gradient tolerance.
#include "common.cuh"

namespace {

constexpr int RING = 16;  // diagonals in flight (cp.async ring in shared memory)

// 4-byte asynchronous global->shared copies: the two log-prob rows of diagonal d+RING-1 are
// requested while diagonal d is computed, so the L2/HBM latency (which bounded the PF=4
// register-prefetch version: ~1.1 us per wavefront step, long_scoreboard) is covered by
// RING-1 steps of work. Every thread reads back only the words it copied itself.
__device__ __forceinline__ void cp_async4(float* dst_smem, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)),
               "l"(src)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

__global__ void __launch_bounds__(1024, 1)
lattice_kernel(const float* __restrict__ lp_blank, const float* __restrict__ lp_label,
               const int32_t* __restrict__ act_lens, const int32_t* __restrict__ label_lens,
               int T, int U1, double* __restrict__ alpha, double* __restrict__ beta,
               double* __restrict__ ll_alpha, double* __restrict__ ll_beta) {
  __shared__ double xch[2][1024 + 2];
  extern __shared__ float ring[];  // [RING][2][blockDim.x]
  const int b = blockIdx.x;
  const int u = threadIdx.x;
  const int nthr = blockDim.x;
  const int Tb = clamp_int(act_lens[b], 1, T);
  const int Ub = clamp_int(label_lens[b], 0, U1 - 1);
  const int nd = Tb + Ub;  // diagonals 0 .. nd-1
  const size_t base = (size_t)b * (size_t)(T + U1 - 1) * (size_t)U1;
  const float* pb = lp_blank + base;
  const float* pl = lp_label + base;
  const bool col_ok = (u <= Ub);
  const double NEG = -INFINITY;
  float* my_ring = ring + u;

  xch[0][u + 1] = NEG;
  xch[1][u + 1] = NEG;
  if (u == 0) { xch[0][0] = NEG; xch[1][0] = NEG; xch[0][blockDim.x + 1] = NEG; xch[1][blockDim.x + 1] = NEG; }
  __syncthreads();

  // request the rows of wavefront step s (alpha: diagonal s; beta: diagonal nd-1-s) into slot s%RING
  const bool is_beta = blockIdx.y != 0;
  auto request = [&](int s) {
    if (s < nd && col_ok) {
      const int d = is_beta ? nd - 1 - s : s;
      const int t = d - u;
      if (t >= 0 && t < Tb) {  // cells outside the valid window are never written by the producer
        float* slot = my_ring + ((s & (RING - 1)) * 2) * nthr;
        cp_async4(slot, pb + (size_t)d * U1 + u);
        cp_async4(slot + nthr, pl + (size_t)d * U1 + u);
      }
    }
    cp_async_commit();  // one group per step, empty or not: keeps wait_group counting uniform
  };
#pragma unroll 1
  for (int s = 0; s < RING - 1; ++s) request(s);

  if (!is_beta) {
    // ---------------- alpha ----------------
    double* out = alpha + base;
    double up = NEG;  // alpha[t-1,u] + lp_blank[t-1,u]
#pragma unroll 1
    for (int d = 0; d < nd; ++d) {
      request(d + RING - 1);
      cp_async_wait<RING - 1>();  // the group of step d has landed
      const float* slot = my_ring + ((d & (RING - 1)) * 2) * nthr;
      const int t = d - u;
      const bool valid = col_ok && t >= 0 && t < Tb;
      const double left = xch[d & 1][u];  // slot u holds the message of thread u-1
      const double a = (d == 0 && u == 0) ? 0.0 : log_add_exp(up, left);
      double right = NEG;
      if (valid) {
        const float lpb = slot[0], lpl = slot[nthr];
        out[(size_t)d * U1 + u] = a;
        up = a + (double)lpb;
        if (u < Ub) right = a + (double)lpl;
      } else {
        up = NEG;
      }
      xch[(d + 1) & 1][u + 1] = right;
      __syncthreads();
    }
    if (u == Ub) ll_alpha[b] = up;  // alpha[Tb-1,Ub] + lp_blank[Tb-1,Ub]
  } else {
    // ---------------- beta ----------------
    double* out = beta + base;
    double own = NEG;  // beta[t+1,u]
#pragma unroll 1
    for (int s = 0; s < nd; ++s) {  // step counter; diagonal d = nd-1-s
      request(s + RING - 1);
      cp_async_wait<RING - 1>();
      const float* slot = my_ring + ((s & (RING - 1)) * 2) * nthr;
      const int d = nd - 1 - s;
      const int t = d - u;
      const bool valid = col_ok && t >= 0 && t < Tb;
      const double right = xch[s & 1][u + 2];  // message of thread u+1 (beta[t,u+1])
      double be = NEG;
      if (valid) {
        const float lpb = slot[0], lpl = slot[nthr];
        if (t == Tb - 1 && u == Ub) {
          be = (double)lpb;
        } else {
          const double viab = own + (double)lpb;
          const double vial = (u < Ub) ? right + (double)lpl : NEG;
          be = log_add_exp(viab, vial);
        }
        out[(size_t)d * U1 + u] = be;
      }
      own = be;
      xch[(s + 1) & 1][u + 1] = be;
      __syncthreads();
    }
    if (u == 0) ll_beta[b] = own;  // beta[0,0]
  }
}

// Elementwise over the valid cells of every utterance; one block row per (b, diagonal).
__global__ void __launch_bounds__(256)
coeff_kernel(const float* __restrict__ lp_blank, const float* __restrict__ lp_label,
             const int32_t* __restrict__ act_lens, const int32_t* __restrict__ label_lens,
             int T, int U1, const double* __restrict__ alpha, const double* __restrict__ beta,
             const double* __restrict__ ll_beta,
             float* __restrict__ c_blank, float* __restrict__ c_label,
             float* __restrict__ loss, double* __restrict__ loss64) {
  const int b = blockIdx.y;
  const int Tb = clamp_int(act_lens[b], 1, T);
  const int Ub = clamp_int(label_lens[b], 0, U1 - 1);
  const int nd = Tb + Ub;
  const double ll = ll_beta[b];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    loss[b] = (float)(-ll);
    if (loss64) loss64[b] = -ll;
  }
  const size_t base = (size_t)b * (size_t)(T + U1 - 1) * (size_t)U1;
  const long long n = (long long)nd * U1;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(i / U1);
    const int u = (int)(i - (long long)d * U1);
    const int t = d - u;
    if (u > Ub || t < 0 || t >= Tb) continue;
    const size_t idx = base + (size_t)i;
    const double a = alpha[idx];
    double cb = 0.0, cl = 0.0;
    if (t + 1 < Tb) cb = exp(a + (double)lp_blank[idx] + beta[idx + U1] - ll);
    else if (u == Ub) cb = exp(a + (double)lp_blank[idx] - ll);
    if (u < Ub) cl = exp(a + (double)lp_label[idx] + beta[idx + U1 + 1] - ll);
    c_blank[idx] = (float)cb;
    c_label[idx] = (float)cl;
  }
}

}  // namespace

extern "C" int rnnt_b200_lattice(const float* lp_blank, const float* lp_label,
                                 const int32_t* act_lens, const int32_t* label_lens,
                                 int B, int T, int U1, double* alpha, double* beta,
                                 float* c_blank, float* c_label, float* loss, double* loss64,
                                 void* stream) {
  if (!lp_blank || !lp_label || !act_lens || !label_lens || !alpha || !beta || !c_blank ||
      !c_label || !loss)
    return RNNT_B200_E_NULL;
  if (B <= 0 || T <= 0 || U1 <= 0 || U1 > 1024) return RNNT_B200_E_SHAPE;
  {
    const int arch = rnnt_check_arch();
    if (arch != RNNT_B200_OK) return arch;
  }
  cudaStream_t st = (cudaStream_t)stream;
  // ll scratch lives in the (never read) tail of the planes?  No: keep it explicit --
  // the last diagonal row D-1 of utterance b is only used when T_b=T and U_b=U1-1, and
  // then only its column U1-1; columns 0/1 of row 0 are taken. We instead reuse loss64
  // when given, else the first two slots past the valid alpha/beta rows are not safe, so
  // require a tiny scratch: alpha/beta planes are allocated with 2*B extra doubles.
  double* ll_a = alpha + rnnt_b200_plane_elems(B, T, U1);
  double* ll_b = beta + rnnt_b200_plane_elems(B, T, U1);
  const int threads = ((U1 + 31) / 32) * 32;
  const size_t ring_bytes = (size_t)RING * 2 * threads * sizeof(float);  // <= 128 KB at 1024 threads
  // static (the diagonal exchange) + dynamic (the ring) shared memory above the 48 KB default needs the opt-in;
  // set it whenever the sum can get there (U+1 >= 225 with the 16 KB exchange array)
  if (ring_bytes > 16 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(lattice_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)ring_bytes);
    if (e != cudaSuccess) return (int)e;
  }
  lattice_kernel<<<dim3(B, 2), threads, ring_bytes, st>>>(lp_blank, lp_label, act_lens, label_lens, T, U1,
                                                          alpha, beta, ll_a, ll_b);
  RNNT_CUDA_RETURN_IF_ERROR();
  const long long n = (long long)(T + U1 - 1) * U1;
  int gx = (int)((n + 255) / 256);
  if (gx > 592) gx = 592;
  coeff_kernel<<<dim3(gx, B), 256, 0, st>>>(lp_blank, lp_label, act_lens, label_lens, T, U1, alpha,
                                            beta, ll_b, c_blank, c_label, loss, loss64);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}

// ---------------------------------------------------------------------------------------
// rnnt_b200_lattice_static: rnnt_b200_lattice for callers that cannot allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc). The float64 alpha / beta workspace lives in a file-scope device array
// of RNNT_B200_STATIC_PLANE_ELEMS + RNNT_B200_STATIC_MAX_B doubles per plane (2 x 16 MiB + 16 KiB: every GPU that
// loads the library holds it until exit). No counterpart in the reference.
// Every call on a device shares the array, which keeps the last call's contents: the kernels write every
// alpha / beta cell (and both log-likelihood slots) before reading it, so nothing has to start at zero -- but two
// calls that overlap on one GPU (two streams, two threads) race on it.
// ---------------------------------------------------------------------------------------
namespace {
constexpr size_t kStaticPlane = (size_t)RNNT_B200_STATIC_PLANE_ELEMS + (size_t)RNNT_B200_STATIC_MAX_B;
}
static __device__ __align__(256) double g_lattice_scratch[2 * kStaticPlane];

// Device address of the scratch array on the CURRENT device. In host code `g_lattice_scratch` is only a stand-in
// for the symbol, not a device address: the address comes from cudaGetSymbolAddress, once per device.
static int lattice_static_scratch(double** out) {
  static double* cached[64];  // per device ordinal; benign race (every writer stores the same value)
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev >= 0 && dev < 64 && cached[dev]) { *out = cached[dev]; return RNNT_B200_OK; }
  double* ptr = nullptr;
  e = cudaGetSymbolAddress((void**)&ptr, g_lattice_scratch);
  if (e != cudaSuccess) return (int)e;
  if (!ptr) return RNNT_B200_E_DRIVER;
  if (dev >= 0 && dev < 64) cached[dev] = ptr;
  *out = ptr;
  return RNNT_B200_OK;
}

extern "C" int rnnt_b200_lattice_static(const float* lp_blank, const float* lp_label,
                                        const int32_t* act_lens, const int32_t* label_lens,
                                        int B, int T, int U1, float* c_blank, float* c_label,
                                        float* loss, double* loss64, void* stream) {
  if (!lp_blank || !lp_label || !act_lens || !label_lens || !c_blank || !c_label || !loss) return RNNT_B200_E_NULL;
  if (B <= 0 || T <= 0 || U1 <= 0 || U1 > 1024) return RNNT_B200_E_SHAPE;
  // the fixed scratch holds B <= RNNT_B200_STATIC_MAX_B utterances and B*(T+U1-1)*U1 <= RNNT_B200_STATIC_PLANE_ELEMS cells
  if (B > RNNT_B200_STATIC_MAX_B || rnnt_b200_plane_elems(B, T, U1) > (size_t)RNNT_B200_STATIC_PLANE_ELEMS)
    return RNNT_B200_E_WORKSPACE;
  double* scratch = nullptr;
  const int rc = lattice_static_scratch(&scratch);
  if (rc != RNNT_B200_OK) return rc;
  // rnnt_b200_lattice keeps its two log-likelihood slots right behind the plane_elems(B,T,U1) cells of each plane
  return rnnt_b200_lattice(lp_blank, lp_label, act_lens, label_lens, B, T, U1, scratch, scratch + kStaticPlane,
                           c_blank, c_label, loss, loss64, stream);
}
