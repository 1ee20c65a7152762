// Boundary 1 (strict drop-in on materialised logits): the two HBM-streaming kernels
// around the lattice.
//
//   logits_to_planes : replaces the log-softmax stage of warprnnt_pytorch.RNNTLoss
//                      (/root/reference rnnt/model.py:70). Reads each
//                      logit exactly once, emits 12 B per cell (lse, lp_blank, lp_label).
//   planes_to_dlogits: the logit gradient (row a8), produced in the backward from the
//                      occupation coefficients instead of being stashed in the forward.
//
// One warp per lattice cell; 16-byte vector loads/stores when V % 4 == 0; grid sized to a
// multiple of the SM count with a grid-stride loop over cells.
#include "common.cuh"
#include "sm100_ptx.cuh"

namespace {

constexpr int WARPS_PER_BLOCK = 8;

// Online log-sum-exp state of a row: m = running maximum (natural domain, exact), sp = sum of
// exp(x_i - m) over the elements seen EXCEPT one occurrence of the maximum, so that
// lse = m + log1p(sp) and lp_k = (x_k - m) - log1p(sp) keep their relative accuracy when one logit
// dominates (trained-like inputs: lp of the dominant symbol ~ -sp, which m + log(1 + sp) - x_k would
// return with the absolute error of fp32 at |x| ~ 10: 1e-6 against per-utterance losses of 1e-3).
__device__ __forceinline__ void online_merge(float& m, float& sp, float m2, float sp2) {
  const float mn = fmaxf(m, m2);
  float r;
  if (mn == -INFINITY) r = 0.f;                                                  // both empty
  else if (m == m2) r = sp + sp2 + 1.f;                                          // one of the two maxima turns ordinary
  else if (m > m2) r = sp + (sp2 + 1.f) * sm100::ex2_approx((m2 - m) * RNNT_LOG2E);          // sm100::ex2_approx(-inf) = 0: empty side adds nothing
  else r = sp2 + (sp + 1.f) * sm100::ex2_approx((m - m2) * RNNT_LOG2E);
  m = mn;
  sp = r;
}
// one more element x into (m, sp)
__device__ __forceinline__ void online_push(float& m, float& sp, float x) {
  if (x <= m) {
    sp += sm100::ex2_approx((x - m) * RNNT_LOG2E);        // ordinary term (x = -inf adds 0)
  } else if (x > m) {                         // new maximum: the old one turns ordinary (m = -inf: (0 + 1) * 0)
    sp = (sp + 1.f) * sm100::ex2_approx((m - x) * RNNT_LOG2E);
    m = x;
  } else {                                    // NaN logit: poison the row like the plain formulation does
    sp = x;
  }
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
logits_to_planes_kernel(const float* __restrict__ logits, const int32_t* __restrict__ labels,
                        const int32_t* __restrict__ act_lens, const int32_t* __restrict__ label_lens,
                        int B, int T, int U1, int V, int blank,
                        float* __restrict__ lse_out, float* __restrict__ lpb_out,
                        float* __restrict__ lpl_out) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * WARPS_PER_BLOCK;
  const long long ncell = (long long)B * T * U1;
  const int U = U1 - 1;
  for (long long cell = warp0; cell < ncell; cell += nwarps) {
    const int b = (int)(cell / ((long long)T * U1));
    const int r = (int)(cell - (long long)b * T * U1);
    const int t = r / U1, u = r - t * U1;
    const int Tb = clamp_int(act_lens[b], 1, T);
    const int Ub = clamp_int(label_lens[b], 0, U);
    if (t >= Tb || u > Ub) continue;  // warp-uniform
    const float* x = logits + (size_t)cell * V;
    float m = -INFINITY, sp = 0.f;
    if ((V & 3) == 0) {
      const float4* x4 = reinterpret_cast<const float4*>(x);
      const int n4 = V >> 2;
      for (int i = lane; i < n4; i += 32) {
        const float4 v = __ldcs(x4 + i);
        const float mn = fmaxf(fmaxf(m, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
        if (mn == -INFINITY) continue;  // only masked (-inf) logits so far
        // five candidates (old maximum, four new logits): the first one equal to mn is the excluded
        // occurrence, the others are ordinary terms exp(c - mn): five exponentials per four logits,
        // branch free (a fast path for "maximum unchanged" diverges within the warp and was slower)
        bool taken = (m == mn);
        float acc = taken ? sp : (sp + 1.f) * sm100::ex2_approx((m - mn) * RNNT_LOG2E);  // m = -inf: (0 + 1) * 0
        const float c[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const bool is = !taken && (c[j] == mn);
          taken = taken || is;
          const float e = sm100::ex2_approx((c[j] - mn) * RNNT_LOG2E);
          acc += is ? 0.f : e;
        }
        sp = acc;
        m = mn;
      }
    } else {
      for (int i = lane; i < V; i += 32) online_push(m, sp, __ldcs(x + i));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
      const float s2 = __shfl_xor_sync(0xffffffffu, sp, o);
      online_merge(m, sp, m2, s2);
    }
    if (lane == 0) {
      const float lg = log1pf(sp);
      const size_t idx = plane_index(b, t, u, T, U1);
      lse_out[idx] = m + lg;
      lpb_out[idx] = (x[blank] - m) - lg;
      float ll = -INFINITY;
      if (u < Ub) {
        const int y = clamp_int(labels[(size_t)b * U + u], 0, V - 1);
        ll = (x[y] - m) - lg;
      }
      lpl_out[idx] = ll;
    }
  }
}

// Exact n / d for n < 2^31 by one multiply-high and a shift (d fixed per launch; the two per-cell
// divisions by T*U1 and U1 were 10 % of the issued instructions at small V).
struct FastDiv {
  uint32_t mul, shr, d;
};
__host__ inline FastDiv make_fastdiv(uint32_t d) {
  FastDiv f;
  f.d = d;
  if (d == 1) { f.mul = 0; f.shr = 0; return f; }
  uint32_t lg = 0;
  while ((1ull << lg) < d) ++lg;  // ceil(log2 d)
  const uint32_t pw = 31 + lg;
  f.mul = (uint32_t)(((1ull << pw) + d - 1) / d);
  f.shr = pw - 32;
  return f;
}
__device__ __forceinline__ uint32_t fastdiv(uint32_t n, const FastDiv& f) {
  return f.d == 1 ? n : (__umulhi(n, f.mul) >> f.shr);
}

// Small vocabularies (V <= 256, e.g. TIMIT's 49): G lanes per lattice cell, 32/G cells per warp. With one
// warp per cell the kernel was bound by issued instructions (83 % issue-active, 5 % of the HBM peak: the
// per-cell index arithmetic, five shuffle-merge rounds and the tail are paid per WARP); packing 2 or 4
// cells into a warp divides them. Same (m, sp) formulation as above, scalar loads (rows of V floats
// are only 4-byte aligned). Cell count < 2^31 (the host falls back to the kernel above otherwise).
template <int G>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
logits_to_planes_small_kernel(const float* __restrict__ logits, const int32_t* __restrict__ labels,
                              const int32_t* __restrict__ act_lens, const int32_t* __restrict__ label_lens,
                              int B, int T, int U1, int V, int blank, FastDiv dTU, FastDiv dU1,
                              float* __restrict__ lse_out, float* __restrict__ lpb_out,
                              float* __restrict__ lpl_out) {
  constexpr int CPW = 32 / G;
  const int lane = threadIdx.x & 31;
  const int sub = lane / G, gl = lane % G;
  const uint32_t ncell = (uint32_t)B * (uint32_t)T * (uint32_t)U1;
  const uint32_t stride = gridDim.x * WARPS_PER_BLOCK * CPW;
  const int U = U1 - 1;
  for (uint32_t c0 = (blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5)) * CPW; c0 < ncell; c0 += stride) {
    const uint32_t cell = c0 + sub;
    bool valid = cell < ncell;
    int b = 0, t = 0, u = 0, Ub = 0;
    if (valid) {
      b = (int)fastdiv(cell, dTU);
      const uint32_t r = cell - (uint32_t)b * dTU.d;
      t = (int)fastdiv(r, dU1);
      u = (int)(r - (uint32_t)t * dU1.d);
      const int Tb = clamp_int(__ldg(act_lens + b), 1, T);
      Ub = clamp_int(__ldg(label_lens + b), 0, U);
      valid = t < Tb && u <= Ub;
    }
    const float* x = logits + (size_t)cell * V;
    float m = -INFINITY, sp = 0.f, xb = 0.f, xl = 0.f;
    int y = -1;
    if (valid) {
      if (u < Ub) y = clamp_int(__ldg(labels + (size_t)b * U + u), 0, V - 1);
      xb = __ldg(x + blank);
      for (int i = gl; i < V; i += G) online_push(m, sp, __ldcs(x + i));
      if (y >= 0) xl = __ldg(x + y);
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {  // stays inside the group of G lanes; every lane takes part
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
      const float s2 = __shfl_xor_sync(0xffffffffu, sp, o);
      online_merge(m, sp, m2, s2);
    }
    if (valid && gl == 0) {
      const float lg = log1pf(sp);
      const size_t idx = plane_index(b, t, u, T, U1);
      lse_out[idx] = m + lg;
      lpb_out[idx] = (xb - m) - lg;
      lpl_out[idx] = (y >= 0) ? (xl - m) - lg : -INFINITY;
    }
    if (c0 > 0xffffffffu - stride) break;  // 32-bit wrap guard
  }
}

__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
planes_to_dlogits_kernel(const float* __restrict__ logits, const int32_t* __restrict__ labels,
                         const int32_t* __restrict__ act_lens, const int32_t* __restrict__ label_lens,
                         int B, int T, int U1, int V, int blank,
                         const float* __restrict__ lse_p, const float* __restrict__ cb_p,
                         const float* __restrict__ cl_p, const float* __restrict__ scale,
                         float* __restrict__ dlogits) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * WARPS_PER_BLOCK;
  const long long ncell = (long long)B * T * U1;
  const int U = U1 - 1;
  const bool vec = ((V & 3) == 0);
  for (long long cell = warp0; cell < ncell; cell += nwarps) {
    const int b = (int)(cell / ((long long)T * U1));
    const int r = (int)(cell - (long long)b * T * U1);
    const int t = r / U1, u = r - t * U1;
    const int Tb = clamp_int(act_lens[b], 1, T);
    const int Ub = clamp_int(label_lens[b], 0, U);
    float* g = dlogits + (size_t)cell * V;
    if (t >= Tb || u > Ub) {  // outside the valid window: exact zeros
      if (vec) {
        float4* g4 = reinterpret_cast<float4*>(g);
        for (int i = lane; i < (V >> 2); i += 32) __stcs(g4 + i, make_float4(0.f, 0.f, 0.f, 0.f));
      } else {
        for (int i = lane; i < V; i += 32) __stcs(g + i, 0.f);
      }
      continue;
    }
    const float* x = logits + (size_t)cell * V;
    const size_t idx = plane_index(b, t, u, T, U1);
    const float sc = scale[b];
    const float cb = cb_p[idx] * sc, cl = cl_p[idx] * sc;
    const float c0 = cb + cl;
    const float nl = -lse_p[idx] * RNNT_LOG2E;
    const int y = (u < Ub) ? clamp_int(labels[(size_t)b * U + u], 0, V - 1) : -1;
    if (vec) {
      const float4* x4 = reinterpret_cast<const float4*>(x);
      float4* g4 = reinterpret_cast<float4*>(g);
      for (int i = lane; i < (V >> 2); i += 32) {
        const float4 v = __ldcs(x4 + i);
        float4 o;
        o.x = exp2f(fmaf(v.x, RNNT_LOG2E, nl)) * c0;
        o.y = exp2f(fmaf(v.y, RNNT_LOG2E, nl)) * c0;
        o.z = exp2f(fmaf(v.z, RNNT_LOG2E, nl)) * c0;
        o.w = exp2f(fmaf(v.w, RNNT_LOG2E, nl)) * c0;
        const int v0 = i << 2;
        if ((unsigned)(blank - v0) < 4u) {
          const int k = blank - v0;
          if (k == 0) o.x -= cb; else if (k == 1) o.y -= cb; else if (k == 2) o.z -= cb; else o.w -= cb;
        }
        if ((unsigned)(y - v0) < 4u) {
          const int k = y - v0;
          if (k == 0) o.x -= cl; else if (k == 1) o.y -= cl; else if (k == 2) o.z -= cl; else o.w -= cl;
        }
        __stcs(g4 + i, o);
      }
    } else {
      for (int i = lane; i < V; i += 32) {
        float o = exp2f(fmaf(__ldcs(x + i), RNNT_LOG2E, nl)) * c0;
        if (i == blank) o -= cb;
        if (i == y) o -= cl;
        __stcs(g + i, o);
      }
    }
  }
}

// Small vocabularies: G lanes per cell, 32/G cells per warp (see logits_to_planes_small_kernel).
template <int G>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
planes_to_dlogits_small_kernel(const float* __restrict__ logits, const int32_t* __restrict__ labels,
                               const int32_t* __restrict__ act_lens, const int32_t* __restrict__ label_lens,
                               int B, int T, int U1, int V, int blank, FastDiv dTU, FastDiv dU1,
                               const float* __restrict__ lse_p, const float* __restrict__ cb_p,
                               const float* __restrict__ cl_p, const float* __restrict__ scale,
                               float* __restrict__ dlogits) {
  constexpr int CPW = 32 / G;
  const int lane = threadIdx.x & 31;
  const int sub = lane / G, gl = lane % G;
  const uint32_t ncell = (uint32_t)B * (uint32_t)T * (uint32_t)U1;
  const uint32_t stride = gridDim.x * WARPS_PER_BLOCK * CPW;
  const int U = U1 - 1;
  for (uint32_t c0 = (blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5)) * CPW; c0 < ncell; c0 += stride) {
    const uint32_t cell = c0 + sub;
    if (cell < ncell) {
      const int b = (int)fastdiv(cell, dTU);
      const uint32_t r = cell - (uint32_t)b * dTU.d;
      const int t = (int)fastdiv(r, dU1);
      const int u = (int)(r - (uint32_t)t * dU1.d);
      const int Tb = clamp_int(__ldg(act_lens + b), 1, T);
      const int Ub = clamp_int(__ldg(label_lens + b), 0, U);
      float* g = dlogits + (size_t)cell * V;
      if (t >= Tb || u > Ub) {  // outside the valid window: exact zeros
        for (int i = gl; i < V; i += G) __stcs(g + i, 0.f);
      } else {
        const float* x = logits + (size_t)cell * V;
        const size_t idx = plane_index(b, t, u, T, U1);
        const float sc = __ldg(scale + b);
        const float cb = __ldg(cb_p + idx) * sc, cl = __ldg(cl_p + idx) * sc;
        const float c0f = cb + cl;
        const float nl = -__ldg(lse_p + idx) * RNNT_LOG2E;
        const int y = (u < Ub) ? clamp_int(__ldg(labels + (size_t)b * U + u), 0, V - 1) : -1;
        for (int i = gl; i < V; i += G) {
          float o = exp2f(fmaf(__ldcs(x + i), RNNT_LOG2E, nl)) * c0f;
          if (i == blank) o -= cb;
          if (i == y) o -= cl;
          __stcs(g + i, o);
        }
      }
    }
    if (c0 > 0xffffffffu - stride) break;  // 32-bit wrap guard
  }
}

int grid_for_cells(long long ncell) {
  long long blocks = (ncell + WARPS_PER_BLOCK - 1) / WARPS_PER_BLOCK;
  const long long cap = 148LL * 8 * 4;  // 4 waves of 8 resident CTAs per SM
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace

extern "C" int rnnt_b200_logits_to_planes(const float* logits, const int32_t* labels,
                                          const int32_t* act_lens, const int32_t* label_lens,
                                          int B, int T, int U1, int V, int blank, float* lse,
                                          float* lp_blank, float* lp_label, void* stream) {
  if (!logits || !act_lens || !label_lens || !lse || !lp_blank || !lp_label) return RNNT_B200_E_NULL;
  if (U1 > 1 && !labels) return RNNT_B200_E_NULL;
  if (B <= 0 || T <= 0 || U1 <= 0 || V <= 0) return RNNT_B200_E_SHAPE;
  if (blank < 0 || blank >= V) return RNNT_B200_E_BLANK;
  if ((V & 3) == 0 && ((uintptr_t)logits & 15)) return RNNT_B200_E_ALIGN;
  if (const int arch = rnnt_check_arch()) return arch;
  const long long ncell = (long long)B * T * U1;
  if (V <= 256 && ncell < 0x7fffffffLL) {  // several cells per warp
    const FastDiv dTU = make_fastdiv((uint32_t)T * (uint32_t)U1), dU1 = make_fastdiv((uint32_t)U1);
    if (V <= 64)
      logits_to_planes_small_kernel<8><<<grid_for_cells((ncell + 3) / 4), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
          logits, labels, act_lens, label_lens, B, T, U1, V, blank, dTU, dU1, lse, lp_blank, lp_label);
    else
      logits_to_planes_small_kernel<16><<<grid_for_cells((ncell + 1) / 2), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
          logits, labels, act_lens, label_lens, B, T, U1, V, blank, dTU, dU1, lse, lp_blank, lp_label);
  } else {
    logits_to_planes_kernel<<<grid_for_cells(ncell), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
        logits, labels, act_lens, label_lens, B, T, U1, V, blank, lse, lp_blank, lp_label);
  }
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}

extern "C" int rnnt_b200_planes_to_dlogits(const float* logits, const int32_t* labels,
                                           const int32_t* act_lens, const int32_t* label_lens,
                                           int B, int T, int U1, int V, int blank, const float* lse,
                                           const float* c_blank, const float* c_label,
                                           const float* scale, float* dlogits, void* stream) {
  if (!logits || !act_lens || !label_lens || !lse || !c_blank || !c_label || !scale || !dlogits)
    return RNNT_B200_E_NULL;
  if (U1 > 1 && !labels) return RNNT_B200_E_NULL;
  if (B <= 0 || T <= 0 || U1 <= 0 || V <= 0) return RNNT_B200_E_SHAPE;
  if (blank < 0 || blank >= V) return RNNT_B200_E_BLANK;
  if ((V & 3) == 0 && (((uintptr_t)logits & 15) || ((uintptr_t)dlogits & 15))) return RNNT_B200_E_ALIGN;
  if (const int arch = rnnt_check_arch()) return arch;
  const long long ncell = (long long)B * T * U1;
  if (V <= 256 && ncell < 0x7fffffffLL) {  // several cells per warp
    const FastDiv dTU = make_fastdiv((uint32_t)T * (uint32_t)U1), dU1 = make_fastdiv((uint32_t)U1);
    if (V <= 64)
      planes_to_dlogits_small_kernel<8><<<grid_for_cells((ncell + 3) / 4), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
          logits, labels, act_lens, label_lens, B, T, U1, V, blank, dTU, dU1, lse, c_blank, c_label, scale, dlogits);
    else
      planes_to_dlogits_small_kernel<16><<<grid_for_cells((ncell + 1) / 2), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
          logits, labels, act_lens, label_lens, B, T, U1, V, blank, dTU, dU1, lse, c_blank, c_label, scale, dlogits);
  } else {
    planes_to_dlogits_kernel<<<grid_for_cells(ncell), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
        logits, labels, act_lens, label_lens, B, T, U1, V, blank, lse, c_blank, c_label, scale, dlogits);
  }
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}
