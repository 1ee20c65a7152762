// Greedy-decoding step of the joint network: for B utterances at
// once, pred[b] = argmax_v ( W2 . tanh(P_enc[b,t] + P_dec[b]) + b2 )[v] -- what rnnt/model.py:90-92
// computes one utterance and one frame at a time (joint on two 1-D vectors, softmax, argmax,
// .item()). fp32 throughout, no logits written anywhere, nothing read back by the host.
//
// Grid: one CTA per 32 vocabulary rows (V=4232: 133 CTAs, one wave). Every CTA rebuilds the B x H
// hidden rows in shared memory (B*H tanhf, cheap next to re-reading them from HBM), each warp
// takes 4 vocabulary rows: the row of W2 lives in registers (H/32 values per lane), the B dot
// products run against shared memory, warp-shuffle reduced. Per-utterance maxima are merged as
// 64-bit keys (orderable logit bits << 32 | ~index: ties resolve to the LOWEST index, as
// torch.argmax does) with shared- then global-memory atomicMax; the last CTA to finish turns the
// keys into int64 predictions and leaves the work area zeroed for the next step.
#include "common.cuh"

namespace {

constexpr int DEC_THREADS = 256;
constexpr int DEC_ROWS = 32;  // vocabulary rows per CTA

__device__ __forceinline__ unsigned long long argmax_key(float v, int idx) {
  const uint32_t u = __float_as_uint(v);
  const uint32_t o = (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // monotone float -> uint
  return ((unsigned long long)o << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)idx);
}

template <int KPL>  // H / 32 values of a W2 row per lane
__global__ void __launch_bounds__(DEC_THREADS)
joint_step_argmax_kernel(const float* __restrict__ pe_t, long long pe_stride, const float* __restrict__ pd,
                         const float* __restrict__ w2, const float* __restrict__ b2, int B, int V,
                         unsigned long long* __restrict__ work, long long* __restrict__ pred) {
  constexpr int H = KPL * 32;
  extern __shared__ __align__(16) uint8_t dec_smem[];
  float* hs = reinterpret_cast<float*>(dec_smem);                                   // B x H
  unsigned long long* best = reinterpret_cast<unsigned long long*>(hs + (size_t)B * H);  // B
  __shared__ int is_last;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < B * H; i += DEC_THREADS) {
    const int b = i / H, k = i - b * H;
    hs[i] = tanhf(pe_t[(size_t)b * pe_stride + k] + pd[(size_t)b * H + k]);
  }
  for (int b = threadIdx.x; b < B; b += DEC_THREADS) best[b] = 0ull;
  __syncthreads();
  const int v0 = blockIdx.x * DEC_ROWS;
  for (int r = warp; r < DEC_ROWS; r += DEC_THREADS / 32) {
    const int v = v0 + r;
    if (v >= V) break;
    float w[KPL];
#pragma unroll
    for (int j = 0; j < KPL; ++j) w[j] = __ldg(w2 + (size_t)v * H + lane + 32 * j);
    const float bias = __ldg(b2 + v);
    for (int b = 0; b < B; ++b) {
      const float* hb = hs + (size_t)b * H + lane;
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < KPL; ++j) acc = fmaf(w[j], hb[32 * j], acc);
#pragma unroll
      for (int sh = 16; sh > 0; sh >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, sh);
      if (lane == 0) atomicMax(best + b, argmax_key(acc + bias, v));
    }
  }
  __syncthreads();
  for (int b = threadIdx.x; b < B; b += DEC_THREADS) atomicMax(work + 1 + b, best[b]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(work, 1ull) == (unsigned long long)(gridDim.x - 1));
  __syncthreads();
  if (is_last) {
    __threadfence();
    for (int b = threadIdx.x; b < B; b += DEC_THREADS) {
      const unsigned long long key = atomicExch(work + 1 + b, 0ull);
      pred[b] = (long long)(0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFull));
    }
    if (threadIdx.x == 0) work[0] = 0ull;
  }
}

template <int KPL>
int launch_step(const float* pe_t, long long pe_stride, const float* pd, const float* w2, const float* b2,
                int B, int V, unsigned long long* work, long long* pred, cudaStream_t st) {
  const size_t smem = (size_t)B * KPL * 32 * sizeof(float) + (size_t)B * sizeof(unsigned long long);
  auto kern = joint_step_argmax_kernel<KPL>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  kern<<<(V + DEC_ROWS - 1) / DEC_ROWS, DEC_THREADS, smem, st>>>(pe_t, pe_stride, pd, w2, b2, B, V, work, pred);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}

}  // namespace

extern "C" int rnnt_b200_joint_step_argmax(const float* pe_t, long long pe_stride, const float* pd,
                                           const float* w2, const float* b2, int B, int H, int V,
                                           unsigned long long* work, long long* pred, void* stream) {
  if (!pe_t || !pd || !w2 || !b2 || !work || !pred) return RNNT_B200_E_NULL;
  if (B <= 0 || B > 64 || V <= 0 || H < 64 || H > 512 || (H & 63) || pe_stride < H) return RNNT_B200_E_SHAPE;
  cudaStream_t st = (cudaStream_t)stream;
  switch (H / 32) {
    case 2: return launch_step<2>(pe_t, pe_stride, pd, w2, b2, B, V, work, pred, st);
    case 4: return launch_step<4>(pe_t, pe_stride, pd, w2, b2, B, V, work, pred, st);
    case 6: return launch_step<6>(pe_t, pe_stride, pd, w2, b2, B, V, work, pred, st);
    case 8: return launch_step<8>(pe_t, pe_stride, pd, w2, b2, B, V, work, pred, st);
    case 10: return launch_step<10>(pe_t, pe_stride, pd, w2, b2, B, V, work, pred, st);
    case 12: return launch_step<12>(pe_t, pe_stride, pd, w2, b2, B, V, work, pred, st);
    case 14: return launch_step<14>(pe_t, pe_stride, pd, w2, b2, B, V, work, pred, st);
    case 16: return launch_step<16>(pe_t, pe_stride, pd, w2, b2, B, V, work, pred, st);
  }
  return RNNT_B200_E_SHAPE;
}
