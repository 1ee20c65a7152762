// Small helper kernels of the fused path (boundary 2): tile table, vocabulary packing,
// and the dh -> (dP_enc, dP_dec) reduction (the backward of the broadcast add that
// replaces the reference's repeat+cat, /root/reference rnnt/model.py:26-31).
#include <cuda_bf16.h>
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(1024)
build_tiles_kernel(const int32_t* __restrict__ act_lens, const int32_t* __restrict__ label_lens,
                   int B, int T, int U1, int32_t* __restrict__ tile_utt,
                   int32_t* __restrict__ tile_cell0, int32_t* __restrict__ n_tiles) {
  int off = 0;
  for (int b = 0; b < B; ++b) {
    const int Tb = clamp_int(act_lens[b], 1, T);
    const int Ub = clamp_int(label_lens[b], 0, U1 - 1);
    const int ncell = Tb * (Ub + 1);
    const int nt = (ncell + RNNT_TILE_M - 1) / RNNT_TILE_M;
    for (int i = threadIdx.x; i < nt; i += blockDim.x) {
      tile_utt[off + i] = b;
      tile_cell0[off + i] = i * RNNT_TILE_M;
    }
    off += nt;
  }
  if (threadIdx.x == 0) n_tiles[0] = off;
}

__global__ void __launch_bounds__(256)
pack_vocab_kernel(const float* __restrict__ w2, const float* __restrict__ b2, int V, int Vpad, int H,
                  __nv_bfloat16* __restrict__ w2q, float* __restrict__ b2l) {
  const long long n = (long long)Vpad * H;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i / H);
    w2q[i] = __float2bfloat16_rn(v < V ? w2[i] : 0.f);
    if (i < Vpad) b2l[i] = (i < V) ? b2[i] * RNNT_LOG2E : RNNT_NEG_BIG;
  }
}

// One CTA per tile, one thread per hidden column k. Rows of a tile are consecutive
// cells (t,u) of one utterance: runs of equal t are pre-reduced in a register before
// one atomic into dP_enc; dP_dec takes one coalesced red per row.
__global__ void __launch_bounds__(512)
dh_reduce_kernel(const float* __restrict__ dh, const __nv_bfloat16* __restrict__ h,
                 const int32_t* __restrict__ act_lens, const int32_t* __restrict__ label_lens,
                 const int32_t* __restrict__ tile_utt, const int32_t* __restrict__ tile_cell0,
                 const int32_t* __restrict__ n_tiles, int T, int U1, int H, int tile_lo,
                 float* __restrict__ dpe, float* __restrict__ dpd) {
  const int tile = tile_lo + blockIdx.x;
  if (tile >= n_tiles[0]) return;
  const int k = threadIdx.x;
  const int b = tile_utt[tile];
  const int cell0 = tile_cell0[tile];
  const int Tb = clamp_int(act_lens[b], 1, T);
  const int Ub1 = clamp_int(label_lens[b], 0, U1 - 1) + 1;
  const int ncell = Tb * Ub1;
  const int rows = min(RNNT_TILE_M, ncell - cell0);
  const int hp = H + 8;
  const float* dhp = dh + (size_t)blockIdx.x * RNNT_TILE_M * H + k;
  const __nv_bfloat16* hpn = h + (size_t)tile * RNNT_TILE_M * hp + k;
  int t = cell0 / Ub1, u = cell0 - t * Ub1;
  float acc_e = 0.f;
  for (int r = 0; r < rows; ++r) {
    const float hv = __bfloat162float(hpn[(size_t)r * hp]);
    const float g = dhp[(size_t)r * H] * (1.f - hv * hv);
    acc_e += g;
    atomicAdd(dpd + ((size_t)b * U1 + u) * H + k, g);
    if (++u == Ub1) {
      atomicAdd(dpe + ((size_t)b * T + t) * H + k, acc_e);
      acc_e = 0.f; u = 0; ++t;
    }
  }
  if (u != 0) atomicAdd(dpe + ((size_t)b * T + t) * H + k, acc_e);
}

}  // namespace

extern "C" int rnnt_b200_max_tiles(int B, int T, int U1) {
  if (B <= 0 || T <= 0 || U1 <= 0) return 0;
  const long long per = ((long long)T * U1 + RNNT_TILE_M - 1) / RNNT_TILE_M;
  const long long tot = per * B;
  return tot > 0x7fffffffLL ? -1 : (int)tot;
}

extern "C" int rnnt_b200_build_tiles(const int32_t* act_lens, const int32_t* label_lens, int B, int T,
                                     int U1, int32_t* tile_utt, int32_t* tile_cell0,
                                     int32_t* n_tiles, void* stream) {
  if (!act_lens || !label_lens || !tile_utt || !tile_cell0 || !n_tiles) return RNNT_B200_E_NULL;
  if (B <= 0 || T <= 0 || U1 <= 0 || rnnt_b200_max_tiles(B, T, U1) < 0) return RNNT_B200_E_SHAPE;
  build_tiles_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(act_lens, label_lens, B, T, U1, tile_utt,
                                                         tile_cell0, n_tiles);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}

extern "C" int rnnt_b200_vocab_tile(int V) { return V <= 64 ? 64 : (V <= 128 ? 128 : 256); }
extern "C" int rnnt_b200_vocab_pad(int V) {
  const int bn = rnnt_b200_vocab_tile(V);
  return ((V + bn - 1) / bn) * bn;
}
extern "C" int rnnt_b200_h_pitch(int H) { return H + 8; }

extern "C" int rnnt_b200_pack_vocab(const float* w2, const float* b2, int V, int H, void* w2_bf16,
                                    float* b2_log2, void* stream) {
  if (!w2 || !b2 || !w2_bf16 || !b2_log2) return RNNT_B200_E_NULL;
  if (V <= 0 || H <= 0) return RNNT_B200_E_SHAPE;
  const int Vpad = rnnt_b200_vocab_pad(V);
  const long long n = (long long)Vpad * H;
  int grid = (int)((n + 255) / 256);
  if (grid > 148 * 16) grid = 148 * 16;
  pack_vocab_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w2, b2, V, Vpad, H,
                                                           (__nv_bfloat16*)w2_bf16, b2_log2);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}

extern "C" int rnnt_b200_dh_reduce(const float* dh, const void* h_stash, const int32_t* act_lens,
                                   const int32_t* label_lens, const int32_t* tile_utt,
                                   const int32_t* tile_cell0, const int32_t* n_tiles, int B, int T,
                                   int U1, int H, int tile_lo, int tile_hi, float* dp_enc,
                                   float* dp_dec, void* stream) {
  if (!dh || !h_stash || !act_lens || !label_lens || !tile_utt || !tile_cell0 || !n_tiles || !dp_enc ||
      !dp_dec)
    return RNNT_B200_E_NULL;
  if (B <= 0 || T <= 0 || U1 <= 0 || H < 64 || H > 512 || (H & 63) || tile_lo < 0 || tile_hi < tile_lo)
    return RNNT_B200_E_SHAPE;
  if (tile_hi == tile_lo) return RNNT_B200_OK;
  dh_reduce_kernel<<<tile_hi - tile_lo, H, 0, (cudaStream_t)stream>>>(
      dh, (const __nv_bfloat16*)h_stash, act_lens, label_lens, tile_utt, tile_cell0, n_tiles, T, U1, H,
      tile_lo, dp_enc, dp_dec);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}
