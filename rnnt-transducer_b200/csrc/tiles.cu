// Small helper kernels of the fused path (boundary 2): tile table, vocabulary packing,
// and the dh -> (dP_enc, dP_dec) reduction (the backward of the broadcast add that
// replaces the reference's repeat+cat, /root/reference rnnt/model.py:26-31).
#include <cuda_bf16.h>
#include <stdlib.h>
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

// dpre = dh * (1 - h^2) and its two reductions. One CTA per (utterance b, block of DH_TT lattice
// rows t). A thread owns 8 consecutive hidden columns (16-byte bf16 loads), H/8 threads span a
// packed row and the 256/(H/8) row lanes of the CTA take u = lane, lane + lanes, ...  The u loop
// is outermost so that the sum over the block's t rows for dP_dec[b,u,:] lives in registers and
// costs two float4 reductions per (row lane, u); the DH_TT partial sums for dP_enc[b,t,:] stay in
// registers for the whole sweep and are reduced into memory at the end. All 2*DH_TT loads of an
// iteration are unconditional (masked cells are redirected to a row that is always inside the
// window) and issue back to back. Cells of an utterance are
// contiguous in the tile-packed row order, so row(t,u) = first_tile_b*128 + t*(U_b+1) + u; a launch
// covers the packed rows of tiles [tile_lo, tile_hi) only (the caller stages dS/dh in chunks),
// hence the atomics on dP_enc.
constexpr int DH_TT = 8;

__device__ __forceinline__ void unpack8(const uint4& w, float (&f)[8]) {
  const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ww[j]));
    f[2 * j] = v.x;
    f[2 * j + 1] = v.y;
  }
}

__global__ void __launch_bounds__(256, 2)
dh_reduce_kernel(const __nv_bfloat16* __restrict__ dh, const __nv_bfloat16* __restrict__ h,
                 const int32_t* __restrict__ act_lens, const int32_t* __restrict__ label_lens,
                 int T, int U1, int H, int tile_lo, int tile_hi, float dh_scale,
                 float* __restrict__ dpe, float* __restrict__ dpd) {
  const int b = blockIdx.y;
  const int Tb = clamp_int(act_lens[b], 1, T);
  const int w = clamp_int(label_lens[b], 0, U1 - 1) + 1;
  const int t0 = blockIdx.x * DH_TT;
  if (t0 >= Tb) return;
  int first_tile = 0;  // tiles of the utterances before b
  for (int i = 0; i < b; ++i) {
    const int n = clamp_int(act_lens[i], 1, T) * (clamp_int(label_lens[i], 0, U1 - 1) + 1);
    first_tile += (n + RNNT_TILE_M - 1) / RNNT_TILE_M;
  }
  // packed-row window of this launch, relative to the utterance's first cell
  const long long lo = (long long)(tile_lo - first_tile) * RNNT_TILE_M;
  const long long hi = (long long)(tile_hi - first_tile) * RNNT_TILE_M;
  const int nt = min(DH_TT, Tb - t0);
  const long long c_first = (long long)t0 * w, c_last = (long long)(t0 + nt) * w;
  if (c_last <= lo || c_first >= hi) return;
  const int cpr = H / 8;              // 16-byte chunks per row
  const int lanes = 256 / cpr;        // row lanes of the CTA
  const int chunk = threadIdx.x % cpr;
  const int rl = threadIdx.x / cpr;
  if (rl >= lanes) return;
  const int hp = rnnt_hp(H);
  const uint4* dh4 = reinterpret_cast<const uint4*>(dh) + chunk;                // row pitch H/8 uint4
  const uint4* h4 = reinterpret_cast<const uint4*>(h + (size_t)first_tile * RNNT_TILE_M * hp) + chunk;  // pitch hp/8
  float acc_e[DH_TT][8];
#pragma unroll
  for (int i = 0; i < DH_TT; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc_e[i][j] = 0.f;
  // a row that is always inside the window
  const long long c_safe = max(c_first, lo);
  for (int u = rl; u < w; u += lanes) {
    float acc_d[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc_d[j] = 0.f;
#pragma unroll
    for (int i0 = 0; i0 < DH_TT; i0 += 4) {  // two batches of 4 rows: 8 x 16 bytes in flight per thread
      uint4 g[4], hv[4];
      bool ok[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const long long c = (long long)(t0 + i0 + i) * w + u;
        ok[i] = (i0 + i < nt) && c >= lo && c < hi;
        const long long cc = ok[i] ? c : c_safe;
        g[i] = __ldg(dh4 + (size_t)(cc - lo) * (H / 8));
        hv[i] = __ldg(h4 + (size_t)cc * (hp / 8));
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float gf[8], hf[8];
        unpack8(g[i], gf);
        unpack8(hv[i], hf);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float v = ok[i] ? gf[j] * (1.f - hf[j] * hf[j]) : 0.f;
          acc_d[j] += v;
          acc_e[i0 + i][j] += v;
        }
      }
    }
    float4* dst = reinterpret_cast<float4*>(dpd + ((size_t)b * U1 + u) * H + chunk * 8);
    atomicAdd(dst, make_float4(acc_d[0] * dh_scale, acc_d[1] * dh_scale, acc_d[2] * dh_scale, acc_d[3] * dh_scale));
    atomicAdd(dst + 1, make_float4(acc_d[4] * dh_scale, acc_d[5] * dh_scale, acc_d[6] * dh_scale, acc_d[7] * dh_scale));
  }
#pragma unroll
  for (int i = 0; i < DH_TT; ++i)
    if (i < nt) {
      float4* dst = reinterpret_cast<float4*>(dpe + ((size_t)b * T + t0 + i) * H + chunk * 8);
      atomicAdd(dst, make_float4(acc_e[i][0] * dh_scale, acc_e[i][1] * dh_scale, acc_e[i][2] * dh_scale, acc_e[i][3] * dh_scale));
      atomicAdd(dst + 1, make_float4(acc_e[i][4] * dh_scale, acc_e[i][5] * dh_scale, acc_e[i][6] * dh_scale, acc_e[i][7] * dh_scale));
    }
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
extern "C" int rnnt_b200_h_pitch(int H) { return rnnt_hp(H); }
extern "C" int rnnt_b200_ds_pitch(int V) {
  // The UMMA N granularity is 16, but a pitch that is not a multiple of 64 elements (128 bytes)
  // slows the library GEMMs that consume dS by more than the narrower tile saves (measured:
  // profiles/r01_notes.md), so the last vocabulary tile is cut at the next multiple of 64.
  int gran = 64;
  if (const char* g = getenv("RNNT_B200_VGRAN")) {  // timing experiments only: 16, 32, 64, ...
    const int v = atoi(g);
    if (v >= 16 && (v & (v - 1)) == 0) gran = v;
  }
  const int v16 = (V + gran - 1) & ~(gran - 1);
  const int pad = rnnt_b200_vocab_pad(V);
  return v16 < pad ? v16 : pad;
}

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

extern "C" int rnnt_b200_dh_reduce(const void* dh_bf16, const void* h_stash, const int32_t* act_lens,
                                   const int32_t* label_lens, int B, int T, int U1, int H, int tile_lo,
                                   int tile_hi, float dh_scale, float* dp_enc, float* dp_dec,
                                   void* stream) {
  if (!dh_bf16 || !h_stash || !act_lens || !label_lens || !dp_enc || !dp_dec) return RNNT_B200_E_NULL;
  if (B <= 0 || B > 65535 || T <= 0 || U1 <= 0 || H < 64 || H > 512 || (H & 63) || tile_lo < 0 ||
      tile_hi < tile_lo)
    return RNNT_B200_E_SHAPE;
  if (tile_hi == tile_lo) return RNNT_B200_OK;
  dh_reduce_kernel<<<dim3((T + DH_TT - 1) / DH_TT, B), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dh_bf16, (const __nv_bfloat16*)h_stash, act_lens, label_lens, T, U1, H, tile_lo,
      tile_hi, dh_scale, dp_enc, dp_dec);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}
