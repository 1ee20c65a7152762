// Small helper kernels of the fused path (boundary 2): tile table, vocabulary packing,
// and the ordered sum of split-K / per-chunk partial results.
#include <cuda_bf16.h>
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(1024)
build_tiles_kernel(const int32_t* __restrict__ act_lens, const int32_t* __restrict__ label_lens,
                   int B, int T, int U1, int tile_cap, int32_t* __restrict__ tile_utt,
                   int32_t* __restrict__ tile_cell0, int32_t* __restrict__ n_tiles) {
  int off = 0;
  for (int b = 0; b < B; ++b) {
    const int Tb = clamp_int(act_lens[b], 1, T);
    const int Ub = clamp_int(label_lens[b], 0, U1 - 1);
    const int ncell = Tb * (Ub + 1);
    const int nt = (ncell + RNNT_TILE_M - 1) / RNNT_TILE_M;
    for (int i = threadIdx.x; i < nt; i += blockDim.x) {
      if (off + i < tile_cap) {  // the tables hold tile_cap rows: a stale or foreign host hint must not overrun them
        tile_utt[off + i] = b;
        tile_cell0[off + i] = i * RNNT_TILE_M;
      }
    }
    off += nt;
  }
  // n_tiles[0]: tiles present in the tables; n_tiles[1]: tiles the lengths ask for (> tile_cap = overflow:
  // the caller's capacity was too small and the trailing tiles are NOT processed)
  if (threadIdx.x == 0) {
    n_tiles[0] = off < tile_cap ? off : tile_cap;
    n_tiles[1] = off;
  }
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

// w2t[j][v] = bf16(w2[v][j]) for v < V, 0 for V <= v < Vpad: the K-major B operand of the dh GEMM (contraction over v).
// 32 x 32 tiles through shared memory: coalesced reads along j, coalesced writes along v.
__global__ void __launch_bounds__(256)
pack_vocab_t_kernel(const float* __restrict__ w2, int V, int Vpad, int H, __nv_bfloat16* __restrict__ w2t) {
  __shared__ float tile[32][33];
  const int v0 = blockIdx.x * 32, j0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  for (int r = ty; r < 32; r += 8) {
    const int v = v0 + r, j = j0 + tx;
    tile[r][tx] = (v < V && j < H) ? w2[(size_t)v * H + j] : 0.f;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int j = j0 + r, v = v0 + tx;
    if (j < H && v < Vpad) w2t[(size_t)j * Vpad + v] = __float2bfloat16_rn(tile[tx][r]);
  }
}

// out[i] = parts[0][i] + parts[1][i] + ... in slot order: the deterministic end of a split-K GEMM
__global__ void __launch_bounds__(256)
sum_partials_kernel(const float* __restrict__ parts, int nparts, long long n, float* __restrict__ out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float a = parts[i];
    for (int s = 1; s < nparts; ++s) a += parts[(size_t)s * n + i];
    out[i] = a;
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
                                     int U1, int tile_cap, int32_t* tile_utt, int32_t* tile_cell0,
                                     int32_t* n_tiles, void* stream) {
  if (!act_lens || !label_lens || !tile_utt || !tile_cell0 || !n_tiles) return RNNT_B200_E_NULL;
  if (B <= 0 || T <= 0 || U1 <= 0 || rnnt_b200_max_tiles(B, T, U1) < 0) return RNNT_B200_E_SHAPE;
  if (tile_cap < 1 || tile_cap > rnnt_b200_max_tiles(B, T, U1)) return RNNT_B200_E_SHAPE;
  if (const int arch = rnnt_check_arch()) return arch;
  build_tiles_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(act_lens, label_lens, B, T, U1, tile_cap, tile_utt,
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
  {
    const int v = rnnt_tune_int("RNNT_B200_VGRAN", 64);  // timing experiments only: 16, 32, 64, ...
    if (v >= 16 && (v & (v - 1)) == 0) gran = v;
  }
  const int v16 = (V + gran - 1) & ~(gran - 1);
  const int pad = rnnt_b200_vocab_pad(V);
  return v16 < pad ? v16 : pad;
}

extern "C" int rnnt_b200_pack_vocab(const float* w2, const float* b2, int V, int H, void* w2_bf16,
                                    float* b2_log2, void* w2t_bf16, void* stream) {
  if (!w2 || !b2 || !w2_bf16 || !b2_log2) return RNNT_B200_E_NULL;
  if (V <= 0 || H <= 0) return RNNT_B200_E_SHAPE;
  const int Vpad = rnnt_b200_vocab_pad(V);
  const long long n = (long long)Vpad * H;
  int grid = (int)((n + 255) / 256);
  if (grid > 148 * 16) grid = 148 * 16;
  pack_vocab_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w2, b2, V, Vpad, H,
                                                           (__nv_bfloat16*)w2_bf16, b2_log2);
  RNNT_CUDA_RETURN_IF_ERROR();
  if (w2t_bf16) {
    pack_vocab_t_kernel<<<dim3((Vpad + 31) / 32, (H + 31) / 32), 256, 0, (cudaStream_t)stream>>>(
        w2, V, Vpad, H, (__nv_bfloat16*)w2t_bf16);
    RNNT_CUDA_RETURN_IF_ERROR();
  }
  return RNNT_B200_OK;
}

extern "C" int rnnt_b200_sum_partials(const float* parts, int nparts, long long n, float* out, void* stream) {
  if (!parts || !out) return RNNT_B200_E_NULL;
  if (nparts < 1 || n < 0) return RNNT_B200_E_SHAPE;
  if (n == 0) return RNNT_B200_OK;
  long long grid = (n + 255) / 256;
  if (grid > 148 * 16) grid = 148 * 16;
  sum_partials_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>(parts, nparts, n, out);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}
