// Feature front end of a batch: Kaldi matrix payloads as stored in
// the ark file -> (CM decompression) -> CMVN -> context splicing -> subsampling -> zero padding,
// for all utterances of a batch in ONE launch. It replaces, per utterance, rnnt/kaldi_io.py:467-554
// (matrix decode) and rnnt/dataset.py:72-75 (cmvn), :77-98 (concat_frame), :100-109 (subsampling),
// :38-56 (pad), which the reference runs in numpy inside DataLoader workers.
//
// Byte/float streaming work, HBM-bound by construction and tiny (AISHELL batch of 32: ~8 MB in,
// ~10 MB out): one thread per VEC consecutive features of one output block, consecutive threads on
// consecutive addresses of the output (coalesced 16-byte stores; float32 payloads are read with
// coalesced 16-byte loads too), grid-stride over a grid sized from the SM count. Every arithmetic
// step uses the rounding numpy uses (IEEE sub/div/mul/add, no contraction into FMA), so the result
// is bit-identical to the numpy path.
#include "common.cuh"

namespace {

constexpr int FEAT_THREADS = 256;
constexpr int FEAT_MAX_BLK = 32;

struct FeatParams {
  const uint8_t* payload;
  const rnnt_b200_utt_desc* desc;
  const float* col_params;
  const void* mean;
  const void* sd;
  float* out;
  int B, F, nblk, interval, t_pad, cmvn_f64;
  int prim[FEAT_MAX_BLK], over[FEAT_MAX_BLK];
};

// One value of a Kaldi compressed matrix (kaldi_io.py:540-549): three linear segments between the
// column's percentiles; products and sums rounded separately, "+ 0" as the masked sum leaves it.
__device__ __forceinline__ float cm_value(const float* __restrict__ cp, unsigned d) {
  float v;
  if (d <= 64u) v = __fadd_rn(cp[0], __fmul_rn(cp[1], (float)d));
  else if (d <= 192u) v = __fadd_rn(cp[2], __fmul_rn(cp[3], (float)(d - 64u)));
  else v = __fadd_rn(cp[4], __fmul_rn(cp[5], (float)(d - 192u)));
  return __fadd_rn(v, 0.0f);
}

__device__ __forceinline__ float cmvn_value(const FeatParams& p, float x, int spk, int f) {
  if (spk < 0) return x;
  const size_t i = (size_t)spk * p.F + f;
  if (p.cmvn_f64) {
    const double m = reinterpret_cast<const double*>(p.mean)[i], s = reinterpret_cast<const double*>(p.sd)[i];
    return (float)(((double)x - m) / s);
  }
  return __fdiv_rn(__fsub_rn(x, reinterpret_cast<const float*>(p.mean)[i]), reinterpret_cast<const float*>(p.sd)[i]);
}

__device__ __forceinline__ float cmvn_value_f64(const FeatParams& p, double x, int spk, int f) {
  if (spk < 0) return (float)x;
  const size_t i = (size_t)spk * p.F + f;
  return (float)((x - reinterpret_cast<const double*>(p.mean)[i]) / reinterpret_cast<const double*>(p.sd)[i]);
}

template <int VEC>
__global__ void __launch_bounds__(FEAT_THREADS) features_batch_kernel(const FeatParams p) {
  const int fg = p.F / VEC;  // feature groups per block
  const long long per_frame = (long long)p.nblk * fg;
  const long long total = (long long)p.B * p.t_pad * per_frame;
  for (long long idx = (long long)blockIdx.x * FEAT_THREADS + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * FEAT_THREADS) {
    const long long frame = idx / per_frame;
    const int within = (int)(idx - frame * per_frame);
    const int j = within / fg, f0 = (within - j * fg) * VEC;
    const int b = (int)(frame / p.t_pad), t = (int)(frame - (long long)b * p.t_pad);
    const rnnt_b200_utt_desc d = p.desc[b];
    float v[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) v[k] = 0.f;
    int src = -1;
    if (t < d.out_rows) {
      const long long centre = (long long)t * p.interval;
      if (p.over[j] != RNNT_B200_FEAT_NO_SHIFT) {
        const long long r = centre + p.over[j];
        if (r >= 0 && r < d.rows) src = (int)r;
      }
      if (src < 0 && p.prim[j] != RNNT_B200_FEAT_NO_SHIFT) {
        const long long r = centre + p.prim[j];
        if (r >= 0 && r < d.rows) src = (int)r;
      }
    }
    if (src >= 0) {
      const uint8_t* base = p.payload + d.payload_off;
      if (d.format == RNNT_B200_FEAT_F32) {
        const float* row = reinterpret_cast<const float*>(base) + (size_t)src * p.F + f0;
        if (VEC == 4) {
          const float4 x = __ldg(reinterpret_cast<const float4*>(row));
          v[0] = x.x; v[1 % VEC] = x.y; v[2 % VEC] = x.z; v[3 % VEC] = x.w;
        } else {
#pragma unroll
          for (int k = 0; k < VEC; ++k) v[k] = __ldg(row + k);
        }
#pragma unroll
        for (int k = 0; k < VEC; ++k) v[k] = cmvn_value(p, v[k], d.spk, f0 + k);
      } else if (d.format == RNNT_B200_FEAT_CM8) {
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
          const unsigned byte = __ldg(base + (size_t)(f0 + k) * d.rows + src);
          v[k] = cmvn_value(p, cm_value(p.col_params + d.col_off + 6 * (f0 + k), byte), d.spk, f0 + k);
        }
      } else {
        const double* row = reinterpret_cast<const double*>(base) + (size_t)src * p.F + f0;
#pragma unroll
        for (int k = 0; k < VEC; ++k) v[k] = cmvn_value_f64(p, __ldg(row + k), d.spk, f0 + k);
      }
    }
    float* o = p.out + (size_t)frame * p.nblk * p.F + (size_t)j * p.F + f0;
    if (VEC == 4) {
      *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1 % VEC], v[2 % VEC], v[3 % VEC]);
    } else {
#pragma unroll
      for (int k = 0; k < VEC; ++k) o[k] = v[k];
    }
  }
}

}  // namespace

extern "C" int rnnt_b200_features_batch(const void* payload, const rnnt_b200_utt_desc* desc,
                                        const float* col_params, const void* cmvn_mean, const void* cmvn_sd,
                                        int cmvn_f64, int B, int F, int nblk, const int* blk_primary,
                                        const int* blk_override, int interval, int t_pad, float* out,
                                        void* stream) {
  if (!payload || !desc || !blk_primary || !blk_override || !out) return RNNT_B200_E_NULL;
  if (B <= 0 || F <= 0 || nblk <= 0 || nblk > FEAT_MAX_BLK || interval <= 0 || t_pad <= 0) return RNNT_B200_E_SHAPE;
  FeatParams p;
  p.payload = static_cast<const uint8_t*>(payload);
  p.desc = desc;
  p.col_params = col_params;
  p.mean = cmvn_mean;
  p.sd = cmvn_sd;
  p.out = out;
  p.B = B; p.F = F; p.nblk = nblk; p.interval = interval; p.t_pad = t_pad; p.cmvn_f64 = cmvn_f64;
  for (int j = 0; j < FEAT_MAX_BLK; ++j) {
    p.prim[j] = j < nblk ? blk_primary[j] : RNNT_B200_FEAT_NO_SHIFT;
    p.over[j] = j < nblk ? blk_override[j] : RNNT_B200_FEAT_NO_SHIFT;
  }
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const bool vec4 = (F % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0) &&
                    ((reinterpret_cast<uintptr_t>(payload) & 15) == 0);
  const long long items = (long long)B * t_pad * nblk * (vec4 ? F / 4 : F);
  long long blocks = (items + FEAT_THREADS - 1) / FEAT_THREADS;
  const long long cap = (long long)sms * 8;  // 8 resident CTAs of 256 threads per SM
  if (blocks > cap) blocks = cap;
  cudaStream_t st = (cudaStream_t)stream;
  if (vec4) features_batch_kernel<4><<<(unsigned)blocks, FEAT_THREADS, 0, st>>>(p);
  else features_batch_kernel<1><<<(unsigned)blocks, FEAT_THREADS, 0, st>>>(p);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}
