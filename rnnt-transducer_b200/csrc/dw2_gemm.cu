// [dW2 | db2] += dS^T . [h | 1] for one staged dS chunk -- the weight/bias gradient of
// project_layer (rnnt/model.py:16,35; the reference gets it from autograd through the
// materialised logits). The contraction runs over the chunk's ROWS (cells), so both operands are
// "MN-major" for the tensor core: dS (rows x Vds) supplies A[m = v][k = row], the h stash
// (rows x pitch) supplies B[n = j][k = row]. The library does this GEMM at N = H + 32 (the ones
// column that yields db2 rides in the h stash) and loses ~25 % to that shape; here db2 is one extra
// UMMA of N = 16 against a constant tile of ones, accumulated in 16 spare TMEM columns.
//
// Work unit = (K split s, vocabulary tile of 256 rows of dW2, half j of the H = 512 columns): one
// CTA pair (cta_group::2, UMMA 256 x 256 x 16), fp32 accumulator in 256 TMEM columns per CTA.
// Per 64-row K block and CTA: two 64(v) x 64(row) and two 64(j) x 64(row) TMA boxes (SWIZZLE_128B:
// a box lands as 64 rows of 128 bytes = eight 8-row swizzle atoms, exactly the canonical MN-major
// layout ((8,n),(8,k)) : ((1,LBO),(8,SBO)) in 16-byte units with SBO = 1024 B between 8-row groups
// and LBO = 8192 B between the two 64-wide blocks). The two units (j halves) of a (split, tile)
// run on neighbouring pairs at the same time, so the second read of dS hits L2.
// Warps: 0 dS TMA producer, 6 h TMA producer, 1 MMA issuer (+ TMEM), 2-5 epilogue (TMEM ->
// red.global.add.v4.f32).
// H = 512 runs dw2_wide_kernel further down instead (256 x 512 tile per pair, collector reuse of the
// dS^T slice, db2 by the epilogue warps); dw2_gemm_kernel serves H = 128 and 256 and the A/B switch.
#include <cuda.h>
#include <cuda_bf16.h>
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "tma_host.cuh"

using namespace sm100;

namespace {

constexpr int GW_THREADS = 224;  // warps: 0 dS TMA, 1 MMA, 2-5 epilogue, 6 h TMA
constexpr int GW_STAGES = 7;
constexpr int GW_BLOCK = 64 * 128;           // one 64 x 64 bf16 box
constexpr int GW_STAGE = 4 * GW_BLOCK;       // A: 2 boxes (128 v), B: 2 boxes (128 j)
constexpr int GW_ONES = 16 * 128;            // 16 k-rows x 128 B of bf16 1.0
constexpr int GW_SMEM = GW_STAGES * GW_STAGE + GW_ONES + 256;

struct GwCtl {
  uint64_t full[GW_STAGES], empty[GW_STAGES], acc_full, acc_empty;
  uint32_t tmem_base;
};

struct GwParams {
  float* dw2;      // partial s at dw2 + s * Vds * H: (Vds, H) fp32, every element stored exactly once
  float* db2;      // partial s at db2 + s * Vds: (Vds) fp32 (dw2_gemm_kernel adds into it: zeroed by the caller)
  int Vds, H;
  int nvt, nj;     // vocabulary tiles of 256, column units of N
  int N;           // UMMA N of a unit: 256 (H a multiple of 256) or 128 (H = 128)
  int ksteps;      // 64-row K blocks in the chunk
  int splits, ks_per_split;
  int interleave;  // dw2_wide_kernel: K split s takes the 64-row blocks s, s + splits, ... instead of a contiguous range
  const float* row_weight;  // dw2_wide_kernel, optional: db2 = sum over rows of row_weight[row] * dS[row, :]
  int debug;  // RNNT_B200_DW2_DEBUG (timing experiments): 1 no ones-MMA, 2 no main MMA, 4 no epilogue adds, 8 no TMA loads, 16 no A reuse through the collector
};

__global__ void __launch_bounds__(GW_THREADS, 1)
dw2_gemm_kernel(const __grid_constant__ CUtensorMap map_ds, const __grid_constant__ CUtensorMap map_h,
                const GwParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* ones = smem + GW_STAGES * GW_STAGE;
  GwCtl* ctl = reinterpret_cast<GwCtl*>(ones + GW_ONES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  // All units of a K split walk the chunk's rows in step: the nj column halves of a vocabulary
  // tile share their dS boxes and all vocabulary tiles share the h boxes through L2. (A stream-K
  // split that gave every pair an equal share of the (tile, K) space but different K offsets ran
  // 25 % slower: h was re-read from HBM once per vocabulary tile.)
  const int cid = blockIdx.x >> 1, ncl = gridDim.x >> 1;
  const int n_units = p.splits * p.nvt * p.nj;

  for (int i = threadIdx.x; i < GW_ONES / 4; i += GW_THREADS) reinterpret_cast<uint32_t*>(ones)[i] = 0x3F803F80u;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_ds);
    tma_prefetch_desc(&map_h);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < GW_STAGES; ++i) {
      mbar_init(&ctl->full[i], 4);  // {dS, h} producer x {leader, peer}
      mbar_init(&ctl->empty[i], 1);
    }
    mbar_init(&ctl->acc_full, 1);
    mbar_init(&ctl->acc_empty, 2 * 4);
    fence_barrier_init();
  }
  fence_proxy_async_smem();  // the ones tile is read by the tensor core (async proxy)
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tmem_alloc_pair(&ctl->tmem_base, 512);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp == 0 || warp == 6) {
    // =========================== TMA PRODUCERS ===========================
    // One thread per operand (a single producer needed longer for its wait + expect_tx + four
    // TMA issues than the 512 clk the four MMAs of a stage take).
    const bool is_a = warp == 0;
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int u = cid; u < n_units; u += ncl) {
        const int jh = u % p.nj, vt = (u / p.nj) % p.nvt, s = u / (p.nj * p.nvt);
        const int k0 = s * p.ks_per_split, k1 = min(p.ksteps, k0 + p.ks_per_split);
        const int v0 = vt * 256 + (int)rank * 128, j0 = jh * p.N + (int)rank * (p.N / 2);
        for (int k = k0; k < k1; ++k) {
          mbar_wait(&ctl->empty[stage], phase ^ 1);
          const uint32_t full = mapa_u32(smem_u32(&ctl->full[stage]), 0);
          if (p.debug & 8) {  // no loads: the MMAs run on whatever the buffers hold
            mbar_arrive_cluster(full);
            if (++stage == GW_STAGES) { stage = 0; phase ^= 1; }
            continue;
          }
          const int nbox = is_a ? 2 : p.N / 128;  // 64-wide boxes of this operand per CTA
          if (rank == 0) mbar_arrive_expect_tx(&ctl->full[stage], 2 * nbox * GW_BLOCK);  // both CTAs
          else mbar_arrive_cluster(full);
          uint8_t* st = smem + stage * GW_STAGE + (is_a ? 0 : 2 * GW_BLOCK);
          const CUtensorMap* map = is_a ? &map_ds : &map_h;
          const int c0 = is_a ? v0 : j0;
          tma_load_2d_pair(st, map, full, c0, k * 64);
          if (nbox == 2) tma_load_2d_pair(st + GW_BLOCK, map, full, c0 + 64, k * 64);
          if (++stage == GW_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA ISSUER (leader CTA) ===========================
    if (rank == 0) {
      const uint32_t idesc_main = umma_idesc_bf16_mn(256, p.N);
      constexpr uint32_t idesc_ones = umma_idesc_bf16_mn(256, 16);
      const uint32_t s0 = smem_u32(smem);
      const uint64_t ones_desc = umma_desc_mn128(smem_u32(ones), GW_BLOCK);
      uint32_t stage = 0, phase = 0;
      int it = 0;
      for (int u = cid; u < n_units; u += ncl, ++it) {
        const int jh = u % p.nj, s = u / (p.nj * p.nvt);
        const int k0 = s * p.ks_per_split, k1 = min(p.ksteps, k0 + p.ks_per_split);
        mbar_wait(&ctl->acc_empty, (it & 1) ^ 1);
        tc_fence_after();
        // The db2 UMMAs are dealt to the nj column halves by K-block parity: the partner units
        // stay in step (the one that did all of them fell ~6 % behind, ~100 MB of dS by the end
        // of a chunk -- more than L2 holds -- and re-read its dS boxes from HBM).
        const int k_ones0 = k0 + ((p.nj == 2 && ((k0 + jh) & 1)) ? 1 : 0);
        for (int k = k0; k < k1; ++k) {
          mbar_wait(&ctl->full[stage], phase);
          tc_fence_after();
          if (elect_one_sync()) {
            const uint32_t a0 = s0 + stage * GW_STAGE, b0 = a0 + 2 * GW_BLOCK;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {  // 16 rows = two 8-row swizzle atoms = 2048 bytes
              const uint64_t adesc = umma_desc_mn128(a0 + kk * 2048, GW_BLOCK);
              const uint64_t bdesc = umma_desc_mn128(b0 + kk * 2048, GW_BLOCK);
              const uint32_t acc = (k > k0 || kk > 0) ? 1u : 0u;
              const bool with_ones = (p.nj == 1 || ((k + jh) & 1) == 0) && !(p.debug & 1);
              // the ones-UMMA multiplies the same A slice: kept in the collector by the main UMMA, so
              // shared memory is read for it once (without this the db2 UMMAs added 25 % to the A reads)
              if (with_ones && !(p.debug & (2 | 16))) {
                umma_bf16_ss_pair_areuse<1>(tmem_base, adesc, bdesc, idesc_main, acc);
                umma_bf16_ss_pair_areuse<0>(tmem_base + 256, adesc, ones_desc, idesc_ones, (k > k_ones0 || kk > 0) ? 1u : 0u);
              } else {
                if (!(p.debug & 2)) umma_bf16_ss_pair(tmem_base, adesc, bdesc, idesc_main, acc);
                if (with_ones) umma_bf16_ss_pair(tmem_base + 256, adesc, ones_desc, idesc_ones, (k > k_ones0 || kk > 0) ? 1u : 0u);
              }
            }
            umma_commit_pair_mc(&ctl->empty[stage], 3);
            if (k == k1 - 1) umma_commit_pair_mc(&ctl->acc_full, 3);
          }
          __syncwarp();
          if (++stage == GW_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp < 6) {
    // =========================== EPILOGUE ===========================
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    int it = 0;
    for (int u = cid; u < n_units; u += ncl, ++it) {
      const int jh = u % p.nj, vt = (u / p.nj) % p.nvt, s = u / (p.nj * p.nvt);
      const int k0 = s * p.ks_per_split, k1 = min(p.ksteps, k0 + p.ks_per_split);
      const bool has_ones = k0 + ((p.nj == 2 && ((k0 + jh) & 1)) ? 1 : 0) < k1;
      const int v = vt * 256 + (int)rank * 128 + q * 32 + lane;
      {
        mbar_wait(&ctl->acc_full, it & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
        float* dst = p.dw2 + ((size_t)s * p.Vds + v) * p.H + jh * p.N;
#pragma unroll 1
        for (int c = 0; c < p.N / 16; ++c) {
          float z[16];
          tmem_ld_32x16(taddr + c * 16, z);
          tmem_ld_wait();
          if (v < p.Vds && !(p.debug & 4)) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              reinterpret_cast<float4*>(dst + c * 16)[i] = make_float4(z[4 * i], z[4 * i + 1], z[4 * i + 2], z[4 * i + 3]);
          }
        }
        if (has_ones) {
          // at most two adds per element of a partial (the two column halves of H, by K-block parity) onto the
          // caller's zeros: commutative, hence deterministic
          float z[16];
          tmem_ld_32x16(taddr + 256, z);
          tmem_ld_wait();
          if (v < p.Vds) atomicAdd(p.db2 + (size_t)s * p.Vds + v, z[0]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&ctl->acc_empty), 0));
    }
  }
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_pair(tmem_base, 512);
}

// ---------------------------------------------------------------------------------------
// H = 512: one CTA pair takes a 256-row vocabulary tile against ALL of H. Per K = 16 slice two UMMAs
// (columns 0..255 and 256..511 of H, TMEM columns 0..255 / 256..511) share the dS^T slice through
// the tensor core's collector (A_KEEP / A_REUSE): shared memory is read for it once, the dS boxes
// are fetched once per tile instead of once per column half (48 + 48 instead of 64 + 64 B/clk of
// shared-memory fill + operand reads, dS out of L2 once). TMEM is full, so db2 cannot ride along as
// an N = 16 UMMA: the four epilogue warps, idle during the K loop, sum the columns of the dS boxes
// from shared memory (16-byte loads along v, 8 row groups, +16 B/clk) once the stage's UMMAs have
// retired, release the stage, and add their partial sums with red.global.add at the end of the unit.
// Warps as above: 0 dS TMA, 6 h TMA, 1 MMA issuer, 2-5 column sums during the K loop + epilogue.
// ---------------------------------------------------------------------------------------
// K blocks of split s: iteration k in [k0, k1) works on 64-row block base + (k - k0) * stride.
struct GwRange { int k0, k1, base, stride; };
__device__ __forceinline__ GwRange gw_range(const GwParams& p, int s) {
  if (p.interleave) return {0, (p.ksteps - s + p.splits - 1) / p.splits, s, p.splits};
  const int k0 = s * p.ks_per_split;
  return {k0, min(p.ksteps, k0 + p.ks_per_split), k0, 1};
}

constexpr int GWW_STAGES = 4;
constexpr int GWW_STAGE = 6 * GW_BLOCK;  // A: 2 boxes (128 v), B: 4 boxes (2 x 128 j)
constexpr int GWW_CS = 128 * 8 * 4;      // column-sum exchange of the four epilogue warps
constexpr int GWW_SMEM = GWW_STAGES * GWW_STAGE + GWW_CS + 256;

struct GwwCtl {
  uint64_t full[GWW_STAGES], mma_done[GWW_STAGES], empty[GWW_STAGES], acc_full, acc_empty;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(GW_THREADS, 1)
dw2_wide_kernel(const __grid_constant__ CUtensorMap map_ds, const __grid_constant__ CUtensorMap map_h,
                const GwParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  float* cs_x = reinterpret_cast<float*>(smem + GWW_STAGES * GWW_STAGE);
  GwwCtl* ctl = reinterpret_cast<GwwCtl*>(smem + GWW_STAGES * GWW_STAGE + GWW_CS);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cid = blockIdx.x >> 1, ncl = gridDim.x >> 1;
  const int n_units = p.splits * p.nvt;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_ds);
    tma_prefetch_desc(&map_h);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < GWW_STAGES; ++i) {
      mbar_init(&ctl->full[i], 4);      // {dS, h} producer x {leader, peer}; lives in the leader CTA
      mbar_init(&ctl->mma_done[i], 1);  // multicast commit behind the stage's UMMAs (both CTAs)
      mbar_init(&ctl->empty[i], 4);     // this CTA's four column-sum warps, after they have read the dS boxes
    }
    mbar_init(&ctl->acc_full, 1);
    mbar_init(&ctl->acc_empty, 2 * 4);
    fence_barrier_init();
  }
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tmem_alloc_pair(&ctl->tmem_base, 512);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp == 0 || warp == 6) {
    const bool is_a = warp == 0;
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int u = cid; u < n_units; u += ncl) {
        const int vt = u % p.nvt, s = u / p.nvt;
        const GwRange kr = gw_range(p, s);
        const int k0 = kr.k0, k1 = kr.k1;
        const int v0 = vt * 256 + (int)rank * 128;
        for (int k = k0; k < k1; ++k) {
          const int kc = kr.base + (k - kr.k0) * kr.stride;  // the 64-row block of this iteration
          mbar_wait(&ctl->empty[stage], phase ^ 1);
          const uint32_t full = mapa_u32(smem_u32(&ctl->full[stage]), 0);
          const int nbox = is_a ? 2 : 4;
          if (rank == 0) mbar_arrive_expect_tx(&ctl->full[stage], 2 * nbox * GW_BLOCK);  // both CTAs
          else mbar_arrive_cluster(full);
          uint8_t* st = smem + stage * GWW_STAGE + (is_a ? 0 : 2 * GW_BLOCK);
          if (is_a) {
            tma_load_2d_pair(st, &map_ds, full, v0, kc * 64);
            tma_load_2d_pair(st + GW_BLOCK, &map_ds, full, v0 + 64, kc * 64);
          } else {
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {  // this CTA's 128 of the 256 columns of each half of H
              const int j0 = hf * 256 + (int)rank * 128;
              tma_load_2d_pair(st + (2 * hf) * GW_BLOCK, &map_h, full, j0, kc * 64);
              tma_load_2d_pair(st + (2 * hf + 1) * GW_BLOCK, &map_h, full, j0 + 64, kc * 64);
            }
          }
          if (++stage == GWW_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16_mn(256, 256);
      const uint32_t s0 = smem_u32(smem);
      uint32_t stage = 0, phase = 0;
      int it = 0;
      for (int u = cid; u < n_units; u += ncl, ++it) {
        const int s = u / p.nvt;
        const GwRange kr = gw_range(p, s);
        const int k0 = kr.k0, k1 = kr.k1;
        mbar_wait(&ctl->acc_empty, (it & 1) ^ 1);
        tc_fence_after();
        for (int k = k0; k < k1; ++k) {
          mbar_wait(&ctl->full[stage], phase);
          tc_fence_after();
          if (elect_one_sync()) {
            const uint32_t a0 = s0 + stage * GWW_STAGE, b0 = a0 + 2 * GW_BLOCK;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const uint64_t adesc = umma_desc_mn128(a0 + kk * 2048, GW_BLOCK);
              const uint64_t bdesc0 = umma_desc_mn128(b0 + kk * 2048, GW_BLOCK);
              const uint64_t bdesc1 = umma_desc_mn128(b0 + 2 * GW_BLOCK + kk * 2048, GW_BLOCK);
              const uint32_t acc = (k > k0 || kk > 0) ? 1u : 0u;
              umma_bf16_ss_pair_areuse<1>(tmem_base, adesc, bdesc0, idesc, acc);
              umma_bf16_ss_pair_areuse<0>(tmem_base + 256, adesc, bdesc1, idesc, acc);
            }
            umma_commit_pair_mc(&ctl->mma_done[stage], 3);
            if (k == k1 - 1) umma_commit_pair_mc(&ctl->acc_full, 3);
          }
          __syncwarp();
          if (++stage == GWW_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp < 6) {
    const int q = warp & 3;          // TMEM lane quadrant this warp may read
    const int tid = (warp - 2) * 32 + lane;
    const int chunk = tid & 15;      // 8 consecutive v of this CTA's 128
    const int rg = tid >> 4;         // rows 8*rg .. 8*rg+7 of a 64-row K block
    const uint8_t* box = smem + (chunk >> 3) * GW_BLOCK;
    uint32_t stage = 0, phase = 0;
    int it = 0;
    for (int u = cid; u < n_units; u += ncl, ++it) {
      const int vt = u % p.nvt, s = u / p.nvt;
      const GwRange kr = gw_range(p, s);
        const int k0 = kr.k0, k1 = kr.k1;
      // ---- column sums of this CTA's dS boxes while the UMMAs run ----
      float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int k = k0; k < k1; ++k) {
        // (stored-E backward: the operand is E and db2 = E^T g, so each row enters with its weight g; the eight
        // weights of this thread's rows are fetched before the wait and have landed when the stage is taken)
        float wr[8] = {1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f};
        if (p.row_weight) {
          const int kc = kr.base + (k - kr.k0) * kr.stride;
          const float4* wp = reinterpret_cast<const float4*>(p.row_weight + (size_t)kc * 64 + rg * 8);
          const float4 w0 = __ldg(wp), w1 = __ldg(wp + 1);
          wr[0] = w0.x; wr[1] = w0.y; wr[2] = w0.z; wr[3] = w0.w;
          wr[4] = w1.x; wr[5] = w1.y; wr[6] = w1.z; wr[7] = w1.w;
        }
        // The full barrier exists only in the leader CTA, so the column-sum warps of both CTAs take the
        // stage behind its UMMAs (commit multicast) and hand it back to the producers themselves.
        mbar_wait(&ctl->mma_done[stage], phase);
        const uint8_t* a = box + stage * GWW_STAGE;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = rg * 8 + i;  // 128-byte row r of the box; 16-byte chunks XOR-swizzled with r % 8
          const uint4 w = *reinterpret_cast<const uint4*>(a + r * 128 + (((chunk & 7) ^ (r & 7)) << 4));
          const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            cs[2 * j] = fmaf(wr[i], __uint_as_float(ww[j] << 16), cs[2 * j]);
            cs[2 * j + 1] = fmaf(wr[i], __uint_as_float(ww[j] & 0xffff0000u), cs[2 * j + 1]);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&ctl->empty[stage]);  // this warp is done with the stage
        if (++stage == GWW_STAGES) { stage = 0; phase ^= 1; }
      }
      {
        // the eight row groups' sums of a column meet in shared memory and are added in row-group order
        // (deterministic); thread tid then owns column tid of this CTA's 128
#pragma unroll
        for (int j = 0; j < 8; ++j) cs_x[rg * 128 + chunk * 8 + j] = cs[j];
        named_bar_sync(1, 128);
        float a = cs_x[tid];
#pragma unroll
        for (int r = 1; r < 8; ++r) a += cs_x[r * 128 + tid];
        const int v = vt * 256 + (int)rank * 128 + tid;
        if (v < p.Vds) p.db2[(size_t)s * p.Vds + v] = a;
        named_bar_sync(1, 128);  // cs_x is free again
      }
      // ---- accumulators -> this split's partial of dW2 ----
      const int v = vt * 256 + (int)rank * 128 + q * 32 + lane;
      mbar_wait(&ctl->acc_full, it & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
      float* dst = p.dw2 + ((size_t)s * p.Vds + v) * p.H;
#pragma unroll 1
      for (int c = 0; c < 512 / 16; ++c) {
        float z[16];
        tmem_ld_32x16(taddr + c * 16, z);
        tmem_ld_wait();
        if (v < p.Vds && !(p.debug & 4)) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            reinterpret_cast<float4*>(dst + c * 16)[i] = make_float4(z[4 * i], z[4 * i + 1], z[4 * i + 2], z[4 * i + 3]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&ctl->acc_empty), 0));
    }
  }
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_pair(tmem_base, 512);
}

}  // namespace

// K splits of one chunk: one wave of (split, vocabulary tile[, column half]) units over the CTA pairs.
// splits = pairs / units-per-split leaves 6 of the 74 pairs idle at the AISHELL shape (68 units). Filling them
// (13 splits = 442 units = 5.97 waves) made this kernel 3 % faster inside the step and the step no faster
// (24.7 / 25.1 against 24.3 / 24.8 ms in the same call): the chip sits at its power cap, the extra SMs only
// lower the clock of everything else; alone, 1.455 against 1.431 ms.
struct GwSplit { int splits, ks_per_split; bool wide; };
static GwSplit gw_split(long long rows, int V, int H, int sm_pairs) {
  GwSplit g;
  g.wide = (H == 512) && rnnt_tune_int("RNNT_B200_DW2_WIDE", 1) != 0;  // the 256 x 256 kernel at H = 512: A/B only
  const int Vds = rnnt_b200_ds_pitch(V);
  const int nvt = (Vds + 255) / 256;
  const int nj = g.wide ? 1 : H / (H >= 256 ? 256 : 128);
  const int ksteps = (int)(rows / 64);
  const int all_pairs = sm_count() / 2;
  const int pairs = (sm_pairs > 0 && sm_pairs < all_pairs) ? sm_pairs : all_pairs;
  int sw = pairs / (nvt * nj);
  if (sw < 1) sw = 1;
  { const int e = rnnt_tune_int("RNNT_B200_DW2_SPLITS", 0); if (e >= 1) sw = e; }
  if (sw > ksteps) sw = ksteps;
  if (sw < 1) sw = 1;
  g.ks_per_split = (ksteps + sw - 1) / sw;
  if (g.ks_per_split < 1) g.ks_per_split = 1;
  g.splits = (ksteps + g.ks_per_split - 1) / g.ks_per_split;  // every unit has K blocks
  if (g.splits < 1) g.splits = 1;
  return g;
}

extern "C" int rnnt_b200_dw2_splits(long long rows, int V, int H, int sm_pairs) {
  if (rows <= 0 || (rows & 63) || V <= 0 || !(H == 128 || H == 256 || H == 512)) return 0;
  return gw_split(rows, V, H, sm_pairs).splits;
}

static int dw2_gemm_launch(const void* ds_bf16, const void* h_rows, const float* row_weight, long long rows, int V,
                           int H, int sm_pairs, float* dw2_parts, float* db2_parts, void* stream) {
  if (!ds_bf16 || !h_rows || !dw2_parts || !db2_parts) return RNNT_B200_E_NULL;
  if (rows < 0 || (rows & 63) || V <= 0 || !(H == 128 || H == 256 || H == 512)) return RNNT_B200_E_SHAPE;
  if ((reinterpret_cast<uintptr_t>(ds_bf16) & 127u) || (reinterpret_cast<uintptr_t>(h_rows) & 15u) ||
      (reinterpret_cast<uintptr_t>(dw2_parts) & 15u))
    return RNNT_B200_E_ALIGN;
  if (rows == 0) return RNNT_B200_OK;
  int rc = rnnt_check_arch();
  if (rc != RNNT_B200_OK) return rc;
  const int Vds = rnnt_b200_ds_pitch(V);
  CUtensorMap mds, mh;
  rc = make_map_bf16(&mds, ds_bf16, (uint64_t)rows, (uint64_t)Vds, (uint64_t)Vds, 64, 64,
                     CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != RNNT_B200_OK) return rc;
  rc = make_map_bf16(&mh, h_rows, (uint64_t)rows, (uint64_t)H, (uint64_t)rnnt_hp(H), 64, 64,
                     CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != RNNT_B200_OK) return rc;
  const GwSplit g = gw_split(rows, V, H, sm_pairs);
  if (row_weight && (!g.wide || (reinterpret_cast<uintptr_t>(row_weight) & 15u))) return RNNT_B200_E_SHAPE;
  GwParams p;
  p.row_weight = row_weight;
  p.dw2 = dw2_parts; p.db2 = db2_parts; p.Vds = Vds; p.H = H;
  p.nvt = (Vds + 255) / 256;
  p.N = H >= 256 ? 256 : 128;
  p.nj = H / p.N;
  p.ksteps = (int)(rows / 64);
  p.splits = g.splits;
  p.ks_per_split = g.ks_per_split;
  p.debug = rnnt_tune_int("RNNT_B200_DW2_DEBUG", 0);
  const bool wide = g.wide;
  const int all_pairs = sm_count() / 2;
  const int pairs = (sm_pairs > 0 && sm_pairs < all_pairs) ? sm_pairs : all_pairs;
  // interleaved K splits: all units of a launch walk the chunk's rows front to back together (h boxes shared
  // through L2 by every unit, and the dS rows by a dh GEMM that runs beside this kernel: DESIGN.md 3.4)
  p.interleave = (wide && rnnt_tune_int("RNNT_B200_DW2_INTERLEAVE", 1)) ? 1 : 0;
  const int units = wide ? p.splits * p.nvt : p.splits * p.nvt * p.nj;
  const int grid = 2 * (units < pairs ? units : pairs);
  cudaError_t e = wide ? cudaFuncSetAttribute(dw2_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GWW_SMEM)
                       : cudaFuncSetAttribute(dw2_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GW_SMEM);
  if (e != cudaSuccess) return (int)e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(GW_THREADS);
  cfg.dynamicSmemBytes = wide ? GWW_SMEM : GW_SMEM;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  e = wide ? cudaLaunchKernelEx(&cfg, dw2_wide_kernel, mds, mh, p) : cudaLaunchKernelEx(&cfg, dw2_gemm_kernel, mds, mh, p);
  if (e != cudaSuccess) return (int)e;
  return RNNT_B200_OK;
}

extern "C" int rnnt_b200_dw2_gemm(const void* ds_bf16, const void* h_rows, long long rows, int V, int H,
                                  int sm_pairs, float* dw2_parts, float* db2_parts, void* stream) {
  return dw2_gemm_launch(ds_bf16, h_rows, nullptr, rows, V, H, sm_pairs, dw2_parts, db2_parts, stream);
}

extern "C" int rnnt_b200_dw2_gemm_weighted(const void* e_bf16, const void* h_scaled_rows, const float* row_weight,
                                           long long rows, int V, int H, int sm_pairs, float* dw2_parts,
                                           float* db2_parts, void* stream) {
  if (!row_weight) return RNNT_B200_E_NULL;
  if (H != 512) return RNNT_B200_E_SHAPE;  // the 256 x 512 kernel only
  return dw2_gemm_launch(e_bf16, h_scaled_rows, row_weight, rows, V, H, sm_pairs, dw2_parts, db2_parts, stream);
}
