// dpre = (dS . W2) * (1 - h^2)  -- the gradient that reaches the pre-activation of the joint's
// tanh (rnnt/model.py:32-35 in reverse: the backward of project_layer followed by the backward of
// tanh; the reference gets both from autograd through the materialised logits, train.py:45).
//
//   dh[row, j]   = sum_v dS[row, v] * W2q[v, j]          (rows = lattice cells, v < ds_pitch(V), j < H)
//   dpre[row, j] = bf16( dh[row, j] * (1 - h[row, j]^2) )
//
// One GEMM per staged dS chunk on the 5th-gen tensor cores: M = rows (256 per CTA pair, cta_group::2),
// N = H (all of it: per K block of 64 four UMMAs of N = 256 into TMEM columns 0..255, then four into columns
// 256..511, so that a dS box is fetched from L2 once for both column halves), K = v. dS is K-major as it lies
// (128 rows x 64 v TMA boxes, SWIZZLE_128B); B[n = j][k = v] comes K-major from a TRANSPOSED bf16 pack of W2
// (H, Vpad) that rnnt_b200_pack_vocab writes next to the (Vpad, H) one (64 j x 64 v boxes), so both operands use
// one descriptor form. Shared memory per K block of 64: fill 48 KB + operand reads 64 KB per 1024 clk of MMA =
// 112 of 128 B/clk/SM (measured with clock64 stamps: 1024-1060 clk per K block; profiles/r02_notes.md).
//
// Order of the UMMAs: a first version alternated the two accumulator halves UMMA by UMMA so that the dS slice
// could be shared through the tensor core's collector (as dw2_wide_kernel does): 145 clk per UMMA without and
// 154 clk with the collector hint, against 128 clk when the four K slices of one half are issued back to back
// (with or without TMA traffic: the cost is in the tensor pipe, not in shared memory).
//
// TMEM is full (512 columns), so the epilogue of a unit cannot overlap the next unit's MMAs; it is
// kept short instead: the unit's h rows (the tanh derivative) are TMA-loaded into ring stages that the
// last K blocks have just released ("H items" of the same ring), eight warps drain TMEM, multiply, pack to
// bf16 and write dpre IN PLACE over the h blocks (same swizzled 16-byte chunks, same thread), and a store warp
// sends every finished 128 x 64 block out with one TMA store (full 128-byte lines).
//
// Ring items of one unit, in order: NK K items (dS box + W2^T boxes; consumed by the MMA thread, released by
// a multicast tcgen05.commit) then NHI H items (up to three 128 x 64 h blocks each; consumed by the CTA's own
// eight epilogue warps, released by the store warp once its stores have read them). Every role walks the same
// item sequence (RingWalk). An mbarrier parity wait is only meaningful within one phase of the barrier, so every
// waiter must follow the barrier it waits on use by use: every producer thread performs the acquire wait of
// EVERY item of the ring, also of the items it loads nothing for (a first version whose h producer skipped the
// K items ran phases ahead of the barriers and its waits passed spuriously). Three producer threads: one wait +
// expect_tx + five TMA issues per 1024 clk of MMA is more than one thread manages.
// Warps: 0-7 epilogue, 8 TMA producer (dS boxes, H items), 9 TMEM alloc + MMA issue (leader CTA), 10 / 11 TMA
// producers of the W2^T boxes of column half 0 / 1, 12 store warp.
#include <cuda.h>
#include <cuda_bf16.h>
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "tma_host.cuh"

using namespace sm100;

#ifdef RNNT_B200_TUNE
#include <stdio.h>
// RNNT_B200_DH_DEBUG: 1 no epilogue work (barriers only), 4 no dpre stores, 8 no TMA loads of the K items (MMA on
// stale buffers), 16 accumulator halves alternating per K slice through the collector (the first version)
#define RNNT_DH_DBG(p) ((p).debug)
#define RNNT_DH_TRACE(p) ((p).trace != nullptr && blockIdx.x == 0)  // RNNT_B200_DH_TRACE=<file>: clock64 stamps of CTA 0
#else
#define RNNT_DH_DBG(p) 0
#define RNNT_DH_TRACE(p) false
#endif

namespace {

constexpr int DH_THREADS = 416;                   // 13 warps (roles in the header comment)
constexpr int DH_STAGES = 4;
constexpr int DH_ABYTES = 128 * 128;              // dS box: 128 rows x 64 v
constexpr int DH_BBOX = 64 * 128;                 // W2^T box: 64 j rows x 64 v
constexpr int DH_HBLK = 128 * 128;                // h block: 128 rows x 64 j
constexpr int DH_STAGE = DH_ABYTES + 4 * DH_BBOX;  // 48 KB = three h blocks
constexpr int DH_SMEM = DH_STAGES * DH_STAGE + 512;

struct DhCtl {
  uint64_t full[DH_STAGES];    // leader CTA: K item landed (the producer threads of both CTAs + the bytes of both)
  uint64_t kempty[DH_STAGES];  // both CTAs: multicast commit behind the K item's UMMAs
  uint64_t hfull[DH_STAGES];   // this CTA: h blocks landed
  uint64_t hempty[DH_STAGES];  // this CTA: the store warp's TMA stores have read the stage's blocks
  uint64_t blk_done[8];        // this CTA: the eight epilogue warps have written dpre over h block i
  uint64_t acc_full, acc_empty;
  uint32_t tmem_base;
};
static_assert(sizeof(DhCtl) <= 512, "ctl region");

struct DhParams {
  int H, HN, NH;          // UMMA N (min(H, 256)) and the number of such column halves (1 or 2)
  int NK;                 // K blocks of 64 (ds_pitch / 64)
  int NHI;                // H items per unit: ceil((H / 64) / 3)
  int tile_lo, tile_hi;   // tiles of this chunk; dS row 0 is row 0 of tile_lo
  int debug;              // tuning build only
  long long* trace;       // tuning build only: 3 roles x 4096 x 3 clock64 stamps
  const float* row_scale; // optional: the product's row r (of ALL tiles' rows) is multiplied by row_scale[r]
};

// Position in the ring's item sequence. bit s of kpar / hpar: parity of the number of K / H items that
// have used stage s so far; last_h: the most recent item of stage s was an H item; used: stage s was used.
struct RingWalk {
  uint32_t n = 0, kpar = 0, hpar = 0, last_h = 0, used = 0;
  __device__ __forceinline__ uint32_t stage() const { return n & (DH_STAGES - 1); }
  // producer side: wait until whatever occupied the stage before has been released
  __device__ __forceinline__ void acquire(DhCtl* ctl) const {
    const uint32_t s = stage(), bit = 1u << s;
    if (!(used & bit)) return;
    if (last_h & bit) mbar_wait(&ctl->hempty[s], ((hpar >> s) & 1u) ^ 1u);
    else mbar_wait(&ctl->kempty[s], ((kpar >> s) & 1u) ^ 1u);
  }
  __device__ __forceinline__ uint32_t k_parity() const { return (kpar >> stage()) & 1u; }
  __device__ __forceinline__ uint32_t h_parity() const { return (hpar >> stage()) & 1u; }
  __device__ __forceinline__ void next_k() {
    const uint32_t bit = 1u << stage();
    kpar ^= bit; last_h &= ~bit; used |= bit; ++n;
  }
  __device__ __forceinline__ void next_h() {
    const uint32_t bit = 1u << stage();
    hpar ^= bit; last_h |= bit; used |= bit; ++n;
  }
};

__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

__global__ void __launch_bounds__(DH_THREADS, 1)
dh_gemm_kernel(const __grid_constant__ CUtensorMap map_ds, const __grid_constant__ CUtensorMap map_w2,
               const __grid_constant__ CUtensorMap map_h, const __grid_constant__ CUtensorMap map_out,
               const DhParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  DhCtl* ctl = reinterpret_cast<DhCtl*>(smem + DH_STAGES * DH_STAGE);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cid = blockIdx.x >> 1, ncl = gridDim.x >> 1;
  const int n_units = (p.tile_hi - p.tile_lo + 1) >> 1;
  const int HB = p.H / 64;  // h blocks of a row tile

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&map_ds);
    tma_prefetch_desc(&map_w2);
    tma_prefetch_desc(&map_h);
    tma_prefetch_desc(&map_out);
  }
  if (warp == 9 && lane == 0) {
    for (int i = 0; i < DH_STAGES; ++i) {
      mbar_init(&ctl->full[i], 2 * (1 + p.NH));  // producers of both CTAs
      mbar_init(&ctl->kempty[i], 1);
      mbar_init(&ctl->hfull[i], 1);
      mbar_init(&ctl->hempty[i], 1);
    }
    for (int i = 0; i < 8; ++i) mbar_init(&ctl->blk_done[i], 8);
    mbar_init(&ctl->acc_full, 1);
    mbar_init(&ctl->acc_empty, 2 * 8);
    fence_barrier_init();
  }
  __syncthreads();
  cluster_sync_all();
  if (warp == 9) {
    tmem_alloc_pair(&ctl->tmem_base, 512);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp == 8 || warp == 10 || warp == 11) {
    // =========================== TMA PRODUCERS ===========================
    // role 0 (warp 8): dS boxes and H items; role 1 / 2 (warps 10 / 11): W2^T boxes of column half 0 / 1.
    const int role = warp == 8 ? 0 : warp - 9;
    if (lane == 0 && role <= p.NH) {
      RingWalk rw;
      const int nbx = p.HN / 128;  // 64-row W2^T boxes per CTA, K item and column half
      for (int u = cid; u < n_units; u += ncl) {
        const int row0 = (2 * u + (int)rank) * RNNT_TILE_M;  // chunk-local dS row of this CTA's tile
        int kb = (u * 5) % p.NK;  // rotated start: the pairs do not pull the same W2^T boxes out of L2 at once
        for (int i = 0; i < p.NK; ++i) {
          const long long tr0 = RNNT_DH_TRACE(p) ? clock64() : 0;
          rw.acquire(ctl);
          const long long tr1 = RNNT_DH_TRACE(p) ? clock64() : 0;
          const uint32_t s = rw.stage();
          const uint32_t full = mapa_u32(smem_u32(&ctl->full[s]), 0);
          const uint32_t bytes = role == 0 ? DH_ABYTES : nbx * DH_BBOX;
          if (RNNT_DH_DBG(p) & 8) {
            mbar_arrive_cluster(full);
            rw.next_k();
            continue;
          }
          if (rank == 0) mbar_arrive_expect_tx(&ctl->full[s], 2 * bytes);  // both CTAs' boxes
          else mbar_arrive_cluster(full);
          uint8_t* st = smem + s * DH_STAGE;
          if (role == 0) {
            tma_load_2d_pair(st, &map_ds, full, kb * 64, row0);
          } else {
            const int hf = role - 1;
            for (int bx = 0; bx < nbx; ++bx) {  // this CTA's HN/2 rows j of W2^T, 64 v wide: K-major like the dS box
              const int j0 = hf * p.HN + (int)rank * (p.HN / 2) + bx * 64;
              tma_load_2d_pair(st + DH_ABYTES + (hf * 2 + bx) * DH_BBOX, &map_w2, full, kb * 64, j0);
            }
          }
          if (RNNT_DH_TRACE(p) && role == 0 && rw.n < 4096) {
            long long* t = p.trace + 3 * 4096 + 3 * rw.n;
            t[0] = tr0; t[1] = tr1; t[2] = clock64();
          }
          rw.next_k();
          if (++kb == p.NK) kb = 0;
        }
        const int tile = p.tile_lo + 2 * u + (int)rank;  // past the stash: TMA fills zeros
        for (int i = 0; i < p.NHI; ++i) {  // H items: this CTA's h rows, into stages the last K blocks have released
          rw.acquire(ctl);                 // (every producer follows every release: see the header comment)
          if (role == 0) {
            const uint32_t s = rw.stage();
            const int nb = min(3, HB - 3 * i);
            mbar_arrive_expect_tx(&ctl->hfull[s], nb * DH_HBLK);
            for (int j = 0; j < nb; ++j)
              tma_load_2d(smem + s * DH_STAGE + j * DH_HBLK, &map_h, &ctl->hfull[s], (3 * i + j) * 64,
                          tile * RNNT_TILE_M);
          }
          rw.next_h();
        }
      }
    }
  } else if (warp == 9) {
    // =========================== MMA ISSUER (leader CTA) ===========================
    // Per K item: the four K = 16 slices into column half 0 (TMEM columns 0..255), then the four into half 1
    // (see the header comment for why not alternating).
    if (rank == 0) {
      const uint32_t idesc = umma_idesc_bf16(256, p.HN);  // both operands K-major
      const uint32_t s0 = smem_u32(smem);
      RingWalk rw;
      int it = 0;
      for (int u = cid; u < n_units; u += ncl, ++it) {
        mbar_wait(&ctl->acc_empty, (it & 1) ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < p.NK; ++kb) {
          const uint32_t s = rw.stage();
          const long long tr0 = RNNT_DH_TRACE(p) ? clock64() : 0;
          mbar_wait(&ctl->full[s], rw.k_parity());
          const long long tr1 = RNNT_DH_TRACE(p) ? clock64() : 0;
          tc_fence_after();
          if (elect_one_sync()) {
            const uint32_t a0 = s0 + s * DH_STAGE, b0 = a0 + DH_ABYTES;
            const uint32_t acc0 = kb > 0 ? 1u : 0u;
            if (p.NH == 2 && (RNNT_DH_DBG(p) & 16)) {  // experiment: halves alternating per K slice through the collector
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {
                umma_bf16_ss_pair_areuse<1>(tmem_base, umma_desc_kmajor<128>(a0) + 2 * kk,
                                            umma_desc_kmajor<128>(b0) + 2 * kk, idesc, (kk > 0) ? 1u : acc0);
                umma_bf16_ss_pair_areuse<0>(tmem_base + 256, umma_desc_kmajor<128>(a0) + 2 * kk,
                                            umma_desc_kmajor<128>(b0 + 2 * DH_BBOX) + 2 * kk, idesc, (kk > 0) ? 1u : acc0);
              }
            } else {
              for (int hf = 0; hf < p.NH; ++hf) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)  // K = 16: 32 bytes along a row of the dS box and of the W2^T boxes
                  umma_bf16_ss_pair(tmem_base + hf * 256, umma_desc_kmajor<128>(a0) + 2 * kk,
                                    umma_desc_kmajor<128>(b0 + hf * 2 * DH_BBOX) + 2 * kk, idesc, (kk > 0) ? 1u : acc0);
              }
            }
            umma_commit_pair_mc(&ctl->kempty[s], 3);
            if (kb == p.NK - 1) umma_commit_pair_mc(&ctl->acc_full, 3);
          }
          __syncwarp();
          if (RNNT_DH_TRACE(p) && lane == 0 && rw.n < 4096) {
            long long* t = p.trace + 3 * rw.n;
            t[0] = tr0; t[1] = tr1; t[2] = clock64();
          }
          rw.next_k();
        }
        for (int i = 0; i < p.NHI; ++i) rw.next_h();
      }
    }
  } else if (warp == 12) {
    // =========================== STORE WARP ===========================
    // dpre lies where h lay (the epilogue warps overwrite the h blocks in place): one 128 x 64 TMA store per
    // block as soon as all eight epilogue warps are done with it; a stage goes back to the producers when the
    // stores of its blocks have read it.
    if (lane == 0) {
      RingWalk rw;
      int it = 0;
      for (int u = cid; u < n_units; u += ncl, ++it) {
        const int tile = p.tile_lo + 2 * u + (int)rank;
        const bool dead = tile >= p.tile_hi;  // odd tail: this CTA only keeps the pair's protocol going
        for (int kb = 0; kb < p.NK; ++kb) rw.next_k();
        for (int i = 0; i < p.NHI; ++i) {
          const uint32_t s = rw.stage();
          const int nb = min(3, HB - 3 * i);
          for (int j = 0; j < nb; ++j) {
            mbar_wait(&ctl->blk_done[3 * i + j], it & 1);
            if (!dead && !(RNNT_DH_DBG(p) & 4)) {
              tma_store_2d(&map_out, smem + s * DH_STAGE + j * DH_HBLK, (3 * i + j) * 64, tile * RNNT_TILE_M);
              bulk_commit_group();
            }
          }
          bulk_wait_group_read0();
          mbar_arrive(&ctl->hempty[s]);
          rw.next_h();
        }
      }
      bulk_wait_group0();  // all stores have landed
    }
  } else if (warp < 8) {
    // =========================== EPILOGUE (warps 0-7) ===========================
    // Warp w reads TMEM lanes 32*(w&3).. (thread = row of the tile) and the 32-column groups g = (w>>2),
    // (w>>2)+2, ...: in its i-th round every warp works on h block i (64 columns = two groups).
    const int q = warp & 3, sub = warp >> 2;
    const int row = q * 32 + lane;
    const int ngroups = p.H / 32;
    RingWalk rw;
    int it = 0;
    for (int u = cid; u < n_units; u += ncl, ++it) {
      for (int kb = 0; kb < p.NK; ++kb) rw.next_k();
      const long long tr0 = RNNT_DH_TRACE(p) ? clock64() : 0;
      // (stored-E backward: the A operand is E, the row factor g of dS = diag(g) . E comes in here)
      float gr = 1.f;
      if (p.row_scale) {
        const int tile = p.tile_lo + 2 * u + (int)rank;
        gr = tile < p.tile_hi ? __ldg(p.row_scale + (size_t)tile * RNNT_TILE_M + row) : 0.f;
      }
      mbar_wait(&ctl->acc_full, it & 1);
      const long long tr1 = RNNT_DH_TRACE(p) ? clock64() : 0;
      tc_fence_after();
      int cur_item = -1;
      uint32_t cur_stage = 0;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
      // One 32-column group: the group's accumulators have landed in zc; the next group's TMEM load (into zn) is
      // issued first and flies while this group is multiplied by the tanh derivative and written over h.
      auto group = [&](float (&zc)[32], float (&zn)[32], int g) {
        const int blk = g >> 1, item = blk / 3;
        while (cur_item < item) {  // H items arrive in order; every warp follows every one of them
          cur_stage = rw.stage();
          mbar_wait(&ctl->hfull[cur_stage], rw.h_parity());
          rw.next_h();
          ++cur_item;
        }
        uint8_t* hrow = smem + cur_stage * DH_STAGE + (blk - 3 * item) * DH_HBLK + row * 128;
        uint4 hv[4];
#pragma unroll
        for (int c = 0; c < 4; ++c)
          hv[c] = *reinterpret_cast<const uint4*>(hrow + ((((g & 1) * 4 + c) ^ (row & 7)) << 4));
        tmem_ld_wait();  // this group's accumulators
        if (g + 2 < ngroups) {
          tmem_ld_32x32(taddr + (g + 2) * 32, zn);
        } else {  // last TMEM read of the unit: hand the accumulators back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&ctl->acc_empty), 0));
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const uint32_t hw[4] = {hv[c].x, hv[c].y, hv[c].z, hv[c].w};
          uint32_t o[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float h0 = __uint_as_float(hw[j] << 16), h1 = __uint_as_float(hw[j] & 0xffff0000u);
            o[j] = pack2_bf16(zc[c * 8 + 2 * j] * (gr * fmaf(-h0, h0, 1.f)), zc[c * 8 + 2 * j + 1] * (gr * fmaf(-h1, h1, 1.f)));
          }
          *reinterpret_cast<uint4*>(hrow + ((((g & 1) * 4 + c) ^ (row & 7)) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
        }
        fence_proxy_async_smem();  // the TMA store (async proxy) reads what this thread wrote
        __syncwarp();
        if (lane == 0) mbar_arrive(&ctl->blk_done[blk]);
      };
      if (!(RNNT_DH_DBG(p) & 1)) {
        float z0[32], z1[32];
        tmem_ld_32x32(taddr + sub * 32, z0);
        for (int g = sub; g < ngroups; g += 4) {  // ngroups / 2 groups per warp: 2, 4 or 8 (even)
          group(z0, z1, g);
          group(z1, z0, g + 2);
        }
      } else {  // tuning build: barriers only (what the unit costs without its epilogue work)
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&ctl->acc_empty), 0));
        for (int b = 0; b < HB; ++b)
          if (lane == 0) mbar_arrive(&ctl->blk_done[b]);
      }
      while (cur_item < p.NHI - 1) {  // (only in the barriers-only mode: every warp reads every item otherwise)
        cur_stage = rw.stage();
        mbar_wait(&ctl->hfull[cur_stage], rw.h_parity());
        rw.next_h();
        ++cur_item;
      }
      if (RNNT_DH_TRACE(p) && threadIdx.x == 0 && it < 4096) {
        long long* t = p.trace + 2 * 3 * 4096 + 3 * it;
        t[0] = tr0; t[1] = tr1; t[2] = clock64();
      }
    }
  }
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  if (warp == 9) tmem_dealloc_pair(tmem_base, 512);
}

// ---------------------------------------------------------------------------------------
// dP_enc[b,t,:] = sum_u dpre,  dP_dec[b,u,:] = sum_t dpre  (the backward of the broadcast add that replaces
// the reference's repeat + cat, rnnt/model.py:26-31), DETERMINISTIC: no atomics, every sum in a fixed order.
// Stage 1: one CTA per (utterance b, block of 8 lattice rows t). A thread owns 8 consecutive hidden columns
// (16-byte bf16 loads), H/8 threads span a packed row, the 256/(H/8) row lanes take u = lane, lane + lanes, ...
// For every u the sum over the block's 8 rows is a complete partial for dP_dec: written once to
// part_dec[b][tblock][u][:]. The sums over u for the block's 8 rows stay in registers per row lane and are
// combined across the row lanes in lane order through shared memory: dP_enc rows are final, written once.
// Stage 2: dP_dec[b,u,:] = sum over the t blocks in order.
// Cells of an utterance are contiguous in the tile-packed row order: row(t,u) = first_tile_b*128 + t*(U_b+1) + u.
// ---------------------------------------------------------------------------------------
constexpr int DR_TT = 8;

__device__ __forceinline__ void unpack8(const uint4& w, float (&f)[8]) {
  const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    f[2 * j] = __uint_as_float(ww[j] << 16);
    f[2 * j + 1] = __uint_as_float(ww[j] & 0xffff0000u);
  }
}

__global__ void __launch_bounds__(256, 2)
dpre_reduce_kernel(const __nv_bfloat16* __restrict__ dpre, const int32_t* __restrict__ act_lens,
                   const int32_t* __restrict__ label_lens, int T, int U1, int H, int n_tb,
                   float* __restrict__ dpe, float* __restrict__ part_dec) {
  __shared__ float4 sm[2 * 256];  // [row lane][H] floats: 256 threads x 8 floats
  const int b = blockIdx.y;
  const int Tb = clamp_int(act_lens[b], 1, T);
  const int w = clamp_int(label_lens[b], 0, U1 - 1) + 1;
  const int t0 = blockIdx.x * DR_TT;
  if (t0 >= Tb) return;
  int first_tile = 0;  // tiles of the utterances before b
  for (int i = 0; i < b; ++i) {
    const int n = clamp_int(act_lens[i], 1, T) * (clamp_int(label_lens[i], 0, U1 - 1) + 1);
    first_tile += (n + RNNT_TILE_M - 1) / RNNT_TILE_M;
  }
  const int nt = min(DR_TT, Tb - t0);
  const int cpr = H / 8;        // 16-byte chunks per row
  const int lanes = 256 / cpr;  // row lanes of the CTA
  const int chunk = threadIdx.x % cpr;
  const int rl = threadIdx.x / cpr;
  const uint4* d4 = reinterpret_cast<const uint4*>(dpre + (size_t)first_tile * RNNT_TILE_M * H) + chunk;
  float acc_e[DR_TT][8];
#pragma unroll
  for (int i = 0; i < DR_TT; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc_e[i][j] = 0.f;
  float* pd_base = part_dec + (((size_t)b * n_tb + blockIdx.x) * U1) * H + chunk * 8;
  // (H / 8 need not divide 256: the threads past the last whole row lane only take part in the barriers)
  for (int u = rl < lanes ? rl : w; u < w; u += lanes) {
    float acc_d[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc_d[j] = 0.f;
#pragma unroll
    for (int i0 = 0; i0 < DR_TT; i0 += 4) {  // two batches of 4 rows, unconditional loads
      uint4 g[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int ti = (i0 + i < nt) ? t0 + i0 + i : t0;  // a row that always exists
        g[i] = ld_stream_u4(d4 + ((size_t)ti * w + u) * cpr);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float gf[8];
        unpack8(g[i], gf);
        const bool ok = i0 + i < nt;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float v = ok ? gf[j] : 0.f;
          acc_d[j] += v;
          acc_e[i0 + i][j] += v;
        }
      }
    }
    float4* dst = reinterpret_cast<float4*>(pd_base + (size_t)u * H);
    dst[0] = make_float4(acc_d[0], acc_d[1], acc_d[2], acc_d[3]);
    dst[1] = make_float4(acc_d[4], acc_d[5], acc_d[6], acc_d[7]);
  }
  // combine the row lanes in lane order (fixed summation order)
  for (int i = 0; i < nt; ++i) {
    sm[2 * threadIdx.x] = make_float4(acc_e[i][0], acc_e[i][1], acc_e[i][2], acc_e[i][3]);
    sm[2 * threadIdx.x + 1] = make_float4(acc_e[i][4], acc_e[i][5], acc_e[i][6], acc_e[i][7]);
    __syncthreads();
    if (rl == 0) {
      float4 a = sm[2 * chunk], c = sm[2 * chunk + 1];
      for (int l = 1; l < lanes; ++l) {
        const float4 x = sm[2 * (l * cpr + chunk)], y = sm[2 * (l * cpr + chunk) + 1];
        a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
        c.x += y.x; c.y += y.y; c.z += y.z; c.w += y.w;
      }
      float4* dst = reinterpret_cast<float4*>(dpe + ((size_t)b * T + t0 + i) * H + chunk * 8);
      dst[0] = a;
      dst[1] = c;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(128)
dpd_final_kernel(const float* __restrict__ part_dec, const int32_t* __restrict__ act_lens,
                 const int32_t* __restrict__ label_lens, int T, int U1, int H, int n_tb,
                 float* __restrict__ dpd) {
  const int b = blockIdx.y, u = blockIdx.x;
  const int Tb = clamp_int(act_lens[b], 1, T);
  const int w = clamp_int(label_lens[b], 0, U1 - 1) + 1;
  const int nb = (Tb + DR_TT - 1) / DR_TT;
  for (int c = threadIdx.x; c < H / 4; c += blockDim.x) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (u < w) {
      const float4* src = reinterpret_cast<const float4*>(part_dec + (((size_t)b * n_tb) * U1 + u) * H) + c;
      for (int tb = 0; tb < nb; ++tb) {
        const float4 x = __ldg(src + (size_t)tb * U1 * (H / 4));
        a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
      }
    }
    reinterpret_cast<float4*>(dpd + ((size_t)b * U1 + u) * H)[c] = a;
  }
}

}  // namespace

extern "C" int rnnt_b200_dh_gemm_supported(int H) { return (H == 128 || H == 256 || H == 512) ? 1 : 0; }

static int dh_gemm_launch(const void* ds_bf16, const void* w2t_bf16, const void* h_stash, const float* row_scale,
                          int V, int H, int tile_cap, int tile_lo, int tile_hi, int sm_pairs, void* dpre_bf16,
                          void* stream) {
  if (!ds_bf16 || !w2t_bf16 || !h_stash || !dpre_bf16) return RNNT_B200_E_NULL;
  if (V <= 0 || !rnnt_b200_dh_gemm_supported(H)) return RNNT_B200_E_SHAPE;
  if (tile_cap < 1 || tile_lo < 0 || tile_hi < tile_lo || tile_hi > tile_cap) return RNNT_B200_E_SHAPE;
  if ((reinterpret_cast<uintptr_t>(ds_bf16) & 127u) || (reinterpret_cast<uintptr_t>(w2t_bf16) & 127u) ||
      (reinterpret_cast<uintptr_t>(h_stash) & 127u) || (reinterpret_cast<uintptr_t>(dpre_bf16) & 127u))
    return RNNT_B200_E_ALIGN;
  if (tile_hi == tile_lo) return RNNT_B200_OK;
  int rc = rnnt_check_arch();
  if (rc != RNNT_B200_OK) return rc;
  const int Vds = rnnt_b200_ds_pitch(V), Vpad = rnnt_b200_vocab_pad(V);
  const uint64_t rows = (uint64_t)(tile_hi - tile_lo) * RNNT_TILE_M, all_rows = (uint64_t)tile_cap * RNNT_TILE_M;
  CUtensorMap mds, mw2, mh, mout;
  rc = make_map_bf16(&mds, ds_bf16, rows, (uint64_t)Vds, (uint64_t)Vds, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != RNNT_B200_OK) return rc;
  rc = make_map_bf16(&mw2, w2t_bf16, (uint64_t)H, (uint64_t)Vpad, (uint64_t)Vpad, 64, 64, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != RNNT_B200_OK) return rc;
  rc = make_map_bf16(&mh, h_stash, all_rows, (uint64_t)H, (uint64_t)rnnt_hp(H), 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != RNNT_B200_OK) return rc;
  rc = make_map_bf16(&mout, dpre_bf16, all_rows, (uint64_t)H, (uint64_t)H, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc != RNNT_B200_OK) return rc;
  DhParams p;
  p.H = H;
  p.HN = H >= 256 ? 256 : 128;
  p.NH = H / p.HN;
  p.NK = Vds / 64;
  p.NHI = (H / 64 + 2) / 3;
  p.tile_lo = tile_lo;
  p.tile_hi = tile_hi;
  p.debug = rnnt_tune_int("RNNT_B200_DH_DEBUG", 0);
  p.row_scale = row_scale;
  const int n_units = (tile_hi - tile_lo + 1) / 2;
  const int all_pairs = sm_count() / 2;
  const int pairs = (sm_pairs > 0 && sm_pairs < all_pairs) ? sm_pairs : all_pairs;
  const int grid = 2 * (n_units < pairs ? n_units : pairs);
  cudaError_t e = cudaFuncSetAttribute(dh_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DH_SMEM);
  if (e != cudaSuccess) return (int)e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(DH_THREADS);
  cfg.dynamicSmemBytes = DH_SMEM;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  p.trace = nullptr;
#ifdef RNNT_B200_TUNE
  const char* trace_path = getenv("RNNT_B200_DH_TRACE");  // tuning build only: allocates, synchronises, writes a file
  const size_t trace_bytes = 3 * 4096 * 3 * sizeof(long long);
  if (trace_path && cudaMalloc(&p.trace, trace_bytes) == cudaSuccess) cudaMemsetAsync(p.trace, 0, trace_bytes, cfg.stream);
#endif
  e = cudaLaunchKernelEx(&cfg, dh_gemm_kernel, mds, mw2, mh, mout, p);
  if (e != cudaSuccess) return (int)e;
#ifdef RNNT_B200_TUNE
  if (p.trace) {
    long long* host = (long long*)malloc(trace_bytes);
    cudaStreamSynchronize(cfg.stream);
    cudaMemcpy(host, p.trace, trace_bytes, cudaMemcpyDeviceToHost);
    FILE* f = fopen(trace_path, "wb");
    if (f) { fwrite(host, 1, trace_bytes, f); fclose(f); }
    free(host);
    cudaFree(p.trace);
  }
#endif
  return RNNT_B200_OK;
}

extern "C" int rnnt_b200_dh_gemm(const void* ds_bf16, const void* w2t_bf16, const void* h_stash, int V, int H,
                                 int tile_cap, int tile_lo, int tile_hi, int sm_pairs, void* dpre_bf16, void* stream) {
  return dh_gemm_launch(ds_bf16, w2t_bf16, h_stash, nullptr, V, H, tile_cap, tile_lo, tile_hi, sm_pairs, dpre_bf16,
                        stream);
}

extern "C" int rnnt_b200_dh_gemm_scaled(const void* e_bf16, const void* w2t_bf16, const void* h_stash,
                                        const float* row_scale, int V, int H, int tile_cap, int tile_lo,
                                        int tile_hi, int sm_pairs, void* dpre_bf16, void* stream) {
  if (!row_scale) return RNNT_B200_E_NULL;
  return dh_gemm_launch(e_bf16, w2t_bf16, h_stash, row_scale, V, H, tile_cap, tile_lo, tile_hi, sm_pairs, dpre_bf16,
                        stream);
}

extern "C" size_t rnnt_b200_dpre_reduce_scratch_elems(int B, int T, int U1, int H) {
  if (B <= 0 || T <= 0 || U1 <= 0 || H <= 0) return 0;
  return (size_t)B * (size_t)((T + DR_TT - 1) / DR_TT) * (size_t)U1 * (size_t)H;
}

extern "C" int rnnt_b200_dpre_reduce(const void* dpre_bf16, const int32_t* act_lens, const int32_t* label_lens,
                                     int B, int T, int U1, int H, float* scratch, float* dp_enc, float* dp_dec,
                                     void* stream) {
  if (!dpre_bf16 || !act_lens || !label_lens || !scratch || !dp_enc || !dp_dec) return RNNT_B200_E_NULL;
  if (B <= 0 || B > 65535 || T <= 0 || U1 <= 0 || U1 > 65535 || H < 64 || H > 512 || (H & 63)) return RNNT_B200_E_SHAPE;
  if ((reinterpret_cast<uintptr_t>(dpre_bf16) & 15u) || (reinterpret_cast<uintptr_t>(scratch) & 15u) ||
      (reinterpret_cast<uintptr_t>(dp_enc) & 15u) || (reinterpret_cast<uintptr_t>(dp_dec) & 15u))
    return RNNT_B200_E_ALIGN;
  const int n_tb = (T + DR_TT - 1) / DR_TT;
  cudaStream_t st = (cudaStream_t)stream;
  dpre_reduce_kernel<<<dim3(n_tb, B), 256, 0, st>>>((const __nv_bfloat16*)dpre_bf16, act_lens, label_lens, T, U1, H,
                                                     n_tb, dp_enc, scratch);
  RNNT_CUDA_RETURN_IF_ERROR();
  dpd_final_kernel<<<dim3(U1, B), 128, 0, st>>>(scratch, act_lens, label_lens, T, U1, H, n_tb, dp_dec);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}
