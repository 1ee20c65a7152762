// Fused joint + log-softmax statistics on the 5th-gen tensor cores (tcgen05 / TMEM / TMA).
//
// Replaces, together, /root/reference rnnt/model.py:68 (JointNet.forward, model.py:18-37)
// and the log-softmax stage of the loss call at model.py:70:
//
//   hidden   : h[cell,:] = bf16(tanh(P_enc[b,t,:] + P_dec[b,u,:])) for the valid lattice cells,
//              tile-packed (M x H, never M x V) -- one streaming pass (hidden_kernel).
//   forward  : S[cell, v] = h[cell,:] . W2[v,:] + b2[v] never leaves the SM: an online
//              log-sum-exp epilogue reads the accumulators out of TMEM and emits 12 bytes per
//              lattice cell (lse, lp_blank, lp_label).
//   backward : the same GEMM is recomputed from h and the epilogue turns it into
//              dS = scale*(softmax*(c_blank+c_label) - onehots), bf16.
//
// Work decomposition: lattice cells are flattened per utterance over the valid (t,u) window
// and cut into tiles of 128 cells. The GEMM kernel runs as CTA PAIRS (cluster of 2,
// tcgen05 cta_group::2, UMMA M = 256): the two CTAs own tiles 2*pt and 2*pt+1 and sweep the
// vocabulary in lockstep in BN-column tiles with K = H; each CTA TMA-loads only its half
// (BN/2 rows) of every W2 stage, which halves the L2->SM traffic per SM -- the measured
// limiter of the single-CTA version (profiles/r01_notes.md). Pairs are persistent: pair c
// walks pair-tiles c, c+gridDim/2, ...  Warp roles (384 threads per CTA):
//   warps 0-7   epilogue: TMEM -> registers; warp w owns TMEM lanes 32*(w&3)..+31 (one
//               lattice cell per thread) and column half (w>>2) of the BN-wide tile
//   warps 8,11  TMA producers: this CTA's half of each W2 K block (even / odd blocks)
//   warp 9      TMEM allocator (both CTAs) and MMA issuer (leader CTA, one elected thread)
//   warp 10     TMA producer of the h tile (A operand, eight 64-wide K blocks, resident
//               for the whole vocabulary sweep)
// Pipelines (mbarriers): W2 smem ring b_full (leader, fed by both CTAs' TMA) / b_empty
// (multicast tcgen05.commit); A tile a_full/a_empty per K block so the next tile's loads
// overlap the last vocabulary tile's MMAs; double-buffered TMEM accumulator acc_full
// (multicast commit) / acc_empty (16 epilogue warps of the pair arrive on the leader) so the
// epilogue of vocabulary tile j overlaps the MMAs of tile j+1.
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdio.h>
#include <type_traits>
#include "common.cuh"
#include "sm100_ptx.cuh"
#include "tma_host.cuh"


using namespace sm100;

// clock64 traces and the debug knock-outs exist only in the tuning build (common.cuh)
#ifdef RNNT_B200_TUNE
#define RNNT_TRACING(p) ((p).trace != nullptr)
#define RNNT_DBG(p) ((p).debug)
#else
#define RNNT_TRACING(p) false
#define RNNT_DBG(p) 0
#endif

namespace {

constexpr int kThreads = 384;  // 8 epilogue warps + W2 TMA (even K blocks) + MMA + A TMA + W2 TMA (odd K blocks)
constexpr int kEpiWarps = 8;
constexpr int kKBlock = 64;              // K block of both operands (128-byte swizzle span of bf16)
constexpr int kABlockBytes = RNNT_TILE_M * 128;
constexpr int kMaxSlots = 6;             // W2 ring slots (one K block of this CTA's half tile each)
constexpr int kSmemLimit = 232448;       // 227 KB opt-in maximum per CTA

struct SharedCtl {
  uint64_t a_full[8], a_empty[8];        // per group of KSUB K blocks of the h tile
  uint64_t b_full[kMaxSlots], b_empty[kMaxSlots];
  uint64_t acc_full[4], acc_empty[4];
  uint32_t tmem_base;
  uint32_t pad[3];
};
static_assert(sizeof(SharedCtl) % 16 == 0, "the regions behind ctl hold float4");

struct JointParams {
  const float* pe;
  const float* pd;
  const __nv_bfloat16* w2q;
  const float* b2l;
  const int32_t* labels;
  const int32_t* act_lens;
  const int32_t* label_lens;
  const int32_t* tile_utt;
  const int32_t* tile_cell0;
  const int32_t* n_tiles;
  // forward outputs (the backward reads them)
  float* lse;
  float* lpb;
  float* lpl;
  float* kref;   // MODE 2: the per-cell exponent reference of the stored exponentials (a plane like lse)
  int32_t* range_flag;  // MODE 2, optional: set to 1 when a cell's logits leave the fixed reference's range
  __nv_bfloat16* h_stash;
  // backward inputs/outputs
  const float* cb;
  const float* cl;
  const float* scale;
  __nv_bfloat16* ds;
  int B, T, U1, V, Vpad, Vds, H, blank;  // Vpad: rows of the packed W2 (NV*BN); Vds: dS pitch
  int tile_lo, tile_hi;  // tile range of this launch
  int slots;             // W2 ring depth
  int kts, nacc;         // V2 kernel: K blocks of the h tile multiplied out of TMEM; accumulator buffers
  long long* trace;  // RNNT_B200_TRACE: clock64 stamps of cluster 0 / rank 0 (debug only)
  int trace_cta;  // RNNT_B200_TRACE_CTA: the CTA whose stamps are recorded (default 0)
  int debug;  // RNNT_B200_DEBUG bitmask (timing experiments only): 1 skip epilogue math, 4 skip MMA issue,
              // 16 skip dS stores
};

// Shared memory of joint_kernel:
//   [h tile: H/64 K blocks x 16 KB | W2 ring: slots x (BN/2 rows x 64 k) |
//    MODE 1, 2: dS / E staging, 2 KB per epilogue warp | ctl |
//    MODE 0: red4, the cross-half (m, s, zb, zl) merge; MODE 2: kref_s, one exponent reference per row (its
//    cross-half merge goes through the staging tiles: 227 KB leave no room for red4 next to five W2 slots) |
//    bias ring: 2 halves x 2 buffers x BN/2 floats]
constexpr int kDsStageBytes = kEpiWarps * 2048;  // one 2 KB (32 cells x 32 columns) tile per epilogue warp  // one 2 KB (32 cells x 32 columns) tile per epilogue warp
constexpr int kRedBytes = RNNT_TILE_M * 16;
constexpr int kKrefBytes = RNNT_TILE_M * 4;
__host__ __device__ inline int mode_extra_bytes(int mode) {
  return mode == 0 ? kRedBytes : mode == 1 ? kDsStageBytes : kDsStageBytes + kKrefBytes;
}
// The MMA issuer consumes ksub(H) consecutive ring slots per loop iteration (two when H allows:
// one wait/fence/elect round per 8 MMAs -- the one-slot loop spent ~620 clk per K block against
// 512 clk of MMA); the slots themselves are filled and released one K block at a time.
__host__ __device__ inline int ksub_for(int H) { return (H % 128 == 0) ? 2 : 1; }
__host__ __device__ inline int smem_bytes_for(int H, int BN, int slots, int mode, int kslot = 1) {
  return (H / kKBlock) * kABlockBytes + slots * kslot * (BN / 2) * 128 + mode_extra_bytes(mode) +
         (int)sizeof(SharedCtl) + 2 * 2 * (BN / 2) * 4;
}

// tanh for hidden_kernel. (tanh.approx.f32 is one MUFU op, but its 2^-11 relative error pushed the
// per-utterance loss of one-cell utterances to 4e-5 relative -- stated tolerance 2e-5 -- for no
// measurable step-time gain.)
// Four tanh at once, tanh(x) = 1 - 2/(exp(2x)+1), with ONE reciprocal for the four denominators
// (1/d0 = d1 * (d2 d3) / (d0 d1 d2 d3), ...): 5 MUFU operations per 4 values instead of 8 for 9
// extra multiplies on the FMA pipe. x is clamped at 10 (tanh(10)
// rounds to 1.0f) so that the product of four denominators stays below 2^116.
__device__ __forceinline__ void tanh4(const float (&x)[4], float (&y)[4]) {
  float d[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) d[i] = ex2_approx((x[i] > 10.f ? 10.f : x[i]) * (2.f * RNNT_LOG2E)) + 1.f;  // NaN stays NaN
  const float p01 = d[0] * d[1], p23 = d[2] * d[3];
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(p01 * p23));
  const float r01 = r * p23, r23 = r * p01;  // 1/(d0 d1), 1/(d2 d3)
  y[0] = fmaf(-2.f, r01 * d[1], 1.f);
  y[1] = fmaf(-2.f, r01 * d[0], 1.f);
  y[2] = fmaf(-2.f, r23 * d[3], 1.f);
  y[3] = fmaf(-2.f, r23 * d[2], 1.f);
}

// z[i & 15] out of a register tile: a four-level select tree inside the rare branch that needs it (a dynamically
// indexed local array put 16 stores and a dependent load into that branch: the epilogue's stall samples sat there).
__device__ __forceinline__ float pick16(const float* z, int i) {
  const bool b0 = i & 1, b1 = i & 2, b2 = i & 4, b3 = i & 8;
  float a[8], b[4];
#pragma unroll
  for (int j = 0; j < 8; ++j) a[j] = b0 ? z[2 * j + 1] : z[2 * j];
#pragma unroll
  for (int j = 0; j < 4; ++j) b[j] = b1 ? a[2 * j + 1] : a[2 * j];
  const float c0 = b2 ? b[1] : b[0], c1 = b2 ? b[3] : b[2];
  return b3 ? c1 : c0;
}

// Two such groups of four side by side on the packed f32x2 pipe (element i of group a with element i of group b):
// the same operations in the same order and rounding as tanh4, in 43 instead of 68 floating-point instructions
// per eight values (hidden_kernel is issue bound: 77 % issue-active under ncu).
__device__ __forceinline__ void tanh4x2(const float (&xa)[4], const float (&xb)[4], float (&ya)[4], float (&yb)[4]) {
  float da[4], db[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float ta, tb;
    fmul2(ta, tb, xa[i] > 10.f ? 10.f : xa[i], xb[i] > 10.f ? 10.f : xb[i], 2.f * RNNT_LOG2E, 2.f * RNNT_LOG2E);
    fadd2(da[i], db[i], ex2_approx(ta), ex2_approx(tb), 1.f, 1.f);
  }
  float p01a, p01b, p23a, p23b, pa, pb, ra, rb;
  fmul2(p01a, p01b, da[0], db[0], da[1], db[1]);
  fmul2(p23a, p23b, da[2], db[2], da[3], db[3]);
  fmul2(pa, pb, p01a, p01b, p23a, p23b);
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ra) : "f"(pa));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rb) : "f"(pb));
  float r01a, r01b, r23a, r23b;
  fmul2(r01a, r01b, ra, rb, p23a, p23b);
  fmul2(r23a, r23b, ra, rb, p01a, p01b);
  float qa[4], qb[4];
  fmul2(qa[0], qb[0], r01a, r01b, da[1], db[1]);
  fmul2(qa[1], qb[1], r01a, r01b, da[0], db[0]);
  fmul2(qa[2], qb[2], r23a, r23b, da[3], db[3]);
  fmul2(qa[3], qb[3], r23a, r23b, da[2], db[2]);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    ya[i] = qa[i];
    yb[i] = qb[i];
    ffma2(ya[i], yb[i], -2.f, 1.f, 1.f);
  }
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t w) {
  return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&w));
}

// Locate row `r` of a tile: returns false when the row lies past the utterance's last cell.
struct RowCoord {
  int b, t, u, Ub;
  bool valid;
};
__device__ __forceinline__ RowCoord locate_row(const JointParams& p, int tile, int r, bool dead) {
  RowCoord c;
  if (dead) { c.b = 0; c.t = 0; c.u = 0; c.Ub = 0; c.valid = false; return c; }
  c.b = p.tile_utt[tile];
  const int cell = p.tile_cell0[tile] + r;
  const int Tb = clamp_int(p.act_lens[c.b], 1, p.T);
  c.Ub = clamp_int(p.label_lens[c.b], 0, p.U1 - 1);
  const int w = c.Ub + 1;
  c.valid = cell < Tb * w;
  c.t = cell / w;
  c.u = cell - c.t * w;
  return c;
}

// ---------------------------------------------------------------------------------------
// MODE 0 = forward (online LSE epilogue), MODE 1 = backward dS (recomputed logits),
// MODE 2 = forward that also STORES the exponentials: E[cell, v] = bf16(exp2(z - kref[cell])) with kref an
// integer near the row's largest logit (from the first 16 columns the row sees, + 86: every later logit may
// exceed them by 213 in log2 units = 147 nats before fp32 overflows -- the loss then comes out infinite, never
// silently wrong -- and what flushes to zero is below 2^-40 of them, under the fp32 sum's resolution), and
// sum(E) in fp32 gives the LSE with that same fixed reference. The backward then needs no second GEMM:
// dS = diag(g) . E with g = scale (c_blank + c_label) exp2(kref - lse log2 e) per cell, after patch_scale_kernel
// has overwritten the two one-hot entries of every row (csrc comment there).
// V2 (experiment, RNNT_B200_V2=<kts>; backward only, BN = 128): the first `kts` K blocks of the h
// tile are copied shared -> TMEM once per lattice tile (tcgen05.cp) and multiplied from there (TS
// form: no shared-memory read of A for them); the accumulator is `nacc` buffers of 128 columns next
// to the 32*kts columns of A; a W2 ring slot holds two K blocks (two TMA boxes, one barrier).
// ---------------------------------------------------------------------------------------
template <int BN, int MODE, int V2 = 0>
__global__ void __launch_bounds__(kThreads, 1)
joint_kernel(const __grid_constant__ CUtensorMap map_w2, const __grid_constant__ CUtensorMap map_h,
             const __grid_constant__ CUtensorMap map_ds, const JointParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];  // no static smem in this kernel: base is 1024-aligned
  if ((smem_u32(smem) & 1023u) != 0) __trap();       // the 128B-swizzle atoms need it
  const int KB = p.H / kKBlock;      // K blocks of 64
  constexpr int KSLOT = V2 ? 2 : 1;  // K blocks per W2 ring slot
  const int KSUB = V2 ? ((p.H % 256 == 0) ? 2 : 1) : ksub_for(p.H);  // ring slots per MMA-issue iteration
  const int KPI = KSUB * KSLOT;      // K blocks per MMA-issue iteration
  const int KS = KB / KPI;           // MMA-issue iterations per vocabulary tile
  const int NV = V2 ? (p.Vds + BN - 1) / BN : p.Vpad / BN;  // vocabulary tiles
  // The last vocabulary tile is issued with UMMA N = n_last (multiple of 16) instead of BN:
  // V=4232 -> 16 tiles of 256 + one of 192 (4288 columns instead of 4352).
  const int n_last = p.Vds - (NV - 1) * BN;
  constexpr int HALF = BN / 2;
  constexpr int kSubBytes = HALF * 128;   // this CTA's half of one W2 K block: BN/2 rows x 64 k
  constexpr int kSlotBytes = KSLOT * kSubBytes;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + KB * kABlockBytes;
  uint8_t* smem_ds = smem_b + p.slots * kSlotBytes;  // MODE 1, 2: 8 warps x 2 KB dS / E staging tiles
  SharedCtl* ctl = reinterpret_cast<SharedCtl*>(smem_ds + (MODE != 0 ? kDsStageBytes : 0));
  float4* red4 = reinterpret_cast<float4*>(ctl + 1);  // MODE 0 only
  float* kref_s = reinterpret_cast<float*>(ctl + 1);  // MODE 2 only
  (void)red4; (void)kref_s;
  float* bias_r = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(ctl + 1) +
                                           (MODE == 0 ? kRedBytes : MODE == 2 ? kKrefBytes : 0));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_tiles = min(p.n_tiles[0], p.tile_hi);
  // CTA pair: the two CTAs of a cluster own tiles 2*pt and 2*pt+1 and sweep the vocabulary in
  // lockstep; rank 0 (leader) issues the cta_group::2 MMAs for both.
  const uint32_t rank = cluster_ctarank();
  const int cid = blockIdx.x >> 1, ncl = gridDim.x >> 1;
  const int n_pairs = (max(n_tiles - p.tile_lo, 0) + 1) >> 1;

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&map_w2);
    tma_prefetch_desc(&map_h);
  }
  if (warp == 9 && lane == 0) {
    for (int i = 0; i < 8; ++i) {
      mbar_init(&ctl->a_full[i], 2);
      mbar_init(&ctl->a_empty[i], 1);
    }
    for (int i = 0; i < kMaxSlots; ++i) {
      mbar_init(&ctl->b_full[i], 2);
      mbar_init(&ctl->b_empty[i], 1);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(&ctl->acc_full[i], 1);
      mbar_init(&ctl->acc_empty[i], 2 * kEpiWarps);
    }
    fence_barrier_init();
  }
  __syncwarp();
  cluster_sync_all();  // barrier inits visible to the peer before any remote arrive / multicast commit
  if (warp == 9) {
    tmem_alloc_pair(&ctl->tmem_base, 512);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();  // both CTAs' TMEM allocations exist before the leader's MMAs write them
  tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp < kEpiWarps) {
    // =========================== EPILOGUE ===========================
    // Warp w reads TMEM lanes 32*(w&3).. (one lattice cell per thread) and the column half w>>2.
    const int q = warp & 3, half = warp >> 2;
    uint8_t* stage_w = smem_ds + warp * 2048;  // MODE 1: this warp's dS staging tile
    (void)stage_w;
    const int row = q * 32 + lane;
    constexpr int NCH = HALF / 16;  // 16-column chunks per half tile
    uint32_t acc_it = 0;
    uint32_t acc_buf = 0, acc_par = 0;  // V2: buffer acc_it % nacc and its phase
    uint32_t gp_hold[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};  // MODE 1: packed even chunk waiting for its odd partner
    (void)gp_hold;
    // Bias ring of this column half: buffer (acc_it & 1) holds b2*log2(e) of the current
    // vocabulary tile's HALF columns. Warp q == 0 of the half fetches the next tile's slice into
    // a register early in a tile, stores it into the other buffer, and one 128-thread named
    // barrier per vocabulary tile hands it to the four warps of the half. The sweep order is
    // vt = (cid + k) % NV for the k-th vocabulary tile this CTA processes (rotated per pair so
    // that the 74 pairs do not pull the same W2 lines out of L2 at the same moment; the online
    // LSE is order independent), across lattice tiles too, so the next slice is always vt+1.
    float* bias_h = bias_r + half * 2 * HALF;
    const bool bias_loader = (q == 0) && (lane < HALF / 4);
    float4 bias_next = make_float4(0.f, 0.f, 0.f, 0.f);
    int vt_fetch = cid % NV;
    if (cid < n_pairs) {
      if (bias_loader) {
        reinterpret_cast<float4*>(bias_h)[lane] =
            __ldg(reinterpret_cast<const float4*>(p.b2l + vt_fetch * BN + half * HALF) + lane);
        vt_fetch = (vt_fetch + 1 == NV) ? 0 : vt_fetch + 1;
        bias_next = __ldg(reinterpret_cast<const float4*>(p.b2l + vt_fetch * BN + half * HALF) + lane);
      }
      named_bar_sync(2 + half, 128);
    }
    for (int pt = cid; pt < n_pairs; pt += ncl) {
      const int tile = p.tile_lo + 2 * pt + (int)rank;
      const bool dead = tile >= n_tiles;  // odd tail: this CTA only keeps the pair's protocol going
      (void)dead;
      const RowCoord rc = locate_row(p, tile, row, dead);
      int label = -1;
      size_t pidx = 0;
      if (rc.valid) {
        pidx = plane_index(rc.b, rc.t, rc.u, p.T, p.U1);
        if (rc.u < rc.Ub) label = clamp_int(p.labels[(size_t)rc.b * (p.U1 - 1) + rc.u], 0, p.V - 1);
      }
      // forward state: running maximum and sum of the online log-sum-exp (log2 domain)
      float m = -INFINITY, s = 0.f, zb = 0.f, zl = 0.f;
      // backward: dS = exp2(z + nl) - onehots with nl = log2(scale*(c_blank+c_label)) - lse*log2(e):
      // the occupation factor rides in the exponent (one MUFU per logit, no multiply). The two
      // one-hot entries of a row are not selected out of the register tile; their exact values
      // (from the forward's lp_blank / lp_label) overwrite the staged bf16 entries instead.
      // (Walking this load chain for the next lattice tile during the current one's last
      // vocabulary tiles was tried: the tile switch is not where the time goes, and the extra
      // live registers slowed the steady state.)
      float nl = RNNT_NEG_BIG;  // invalid rows: exp2(z + nl) == 0 exactly (MODE 2: nl = -kref for valid rows)
      float kref = 0.f;         // MODE 2
      (void)kref;
      float g_blank = 0.f, g_label = 0.f;
      uint32_t sgn = 0u;  // MODE 1: a negative upstream gradient flips the sign bits of the packed exponentials
      (void)sgn;
      if (MODE == 1 && rc.valid) {
        const float sc = p.scale[rc.b];
        const float cbs = p.cb[pidx] * sc;
        const float cls = (rc.u < rc.Ub) ? p.cl[pidx] * sc : 0.f;
        const float c0 = cbs + cls;
        nl = (c0 != 0.f) ? log2f(fabsf(c0)) - p.lse[pidx] * RNNT_LOG2E : RNNT_NEG_BIG;
        sgn = (c0 < 0.f) ? 0x80008000u : 0u;
        g_blank = c0 * __expf(p.lpb[pidx]) - cbs;
        if (label >= 0) g_label = c0 * __expf(p.lpl[pidx]) - cls;
        if (label == p.blank) { g_blank -= cls; g_label = g_blank; }
      }
      for (int vi = 0; vi < NV; ++vi, ++acc_it) {
        const int vt = (vi + cid) % NV;
        const uint32_t buf = V2 ? acc_buf : (acc_it & 1);
        const long long tr0 = RNNT_TRACING(p) ? clock64() : 0;
        mbar_wait(&ctl->acc_full[buf], V2 ? acc_par : ((acc_it >> 1) & 1));
        if (V2 && ++acc_buf == (uint32_t)p.nacc) { acc_buf = 0; acc_par ^= 1; }
        const long long tr1 = RNNT_TRACING(p) ? clock64() : 0;
        tc_fence_after();
        if (bias_loader) {  // next tile's slice -> the buffer nobody reads during this tile
          reinterpret_cast<float4*>(bias_h + ((acc_it + 1) & 1) * HALF)[lane] = bias_next;
          vt_fetch = (vt_fetch + 1 == NV) ? 0 : vt_fetch + 1;
          bias_next = __ldg(reinterpret_cast<const float4*>(p.b2l + vt_fetch * BN + half * HALF) + lane);
        }
        const float* bias_c = bias_h + (acc_it & 1) * HALF;
        const uint32_t taddr0 = tmem_base + ((uint32_t)(q * 32) << 16) + buf * BN + half * HALF;
        // columns of this tile that belong to this warp's half (the last tile is narrower; the
        // UMMA N granularity makes every chunk a full 16 columns)
        const int ncol = min(HALF, max(0, ((vt == NV - 1) ? n_last : BN) - half * HALF));
        const int nch = (RNNT_DBG(p) & 1) ? 0 : ncol >> 4;
        const int colbase = vt * BN + half * HALF;
        if (MODE == 2 && vi == 0) {
          // The row's exponent reference, from the first 16 columns of column half 0 of the first vocabulary tile
          // this pair sweeps (one extra TMEM load per lattice tile), handed to the other half's thread of the row.
          if (half == 0) {
            float z0[16];
            tmem_ld_32x16(taddr0, z0);
            tmem_ld_wait();
            float cm = -INFINITY;
#pragma unroll
            for (int j = 0; j < 16; ++j) cm = fmaxf(cm, fmaf(z0[j], RNNT_LOG2E, bias_c[j]));
            // NaN / infinite logits: any reference will do, the sum carries them into the loss
            kref = (cm > -1.0e29f && cm < 1.0e29f) ? ceilf(cm) + 86.f : 0.f;
            kref_s[row] = kref;
            __threadfence_block();
            named_bar_arrive(4 + q, 64);
          } else {
            named_bar_sync(4 + q, 64);
            kref = kref_s[row];
          }
          nl = rc.valid ? -kref : RNNT_NEG_BIG;
        }
        auto release_acc = [&]() {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&ctl->acc_empty[buf]), 0));
        };
        // stage B: z = acc * log2(e) + b2 * log2(e)  (+ the chunk maximum in the forward pass)
        auto bias_stage = [&](float* zc, int c) -> float {
          const float4* b4 = reinterpret_cast<const float4*>(bias_c + c * 16);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 bb = b4[j];
            ffma2(zc[4 * j + 0], zc[4 * j + 1], RNNT_LOG2E, bb.x, bb.y);
            ffma2(zc[4 * j + 2], zc[4 * j + 3], RNNT_LOG2E, bb.z, bb.w);
          }
          if (MODE != 0) return 0.f;
          float m0 = fmaxf(zc[0], zc[4]), m1 = fmaxf(zc[1], zc[5]), m2 = fmaxf(zc[2], zc[6]), m3 = fmaxf(zc[3], zc[7]);
          m0 = fmaxf(m0, fmaxf(zc[8], zc[12])); m1 = fmaxf(m1, fmaxf(zc[9], zc[13]));
          m2 = fmaxf(m2, fmaxf(zc[10], zc[14])); m3 = fmaxf(m3, fmaxf(zc[11], zc[15]));
          return fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
        };
        // stage C on one 16-column chunk c of this half tile
        auto exp_stage = [&](const float* zc, int c, float cmax, int nchk) {
          const int col0 = colbase + c * 16;
          if (MODE == 0) {
            const float mn = fmaxf(m, cmax);
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;      // four independent sum chains (two packed pairs)
            const float nmn = -mn;
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
              float d0, d1, d2, d3;
              fadd2(d0, d1, zc[j], zc[j + 1], nmn, nmn);
              fadd2(d2, d3, zc[j + 2], zc[j + 3], nmn, nmn);
              fadd2(a0, a1, a0, a1, ex2_approx(d0), ex2_approx(d1));
              fadd2(a2, a3, a2, a3, ex2_approx(d2), ex2_approx(d3));
            }
            s = s * ex2_approx(m - mn) + ((a0 + a1) + (a2 + a3));
            m = mn;
            // the blank / label logits: picked out of the chunk's registers only in the rare chunk that
            // holds one of them (selecting with 16 compare+select pairs in EVERY chunk cost ~7 % of the
            // epilogue's issue slots)
            const bool hit_b = (unsigned)(p.blank - col0) < 16u, hit_l = (unsigned)(label - col0) < 16u;
            if (hit_b || hit_l) {
              if (hit_b) zb = pick16(zc, p.blank - col0);
              if (hit_l) zl = pick16(zc, label - col0);
            }
          } else {
            uint32_t gp[8];
            if (MODE == 1) {
#pragma unroll
              for (int j = 0; j < 8; ++j) {  // softmax * (c_blank + c_label) * scale, two columns per word
                float d0, d1;
                fadd2(d0, d1, zc[2 * j], zc[2 * j + 1], nl, nl);
                gp[j] = pack_bf16x2(ex2_approx(d0), ex2_approx(d1)) ^ sgn;
              }
            } else {
              // exp2(z - kref): summed in fp32 (the LSE with a fixed reference), stored as bf16
              float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
              for (int j = 0; j < 8; j += 2) {
                float d0, d1, d2, d3;
                fadd2(d0, d1, zc[2 * j], zc[2 * j + 1], nl, nl);
                fadd2(d2, d3, zc[2 * j + 2], zc[2 * j + 3], nl, nl);
#ifdef RNNT_B200_TUNE
                const bool no_mufu = RNNT_DBG(p) & 256;  // timing experiments: what the MUFU / the conversions cost
                const float e0 = no_mufu ? d0 : ex2_approx(d0), e1 = no_mufu ? d1 : ex2_approx(d1);
                const float e2 = no_mufu ? d2 : ex2_approx(d2), e3 = no_mufu ? d3 : ex2_approx(d3);
#else
                const float e0 = ex2_approx(d0), e1 = ex2_approx(d1), e2 = ex2_approx(d2), e3 = ex2_approx(d3);
#endif
                fadd2(a0, a1, a0, a1, e0, e1);
                fadd2(a2, a3, a2, a3, e2, e3);
#ifdef RNNT_B200_TUNE
                if (RNNT_DBG(p) & 128) {  // truncating pack on the integer pipe instead of cvt.rn.bf16x2
                  gp[j] = __byte_perm(__float_as_uint(e0), __float_as_uint(e1), 0x7632);
                  gp[j + 1] = __byte_perm(__float_as_uint(e2), __float_as_uint(e3), 0x7632);
                } else
#endif
                {
                  gp[j] = pack_bf16x2(e0, e1);
                  gp[j + 1] = pack_bf16x2(e2, e3);
                }
              }
              s += (a0 + a1) + (a2 + a3);
              const bool hit_b = (unsigned)(p.blank - col0) < 16u, hit_l = (unsigned)(label - col0) < 16u;
              if (hit_b || hit_l) {
                if (hit_b) zb = pick16(zc, p.blank - col0);
                if (hit_l) zl = pick16(zc, label - col0);
              }
            }
            // 32 cells x 32 columns of bf16 (two chunks) -> this warp's 2 KB staging tile (64-byte
            // rows, SWIZZLE_64B: conflict-free 16-byte stores) -> one TMA store (full-sector
            // writes). The tile is single buffered (the 16 KB a second one would take hold the
            // fifth W2 ring slot), so an even chunk is only packed and HELD in registers; both
            // chunks are written behind the wait on the previous store's smem read, which by then
            // is two exponential stages old. (Tried instead of the TMA store: transposed 16-byte
            // st.global from the staged tile, 6.6 ms, and one 32-byte st.global per lane and chunk
            // straight from registers, 8.7 ms, against 5.65 ms.)
            const bool last = c + 1 == nchk;
            if (RNNT_DBG(p) & 512) {  // timing experiments: no staging at all (results invalid)
            } else if ((c & 1) == 0 && !last) {
#pragma unroll
              for (int j = 0; j < 8; ++j) gp_hold[j] = gp[j];
            } else {
              if (lane == 0 && !(RNNT_DBG(p) & 32)) bulk_wait_group_read0();
              __syncwarp();
              uint8_t* srow = stage_w + lane * 64;
              const int swz = (lane >> 1) & 3;
              const int cbase = col0 - (c & 1) * 16;  // first column of the staged tile
              if (c & 1) {
                *reinterpret_cast<uint4*>(srow + ((0 ^ swz) << 4)) = make_uint4(gp_hold[0], gp_hold[1], gp_hold[2], gp_hold[3]);
                *reinterpret_cast<uint4*>(srow + ((1 ^ swz) << 4)) = make_uint4(gp_hold[4], gp_hold[5], gp_hold[6], gp_hold[7]);
              }
              *reinterpret_cast<uint4*>(srow + (((2 * (c & 1)) ^ swz) << 4)) = make_uint4(gp[0], gp[1], gp[2], gp[3]);
              *reinterpret_cast<uint4*>(srow + (((2 * (c & 1) + 1) ^ swz) << 4)) = make_uint4(gp[4], gp[5], gp[6], gp[7]);
              // the row's one-hot entries: exact values over the staged ones (same thread, program order)
              const int ncols = 16 + (c & 1) * 16;
              if (MODE == 1 && (unsigned)(p.blank - cbase) < (unsigned)ncols && rc.valid) {
                const int off = (p.blank - cbase) * 2;
                *reinterpret_cast<__nv_bfloat16*>(srow + (((off >> 4) ^ swz) << 4) + (off & 15)) = __float2bfloat16_rn(g_blank);
              }
              if (MODE == 1 && (unsigned)(label - cbase) < (unsigned)ncols) {
                const int off = (label - cbase) * 2;
                *reinterpret_cast<__nv_bfloat16*>(srow + (((off >> 4) ^ swz) << 4) + (off & 15)) = __float2bfloat16_rn(g_label);
              }
              if (!(RNNT_DBG(p) & 64)) fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0 && !dead && !(RNNT_DBG(p) & 16)) {
                tma_store_2d(&map_ds, stage_w, cbase, (tile - p.tile_lo) * RNNT_TILE_M + q * 32);
                bulk_commit_group();
              }
            }
          }
        };

        if ((NCH == 8 || (V2 && NCH == 4)) && nch == NCH) {
          // Software pipeline over the eight 16-column chunks, three register buffers deep:
          // while chunk c goes through the MUFU-bound exponential stage, chunk c+1 gets its bias
          // FMA (+ running max) in the same basic block and the TMEM load of chunk c+2 is in
          // flight. The accumulator buffer is handed back to the MMA warp as soon as its last
          // TMEM load has completed, before the last two chunks are reduced.
          float z[3][16];
          tmem_ld_32x16(taddr0, z[0]);
          tmem_ld_32x16(taddr0 + 16, z[1]);
          tmem_ld_wait();
          float cmax = bias_stage(z[0], 0);
#pragma unroll
          for (int c = 0; c < NCH; ++c) {
            if (c + 2 < NCH) tmem_ld_32x16(taddr0 + (c + 2) * 16, z[(c + 2) % 3]);
            float cmax_next = -INFINITY;
            if (c + 1 < NCH) cmax_next = bias_stage(z[(c + 1) % 3], c + 1);
            exp_stage(z[c % 3], c, cmax, NCH);
            cmax = cmax_next;
            if (c + 2 < NCH) {
              tmem_ld_wait();  // chunk c+2 has landed (it had this whole iteration to do so)
              if (c + 3 == NCH) release_acc();
            }
          }
        } else {
          // narrow last vocabulary tile, small BN, or nothing to do: one chunk at a time
          float z0[16];
          if (nch == 0) release_acc();
#pragma unroll 1
          for (int c = 0; c < nch; ++c) {
            tmem_ld_32x16(taddr0 + c * 16, z0);
            tmem_ld_wait();
            if (c + 1 == nch) release_acc();
            const float cmax = bias_stage(z0, c);
            exp_stage(z0, c, cmax, nch);
          }
        }
        named_bar_sync(2 + half, 128);  // bias ring hand-over (see above)
        if (RNNT_TRACING(p) && blockIdx.x == p.trace_cta && threadIdx.x == 0 && acc_it < 4096) {
          long long* t = p.trace + 6 * 4096;
          t[3 * acc_it + 0] = tr0; t[3 * acc_it + 1] = tr1; t[3 * acc_it + 2] = clock64();
        }
      }
      if (MODE == 0) {
        // merge the two column halves (each holds zb / zl only if it swept that column)
        if (half == 1) red4[row] = make_float4(m, s, zb, zl);
        named_bar_sync(1, kEpiWarps * 32);
        if (half == 0 && rc.valid) {
          const float4 o = red4[row];
          const float mn = fmaxf(m, o.x);
          const float st = s * ex2_approx(m - mn) + o.y * ex2_approx(o.x - mn);
          const float lg = log2f(st);
          p.lse[pidx] = (mn + lg) * RNNT_LN2;
          // (z - max) - log2(sum), not z - lse: exact difference first, so that a dominant blank / label
          // (z == max) keeps lp = -log2(sum) instead of the rounding of two numbers of size |lse|
          p.lpb[pidx] = ((zb + o.z - mn) - lg) * RNNT_LN2;
          p.lpl[pidx] = (label >= 0) ? ((zl + o.w - mn) - lg) * RNNT_LN2 : -INFINITY;
        }
        named_bar_sync(1, kEpiWarps * 32);  // red4[] is free again
      }
      if (MODE == 2) {
        // merge the two column halves through the staging tile of the half-1 warp (its last store has read it)
        if (half == 1) {
          if (lane == 0) bulk_wait_group_read0();
          __syncwarp();
          reinterpret_cast<float4*>(stage_w)[lane] = make_float4(0.f, s, zb, zl);
        }
        named_bar_sync(1, kEpiWarps * 32);
        if (half == 0 && rc.valid) {
          const float4 o = reinterpret_cast<const float4*>(smem_ds + (warp + 4) * 2048)[lane];
          const float st = s + o.y;
          const float lg = log2f(st);
          if (st < 3.0e38f) {
            p.lse[pidx] = (kref + lg) * RNNT_LN2;
            p.lpb[pidx] = ((zb + o.z - kref) - lg) * RNNT_LN2;
            p.lpl[pidx] = (label >= 0) ? ((zl + o.w - kref) - lg) * RNNT_LN2 : -INFINITY;
          } else {
            // a logit beyond the fixed reference's range (or NaN): log-probabilities of -inf would only close this
            // cell to the lattice and leave a finite, wrong loss -- +inf makes every path through it, hence the
            // utterance's loss, infinite
            if (p.range_flag) *p.range_flag = 1;
            p.lse[pidx] = INFINITY;
            p.lpb[pidx] = INFINITY;
            p.lpl[pidx] = (label >= 0) ? INFINITY : -INFINITY;
          }
          p.kref[pidx] = kref;
        }
        named_bar_sync(1, kEpiWarps * 32);  // the staging tile goes back to the next lattice tile's stores
      }
    }
    if (MODE != 0) {
      if (lane == 0) bulk_wait_group0();  // all of this warp's dS / E stores have landed
      // tiles of this launch past the end of the tile table: dS rows are defined as zero
      const int first_dead = max(n_tiles, p.tile_lo);
      const int epi_tid = threadIdx.x;  // 0..255
      for (int tile = first_dead + blockIdx.x; tile < p.tile_hi; tile += gridDim.x) {
        uint4* dst = reinterpret_cast<uint4*>(p.ds + (size_t)(tile - p.tile_lo) * RNNT_TILE_M * p.Vds);
        const size_t n16 = (size_t)RNNT_TILE_M * p.Vds / 8;
        for (size_t i = epi_tid; i < n16; i += kEpiWarps * 32) dst[i] = make_uint4(0, 0, 0, 0);
      }
    }
  } else if (warp == 8 || warp == 11) {
    // =========================== TMA PRODUCERS (W2) ===========================
    // One ring slot = this CTA's BN/2 rows of one 64-wide K block; the leader's b_full collects
    // both CTAs' bytes (2 arrivals + 2 x kSlotBytes). Two producer warps take alternate K blocks
    // (a single producer needed ~600 clk per slot -- wait, expect_tx, TMA issue -- against the
    // 512 clk the MMAs of a slot take, and starved the issuer: profiles/r01_notes.md).
    if (lane == 0) {
      const int parity = (warp == 8) ? 0 : 1;
      const uint32_t nslot = (uint32_t)p.slots;
      uint32_t slot = (uint32_t)parity % nslot, phase = ((uint32_t)parity / nslot) & 1;
      int kb = parity * KSLOT, pt = cid, vt = cid % NV, vleft = NV;  // vleft: vocabulary tiles left in this lattice tile
      while (kb >= KB) { kb -= KB; vt = (vt + 1 == NV) ? 0 : vt + 1; if (--vleft == 0) { vleft = NV; pt += ncl; } }
      int tr_n = 0;
      while (pt < n_pairs) {
        const int row0 = vt * BN + (int)rank * (((vt == NV - 1) ? n_last : BN) / 2);
        const long long tr0 = RNNT_TRACING(p) ? clock64() : 0;
        mbar_wait(&ctl->b_empty[slot], phase ^ 1);
        const long long tr1 = RNNT_TRACING(p) ? clock64() : 0;
        const uint32_t full = mapa_u32(smem_u32(&ctl->b_full[slot]), 0);
        if (rank == 0) mbar_arrive_expect_tx(&ctl->b_full[slot], 2 * kSlotBytes);
        else mbar_arrive_cluster(full);
#pragma unroll
        for (int j = 0; j < KSLOT; ++j)
          tma_load_2d_pair(smem_b + slot * kSlotBytes + j * kSubBytes, &map_w2, full, (kb + j) * kKBlock, row0);
        if (RNNT_TRACING(p) && blockIdx.x == p.trace_cta && parity == 0 && tr_n < 4096) {
          long long* t = p.trace + 3 * 4096;
          t[3 * tr_n + 0] = tr0; t[3 * tr_n + 1] = tr1; t[3 * tr_n + 2] = clock64();
          ++tr_n;
        }
        slot += 2;
        while (slot >= nslot) { slot -= nslot; phase ^= 1; }
        kb += 2 * KSLOT;
        while (kb >= KB) { kb -= KB; vt = (vt + 1 == NV) ? 0 : vt + 1; if (--vleft == 0) { vleft = NV; pt += ncl; } }
      }
    }
  } else if (warp == 10) {
    // =========================== A TILE TMA (stashed bf16 hidden tile) ===========================
    if (lane == 0) {
      int it = 0;
      for (int pt = cid; pt < n_pairs; pt += ncl, ++it) {
        const int tile = p.tile_lo + 2 * pt + (int)rank;  // a dead odd-tail tile loads zeros or stale rows: unused
        for (int ks = 0; ks < KS; ++ks) {  // one barrier pair per group of KSUB K blocks
          mbar_wait(&ctl->a_empty[ks], (it & 1) ^ 1);
          const uint32_t full = mapa_u32(smem_u32(&ctl->a_full[ks]), 0);
          if (rank == 0) mbar_arrive_expect_tx(&ctl->a_full[ks], 2 * KPI * kABlockBytes);
          else mbar_arrive_cluster(full);
          for (int sub = 0; sub < KPI; ++sub) {
            const int kb = ks * KPI + sub;
            tma_load_2d_pair(smem_a + kb * kABlockBytes, &map_h, full, kb * kKBlock, tile * RNNT_TILE_M);
          }
        }
      }
    }
  } else if (warp == 9) {
    // =========================== MMA ISSUER ===========================
    // The whole warp walks the loop with warp-uniform control flow (cheap uniform-datapath
    // descriptor arithmetic, no single-lane waterfall code); one elected lane issues. Each
    // iteration takes KSUB ring slots: 4 MMAs (K = 64) per slot, the slot's b_empty commit right
    // behind them so that the producer can refill it while the next slot is multiplied.
    if (rank == 0) {
      constexpr uint32_t idesc_full = umma_idesc_bf16(2 * RNNT_TILE_M, BN);
      const uint32_t idesc_last = umma_idesc_bf16(2 * RNNT_TILE_M, n_last);
      uint32_t slot = 0, phase = 0, acc_it = 0;
      uint32_t abuf = 0, apar = 0;  // V2: accumulator buffer acc_it % nacc and its phase
      const uint32_t tmem_a = tmem_base + (uint32_t)p.nacc * BN;  // V2: the h tile's K blocks in TMEM
      (void)tmem_a;
      int it = 0, tr_n = 0;
      const uint32_t a0 = smem_u32(smem_a), b0 = smem_u32(smem_b);
      const bool tracing = RNNT_TRACING(p) && blockIdx.x == p.trace_cta;
      for (int pt = cid; pt < n_pairs; pt += ncl, ++it) {
        for (int vi = 0; vi < NV; ++vi, ++acc_it) {
          const int vt = (vi + cid) % NV;  // same rotated order as the producer and the epilogue
          const uint32_t idesc = (vt == NV - 1) ? idesc_last : idesc_full;
          const uint32_t buf = V2 ? abuf : (acc_it & 1);
          mbar_wait(&ctl->acc_empty[buf], (V2 ? apar : ((acc_it >> 1) & 1)) ^ 1);
          if (V2 && ++abuf == (uint32_t)p.nacc) { abuf = 0; apar ^= 1; }
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + buf * BN;
          for (int ks = 0; ks < KS; ++ks) {
            if (vi == 0) mbar_wait(&ctl->a_full[ks], it & 1);
            const long long tr0 = tracing ? clock64() : 0;
            const uint32_t s0 = slot, ph0 = phase;
            if (++slot == (uint32_t)p.slots) { slot = 0; phase ^= 1; }
            uint32_t s1 = s0;
            mbar_wait(&ctl->b_full[s0], ph0);
            if (KSUB == 2) {
              s1 = slot;
              const uint32_t ph1 = phase;
              if (++slot == (uint32_t)p.slots) { slot = 0; phase ^= 1; }
              mbar_wait(&ctl->b_full[s1], ph1);
            }
            const long long tr1 = tracing ? clock64() : 0;
            tc_fence_after();
            if (elect_one_sync()) {
              for (int sub = 0; sub < KSUB; ++sub) {
                const uint32_t sl = sub ? s1 : s0;
                if (!(RNNT_DBG(p) & 4)) {
#pragma unroll
                  for (int j = 0; j < KSLOT; ++j) {
                    const int kb = (ks * KSUB + sub) * KSLOT + j;
                    const uint64_t adesc = umma_desc_kmajor<128>(a0 + kb * kABlockBytes);
                    const uint64_t bdesc = umma_desc_kmajor<128>(b0 + sl * kSlotBytes + j * kSubBytes);
                    const bool a_tmem = V2 && kb < p.kts;
                    const uint32_t ta = tmem_a + (uint32_t)kb * 32;
                    if (a_tmem && vi == 0) {  // this lattice tile's K block -> TMEM (in order with the MMAs)
#pragma unroll
                      for (int k = 0; k < kKBlock / 16; ++k) tmem_cp_128x256b_pair(ta + 8 * k, adesc + 2 * k);
                    }
#pragma unroll
                    for (int k = 0; k < kKBlock / 16; ++k) {
                      if (a_tmem) umma_bf16_ts_pair(tmem_d, ta + 8 * k, bdesc + 2 * k, idesc, (ks | sub | j | k) != 0);
                      else umma_bf16_ss_pair(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (ks | sub | j | k) != 0);
                    }
                  }
                }
                umma_commit_pair_mc(&ctl->b_empty[sl], 3);
              }
              if (vi == NV - 1) umma_commit_pair_mc(&ctl->a_empty[ks], 3);
              if (ks == KS - 1) umma_commit_pair_mc(&ctl->acc_full[buf], 3);
            }
            __syncwarp();
            if (tracing && lane == 0 && tr_n < 4096) {
              p.trace[3 * tr_n + 0] = tr0; p.trace[3 * tr_n + 1] = tr1; p.trace[3 * tr_n + 2] = clock64();
              ++tr_n;
            }
          }
        }
      }
    }
  }

  __syncwarp();  // single-lane roles: reconverge before the aligned cluster barrier
  tc_fence_before();
  cluster_sync_all();  // peer may still target this CTA's barriers / TMEM until here
  if (warp == 9) tmem_dealloc_pair(tmem_base, 512);
}

// ---------------------------------------------------------------------------------------
// h = bf16(tanh(P_enc[t] + P_dec[u])) for every tile row; the decomposed first joint layer
// (model.py:26-34 without the repeat/cat). One CTA per tile, H/8 threads per row (each 8
// consecutive k: 2x16 B loads from each addend, one 16 B store), 256/(H/8) rows per pass.
// A tile's rows are consecutive cells (t, u) of ONE utterance, u fastest: the tile table and the
// lengths are read once per tile, (t, u) advance incrementally (no division per row) and the
// P_enc slice stays in registers while t does not change (every ~(U+1)/4 passes).
// Rows past an utterance's last cell and tiles past the table are written as zeros;
// column H of the 8-wide tail is 1.0 for valid cells (the "ones" column that makes the library
// GEMM h_ext^T.dS produce d b2 next to dW2).
// (Taking the blank/label logits here as dot products with two W2 rows doubled this kernel's
// time -- 0.45 -> 1.03 ms -- and was dropped again.)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
hidden_kernel(const JointParams p, int max_tiles) {
  const int n_tiles = min(p.n_tiles[0], max_tiles);
  const int hp = rnnt_hp(p.H);
  const int cpr = p.H / 8;            // 16-byte chunks per row
  const int rows_per_pass = 256 / cpr;  // H=512: 4 rows per pass
  const int chunk = threadIdx.x % cpr;
  const int rsub = threadIdx.x / cpr;
  for (int tile = blockIdx.x; tile < max_tiles; tile += gridDim.x) {
    const bool dead = tile >= n_tiles;
    __nv_bfloat16* dst_tile = p.h_stash + (size_t)tile * RNNT_TILE_M * hp;
    int b = 0, cell0 = 0, w = 1, ncell = 0;  // dead tile: no valid cell
    if (!dead) {
      b = __ldg(p.tile_utt + tile);
      cell0 = __ldg(p.tile_cell0 + tile);
      w = clamp_int(__ldg(p.label_lens + b), 0, p.U1 - 1) + 1;
      ncell = clamp_int(__ldg(p.act_lens + b), 1, p.T) * w;
    }
    if (hp > p.H && threadIdx.x < RNNT_TILE_M)  // the ones column exists only where the library GEMM needs it
      *reinterpret_cast<uint4*>(dst_tile + (size_t)threadIdx.x * hp + p.H) =
          make_uint4(cell0 + (int)threadIdx.x < ncell ? 0x00003F80u : 0u, 0u, 0u, 0u);
    if (rsub < rows_per_pass) {
      int cell = cell0 + rsub;
      int t = cell / w, u = cell - t * w;
      int t_have = -1;
      float4 e0 = make_float4(0.f, 0.f, 0.f, 0.f), e1 = e0;
      const float4* pe4 = reinterpret_cast<const float4*>(p.pe + (size_t)b * p.T * p.H) + chunk * 2;
      const float4* pd4 = reinterpret_cast<const float4*>(p.pd + (size_t)b * p.U1 * p.H) + chunk * 2;
      const int h4 = p.H / 4;  // float4 per row of P_enc / P_dec
#pragma unroll 2
      for (int row = rsub; row < RNNT_TILE_M; row += rows_per_pass) {
        uint4 o = make_uint4(0u, 0u, 0u, 0u);
        if (cell < ncell) {
          const float4 d0 = __ldg(pd4 + (size_t)u * h4), d1 = __ldg(pd4 + (size_t)u * h4 + 1);
          if (t != t_have) {
            e0 = __ldg(pe4 + (size_t)t * h4);
            e1 = __ldg(pe4 + (size_t)t * h4 + 1);
            t_have = t;
          }
          float xa[4], xb[4];
          fadd2(xa[0], xb[0], e0.x, e1.x, d0.x, d1.x);
          fadd2(xa[1], xb[1], e0.y, e1.y, d0.y, d1.y);
          fadd2(xa[2], xb[2], e0.z, e1.z, d0.z, d1.z);
          fadd2(xa[3], xb[3], e0.w, e1.w, d0.w, d1.w);
          float ya[4], yb[4];
          tanh4x2(xa, xb, ya, yb);
          o.x = pack_bf16x2(ya[0], ya[1]);
          o.y = pack_bf16x2(ya[2], ya[3]);
          o.z = pack_bf16x2(yb[0], yb[1]);
          o.w = pack_bf16x2(yb[2], yb[3]);
        }
        *reinterpret_cast<uint4*>(dst_tile + (size_t)row * hp + chunk * 8) = o;
        cell += rows_per_pass;
        u += rows_per_pass;
        while (u >= w) { u -= w; ++t; }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// Between the lattice and the two backward GEMMs when the forward stored E (joint_kernel MODE 2):
//   dS[row, v] = c0 softmax[row, v] - [v = blank] scale c_blank - [v = label] scale c_label,   c0 = scale (c_blank + c_label)
//             = g E[row, v]   with g = c0 exp2(kref - lse log2 e)   except in the two one-hot columns,
// so those two entries of every row are overwritten with (their exact dS) / g, from the forward's fp32 lp_blank /
// lp_label -- the value the recomputing backward stages there as well -- and from then on dS = diag(g) . E:
//   dh = diag(g) (E . W2)      -> row_scale[row] = g, applied in the dh GEMM's epilogue
//   dW2 = E^T (diag(g) h)      -> h_scaled = bf16(g h), the dW2 GEMM's B operand
//   db2 = E^T g                -> the dW2 kernel's column sums, weighted with row_scale
// One CTA of four warps per tile, one warp per 32 rows: the first rows' h chunks are requested first (they do not
// depend on g), every lane then works out g and the two patches of one row, and the factors reach the scaling
// through shuffles, eight rows at a time. (Four CTAs per SM: four tiles' dependent plane reads in flight.)
// A row whose c0 is below 1e-30 (or whose g leaves the fp32 range) contributes nothing: g = 0, no patch.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
patch_scale_kernel(const JointParams p, float* __restrict__ row_scale, __nv_bfloat16* __restrict__ h_scaled) {
  const int n_tiles = min(p.n_tiles[0], p.tile_hi);
  const int cpr = p.H / 8;  // 16-byte chunks per row (16, 32 or 64)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r0 = warp * 32;
  for (int tile = p.tile_lo + blockIdx.x; tile < p.tile_hi; tile += gridDim.x) {
    const bool dead = tile >= n_tiles;
    const uint4* src = reinterpret_cast<const uint4*>(p.h_stash + ((size_t)tile * RNNT_TILE_M + r0) * p.H);
    uint4* dst = reinterpret_cast<uint4*>(h_scaled + ((size_t)tile * RNNT_TILE_M + r0) * p.H);
    uint4 hv[8][2];
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c = lane + 32 * cc;
        hv[j][cc] = (c < cpr && !dead) ? ld_stream_u4(src + (size_t)j * cpr + c) : make_uint4(0u, 0u, 0u, 0u);
      }
    float g = 0.f;
    {
      const int r = r0 + lane;
      const RowCoord rc = locate_row(p, tile, r, dead);
      if (rc.valid) {
        const size_t pidx = plane_index(rc.b, rc.t, rc.u, p.T, p.U1);
        const float sc = p.scale[rc.b];
        const float cbs = p.cb[pidx] * sc;
        const float cls = (rc.u < rc.Ub) ? p.cl[pidx] * sc : 0.f;
        const float c0 = cbs + cls;
        if (fabsf(c0) > 1.0e-30f) {
          g = c0 * exp2f(p.kref[pidx] - p.lse[pidx] * RNNT_LOG2E);
          if (!(fabsf(g) > 1.0e-36f && fabsf(g) < 1.0e36f)) g = 0.f;
        }
        if (g != 0.f) {
          int label = -1;
          if (rc.u < rc.Ub) label = clamp_int(p.labels[(size_t)rc.b * (p.U1 - 1) + rc.u], 0, p.V - 1);
          float g_blank = c0 * __expf(p.lpb[pidx]) - cbs, g_label = 0.f;
          if (label >= 0) g_label = c0 * __expf(p.lpl[pidx]) - cls;
          if (label == p.blank) { g_blank -= cls; g_label = g_blank; }
          __nv_bfloat16* erow = p.ds + ((size_t)tile * RNNT_TILE_M + r) * p.Vds;
          const float inv = 1.f / g;
          erow[p.blank] = __float2bfloat16_rn(g_blank * inv);
          if (label >= 0) erow[label] = __float2bfloat16_rn(g_label * inv);
        }
      }
      row_scale[(size_t)tile * RNNT_TILE_M + r] = g;
    }
    auto scale_rows = [&](int jbase) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float gj = __shfl_sync(0xffffffffu, g, jbase + j);
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c = lane + 32 * cc;
          if (c < cpr) {
            const uint32_t ww[4] = {hv[j][cc].x, hv[j][cc].y, hv[j][cc].z, hv[j][cc].w};
            uint32_t oo[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
              oo[k] = pack_bf16x2(gj * __uint_as_float(ww[k] << 16), gj * __uint_as_float(ww[k] & 0xffff0000u));
            dst[(size_t)(jbase + j) * cpr + c] = make_uint4(oo[0], oo[1], oo[2], oo[3]);
          }
        }
      }
    };
    scale_rows(0);
#pragma unroll 1
    for (int jb = 8; jb < 32; jb += 8) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c = lane + 32 * cc;
          hv[j][cc] = (c < cpr && !dead) ? ld_stream_u4(src + (size_t)(jb + j) * cpr + c) : make_uint4(0u, 0u, 0u, 0u);
        }
      scale_rows(jb);
    }
  }
}

// ---------------------------------------------------------------------------------------
// Plain C = A . B^T (bf16 in, fp32 out): exercises exactly the descriptors, swizzles, TMA
// maps, TMEM addressing and barrier protocol used above, with nothing else in the way.
// One CTA per 128 x BN output tile. 6 warps: 0 TMA, 1 MMA, 2-5 epilogue (+TMEM alloc).
// ---------------------------------------------------------------------------------------
template <int BN>
__global__ void __launch_bounds__(192, 1)
selftest_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                     float* __restrict__ C, int N, int K, int ts, const __nv_bfloat16* __restrict__ A) {
  // ts: 0 = both operands from shared memory; 1 = A copied shared -> TMEM with tcgen05.cp and multiplied
  // from there (the form the TS variant of joint_kernel uses); 2 = A written to TMEM from registers
  // (tcgen05.st) -- tells a TMEM operand-layout error from a tcgen05.cp descriptor error. K <= 512 for ts.
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr int STAGES = 4;
  constexpr int kAStage = RNNT_TILE_M * 128;  // 128 x 64 bf16, SW128
  constexpr int kBStage = BN * 64;            // BN x 32 bf16, SW64 (two per A stage)
  uint8_t* sa = smem;
  uint8_t* sb = smem + STAGES * kAStage;
  struct Ctl { uint64_t full[STAGES], empty[STAGES], acc; uint32_t tmem; };
  Ctl* ctl = reinterpret_cast<Ctl*>(sb + STAGES * 2 * kBStage);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * RNNT_TILE_M, n0 = blockIdx.y * BN;
  const int KBn = K / 64;
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&ctl->full[i], 1); mbar_init(&ctl->empty[i], 1); }
    mbar_init(&ctl->acc, 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(&ctl->tmem, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = ctl->tmem;
  const uint32_t tmem_a = tmem + 256;  // A operand: K/2 columns behind the accumulator
  if (ts == 2) {
    if (warp >= 2) {
      const int q = warp & 3;
      const uint4* src = reinterpret_cast<const uint4*>(A + (size_t)(m0 + q * 32 + lane) * K);
      for (int kk = 0; kk < K / 16; ++kk) {
        const uint4 v0 = src[2 * kk], v1 = src[2 * kk + 1];
        const uint32_t r[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        tmem_st_32x8(tmem_a + ((uint32_t)(q * 32) << 16) + kk * 8, r);
      }
      tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  if (warp == 0) {
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int kb = 0; kb < KBn; ++kb) {
        mbar_wait(&ctl->empty[stage], phase ^ 1);
        mbar_arrive_expect_tx(&ctl->full[stage], kAStage + 2 * kBStage);
        tma_load_2d(sa + stage * kAStage, &map_a, &ctl->full[stage], kb * 64, m0);
        tma_load_2d(sb + (stage * 2 + 0) * kBStage, &map_b, &ctl->full[stage], kb * 64, n0);
        tma_load_2d(sb + (stage * 2 + 1) * kBStage, &map_b, &ctl->full[stage], kb * 64 + 32, n0);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(RNNT_TILE_M, BN);
      uint32_t stage = 0, phase = 0;
      for (int kb = 0; kb < KBn; ++kb) {
        mbar_wait(&ctl->full[stage], phase);
        tc_fence_after();
        for (int h2 = 0; h2 < 2; ++h2) {
          const uint64_t adesc = umma_desc_kmajor<128>(smem_u32(sa + stage * kAStage) + h2 * 64);
          const uint64_t bdesc = umma_desc_kmajor<64>(smem_u32(sb + (stage * 2 + h2) * kBStage));
          for (int k = 0; k < 2; ++k) {
            const uint32_t ta = tmem_a + (uint32_t)((kb * 4 + h2 * 2 + k) * 8);
            if (ts == 1) tmem_cp_128x256b(ta, adesc + 2 * k);
            if (ts) umma_bf16_ts(tmem, ta, bdesc + 2 * k, idesc, (kb | h2 | k) != 0);
            else umma_bf16_ss(tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | h2 | k) != 0);
          }
        }
        umma_commit(&ctl->empty[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      umma_commit(&ctl->acc);
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    mbar_wait(&ctl->acc, 0);
    tc_fence_after();
    for (int c = 0; c < BN / 32; ++c) {
      float v[32];
      tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + c * 32, v);
      tmem_ld_wait();
      float* dst = C + (size_t)(m0 + row) * N + n0 + c * 32;
#pragma unroll
      for (int j = 0; j < 32; ++j) dst[j] = v[j];
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------ host side
int pick_slots(int H, int BN, int mode, int kslot = 1) {
  int s = kMaxSlots;
  while (s > 2 && smem_bytes_for(H, BN, s, mode, kslot) > kSmemLimit) --s;
  return s;
}

template <int BN, int MODE, int V2 = 0>
int launch_joint(const CUtensorMap& mw, const CUtensorMap& mh, const CUtensorMap& md, JointParams& p,
                 cudaStream_t st) {
  // the bias table goes to shared memory when it leaves room for >= 3 ring stages, else it is
  // read through L1 (very large vocabularies)
  constexpr int kslot = V2 ? 2 : 1;
  p.slots = pick_slots(p.H, BN, MODE, kslot);
  p.debug = rnnt_tune_int("RNNT_B200_DEBUG", 0);
  if (rnnt_tune_int("RNNT_B200_DEBUG_MODE", MODE) != MODE) p.debug = 0;
  {
    const int stg = rnnt_tune_int("RNNT_B200_SLOTS", 0);
    if (stg >= 2 && stg < p.slots) p.slots = stg;
  }
  const int smem = smem_bytes_for(p.H, BN, p.slots, MODE, kslot);
  if (smem > kSmemLimit) return RNNT_B200_E_SHAPE;
  cudaError_t e = cudaFuncSetAttribute(joint_kernel<BN, MODE, V2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return (int)e;
  int grid = sm_count() & ~1;
  const int ntile = p.tile_hi - p.tile_lo;
  if (ntile < 1) return RNNT_B200_OK;
  if (grid > ((ntile + 1) & ~1)) grid = (ntile + 1) & ~1;
  {
    const int g = rnnt_tune_int("RNNT_B200_GRID", 0);  // timing experiments only
    if (g >= 2 && g < grid) grid = g & ~1;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  p.trace = nullptr;
  p.trace_cta = 0;
#ifdef RNNT_B200_TUNE
  // tuning build only: clock64 stamps of one CTA, written to a file -- allocates and synchronises
  const char* trace_path = getenv("RNNT_B200_TRACE");
  if (trace_path && rnnt_tune_int("RNNT_B200_DEBUG_MODE", MODE) != MODE) trace_path = nullptr;
  p.trace_cta = rnnt_tune_int("RNNT_B200_TRACE_CTA", 0);
  const size_t trace_bytes = 9 * 4096 * sizeof(long long);
  if (trace_path) {
    if (cudaMalloc(&p.trace, trace_bytes) != cudaSuccess) p.trace = nullptr;
    else cudaMemsetAsync(p.trace, 0, trace_bytes, st);
  }
#endif
  e = cudaLaunchKernelEx(&cfg, joint_kernel<BN, MODE, V2>, mw, mh, md, p);
  if (e != cudaSuccess) return (int)e;
#ifdef RNNT_B200_TUNE
  if (p.trace) {
    long long* host = (long long*)malloc(trace_bytes);
    cudaStreamSynchronize(st);
    cudaMemcpy(host, p.trace, trace_bytes, cudaMemcpyDeviceToHost);
    FILE* f = fopen(trace_path, "wb");
    if (f) { fwrite(host, 1, trace_bytes, f); fclose(f); }
    free(host);
    cudaFree(p.trace);
  }
#endif
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}

template <int MODE>
int dispatch_joint(JointParams& p, const void* w2q, const void* h_stash, int max_tiles, cudaStream_t st) {
  {
    const int arch = rnnt_check_arch();
    if (arch != RNNT_B200_OK) return arch;
  }
  int BN = rnnt_b200_vocab_tile(p.V);
  p.kts = 0; p.nacc = 2;
  bool v2 = false;
#ifdef RNNT_B200_TUNE
  if (MODE == 1 && BN == 256 && p.H % 128 == 0) {  // experiment: A operand (partly) from TMEM, 128-column tiles
    const int e = rnnt_tune_int("RNNT_B200_V2", -1);
    if (e >= 0) {
      v2 = true;
      BN = 128;
      p.kts = e < p.H / kKBlock ? e : p.H / kKBlock;
      p.nacc = (512 - 32 * p.kts) / 128 < 4 ? (512 - 32 * p.kts) / 128 : 4;
    }
  }
#endif
  CUtensorMap mw, mh;
  int rc = make_map_bf16(&mw, w2q, p.Vpad, p.H, p.H, kKBlock, BN / 2, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_map_bf16(&mh, h_stash, (uint64_t)max_tiles * RNNT_TILE_M, p.H, rnnt_hp(p.H), kKBlock, RNNT_TILE_M,
                     CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  CUtensorMap md = mw;
  if (MODE != 0) {  // dS chunk / E (rows x Vds): 32 x 32 store boxes out of 64-byte-swizzled staging tiles
    rc = make_map_bf16(&md, p.ds, (uint64_t)(p.tile_hi - p.tile_lo) * RNNT_TILE_M, p.Vds, p.Vds, 32, 32,
                       CU_TENSOR_MAP_SWIZZLE_64B);
    if (rc) return rc;
  }
#ifdef RNNT_B200_TUNE
  if (v2) {
    if constexpr (MODE == 1) return launch_joint<128, 1, 1>(mw, mh, md, p, st);
  }
#endif
  (void)v2;
  switch (BN) {
    case 64: return launch_joint<64, MODE>(mw, mh, md, p, st);
    case 128: return launch_joint<128, MODE>(mw, mh, md, p, st);
    default: return launch_joint<256, MODE>(mw, mh, md, p, st);
  }
}

int check_common(int B, int T, int U1, int V, int H, int blank) {
  if (B <= 0 || T <= 0 || U1 <= 0 || V <= 0) return RNNT_B200_E_SHAPE;
  if (H < 64 || H > 512 || (H & 63)) return RNNT_B200_E_SHAPE;
  if (blank < 0 || blank >= V) return RNNT_B200_E_BLANK;
  if ((long long)B * T * H > 0x7fffffffLL || (long long)B * U1 * H > 0x7fffffffLL) return RNNT_B200_E_SHAPE;
  if (rnnt_b200_max_tiles(B, T, U1) < 0) return RNNT_B200_E_SHAPE;
  return RNNT_B200_OK;
}

}  // namespace

extern "C" int rnnt_b200_joint_hidden(const float* p_enc, const float* p_dec, const int32_t* act_lens,
                                      const int32_t* label_lens, const int32_t* tile_utt,
                                      const int32_t* tile_cell0, const int32_t* n_tiles, int B, int T, int U1,
                                      int H, int tile_cap, void* h_stash, void* stream) {
  if (!p_enc || !p_dec || !act_lens || !label_lens || !tile_utt || !tile_cell0 || !n_tiles || !h_stash)
    return RNNT_B200_E_NULL;
  int rc = check_common(B, T, U1, 1, H, 0);
  if (rc) return rc;
  if (((uintptr_t)p_enc & 15) || ((uintptr_t)p_dec & 15) || ((uintptr_t)h_stash & 127)) return RNNT_B200_E_ALIGN;
  if (tile_cap < 1 || tile_cap > rnnt_b200_max_tiles(B, T, U1)) return RNNT_B200_E_SHAPE;
  if (const int arch = rnnt_check_arch()) return arch;
  JointParams p = {};
  p.pe = p_enc; p.pd = p_dec; p.act_lens = act_lens; p.label_lens = label_lens;
  p.tile_utt = tile_utt; p.tile_cell0 = tile_cell0; p.n_tiles = n_tiles; p.h_stash = (__nv_bfloat16*)h_stash;
  p.B = B; p.T = T; p.U1 = U1; p.H = H;
  const int sms = sm_count();
  const int grid = tile_cap < sms * 8 ? tile_cap : sms * 8;
  hidden_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(p, tile_cap);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}

extern "C" int rnnt_b200_joint_fwd(const void* h_stash, const void* w2_bf16,
                                   const float* b2_log2, const int32_t* labels, const int32_t* act_lens,
                                   const int32_t* label_lens, const int32_t* tile_utt,
                                   const int32_t* tile_cell0, const int32_t* n_tiles, int B, int T, int U1,
                                   int V, int H, int blank, int tile_cap, float* lse, float* lp_blank,
                                   float* lp_label, void* stream) {
  if (!h_stash || !w2_bf16 || !b2_log2 || !act_lens || !label_lens || !tile_utt || !tile_cell0 ||
      !n_tiles || !lse || !lp_blank || !lp_label)
    return RNNT_B200_E_NULL;
  if (U1 > 1 && !labels) return RNNT_B200_E_NULL;
  int rc = check_common(B, T, U1, V, H, blank);
  if (rc) return rc;
  if (((uintptr_t)w2_bf16 & 127) || ((uintptr_t)b2_log2 & 15) || ((uintptr_t)h_stash & 127))
    return RNNT_B200_E_ALIGN;
  JointParams p = {};
  p.w2q = (const __nv_bfloat16*)w2_bf16; p.b2l = b2_log2; p.labels = labels; p.act_lens = act_lens;
  p.label_lens = label_lens; p.tile_utt = tile_utt; p.tile_cell0 = tile_cell0; p.n_tiles = n_tiles;
  p.lse = lse; p.lpb = lp_blank; p.lpl = lp_label;
  p.h_stash = (__nv_bfloat16*)const_cast<void*>(h_stash);
  p.B = B; p.T = T; p.U1 = U1; p.V = V; p.Vpad = rnnt_b200_vocab_pad(V); p.Vds = rnnt_b200_ds_pitch(V); p.H = H; p.blank = blank;
  if (tile_cap < 1 || tile_cap > rnnt_b200_max_tiles(B, T, U1)) return RNNT_B200_E_SHAPE;
  p.tile_lo = 0; p.tile_hi = tile_cap;
  return dispatch_joint<0>(p, w2_bf16, h_stash, p.tile_hi, (cudaStream_t)stream);
}

extern "C" int rnnt_b200_joint_fwd_store(const void* h_stash, const void* w2_bf16, const float* b2_log2,
                                         const int32_t* labels, const int32_t* act_lens,
                                         const int32_t* label_lens, const int32_t* tile_utt,
                                         const int32_t* tile_cell0, const int32_t* n_tiles, int B, int T, int U1,
                                         int V, int H, int blank, int tile_cap, float* lse, float* lp_blank,
                                         float* lp_label, float* kref, void* e_out, int32_t* range_flag,
                                         void* stream) {
  if (!h_stash || !w2_bf16 || !b2_log2 || !act_lens || !label_lens || !tile_utt || !tile_cell0 ||
      !n_tiles || !lse || !lp_blank || !lp_label || !kref || !e_out)
    return RNNT_B200_E_NULL;
  if (U1 > 1 && !labels) return RNNT_B200_E_NULL;
  int rc = check_common(B, T, U1, V, H, blank);
  if (rc) return rc;
  if (((uintptr_t)w2_bf16 & 127) || ((uintptr_t)b2_log2 & 15) || ((uintptr_t)h_stash & 127) ||
      ((uintptr_t)e_out & 15))
    return RNNT_B200_E_ALIGN;
  if (tile_cap < 1 || tile_cap > rnnt_b200_max_tiles(B, T, U1)) return RNNT_B200_E_SHAPE;
  JointParams p = {};
  p.w2q = (const __nv_bfloat16*)w2_bf16; p.b2l = b2_log2; p.labels = labels; p.act_lens = act_lens;
  p.label_lens = label_lens; p.tile_utt = tile_utt; p.tile_cell0 = tile_cell0; p.n_tiles = n_tiles;
  p.lse = lse; p.lpb = lp_blank; p.lpl = lp_label; p.kref = kref; p.range_flag = range_flag;
  p.h_stash = (__nv_bfloat16*)const_cast<void*>(h_stash);
  p.ds = (__nv_bfloat16*)e_out;
  p.B = B; p.T = T; p.U1 = U1; p.V = V; p.Vpad = rnnt_b200_vocab_pad(V); p.Vds = rnnt_b200_ds_pitch(V); p.H = H; p.blank = blank;
  p.tile_lo = 0; p.tile_hi = tile_cap;
  return dispatch_joint<2>(p, w2_bf16, h_stash, p.tile_hi, (cudaStream_t)stream);
}

// g = scale (c_blank + c_label) exp2(kref - lse log2 e) per tile row (0 for rows without a cell and for cells the
// lattice does not reach), the row's two one-hot entries of E overwritten so that g . E is dS there too, and the
// hidden rows scaled by g for the dW2 GEMM (E^T . diag(g) . h = E^T . h_scaled).
extern "C" int rnnt_b200_patch_scale(const void* h_stash, const int32_t* labels, const int32_t* act_lens,
                                     const int32_t* label_lens, const int32_t* tile_utt,
                                     const int32_t* tile_cell0, const int32_t* n_tiles, const float* lse,
                                     const float* lp_blank, const float* lp_label, const float* kref,
                                     const float* c_blank, const float* c_label, const float* scale, int B,
                                     int T, int U1, int V, int H, int blank, int tile_cap, int tile_lo,
                                     int tile_hi, int max_ctas, void* e_inout, float* row_scale, void* h_scaled,
                                     void* stream) {
  if (!h_stash || !act_lens || !label_lens || !tile_utt || !tile_cell0 || !n_tiles || !lse || !lp_blank ||
      !lp_label || !kref || !c_blank || !c_label || !scale || !e_inout || !row_scale || !h_scaled)
    return RNNT_B200_E_NULL;
  if (U1 > 1 && !labels) return RNNT_B200_E_NULL;
  int rc = check_common(B, T, U1, V, H, blank);
  if (rc) return rc;
  if (rnnt_hp(H) != H) return RNNT_B200_E_SHAPE;  // H in {128, 256, 512}: the stash has no tail columns
  if (tile_cap < 1 || tile_cap > rnnt_b200_max_tiles(B, T, U1)) return RNNT_B200_E_SHAPE;
  if (tile_lo < 0 || tile_hi < tile_lo || tile_hi > tile_cap) return RNNT_B200_E_SHAPE;
  if (((uintptr_t)h_stash & 15) || ((uintptr_t)h_scaled & 15) || ((uintptr_t)e_inout & 1)) return RNNT_B200_E_ALIGN;
  if (tile_hi == tile_lo) return RNNT_B200_OK;
  if (const int arch = rnnt_check_arch()) return arch;
  JointParams p = {};
  p.labels = labels; p.act_lens = act_lens; p.label_lens = label_lens;
  p.tile_utt = tile_utt; p.tile_cell0 = tile_cell0; p.n_tiles = n_tiles;
  p.lse = const_cast<float*>(lse); p.lpb = const_cast<float*>(lp_blank); p.lpl = const_cast<float*>(lp_label);
  p.kref = const_cast<float*>(kref);
  p.cb = c_blank; p.cl = c_label; p.scale = scale;
  p.h_stash = (__nv_bfloat16*)const_cast<void*>(h_stash);
  p.ds = (__nv_bfloat16*)e_inout;
  p.B = B; p.T = T; p.U1 = U1; p.V = V; p.Vds = rnnt_b200_ds_pitch(V); p.H = H; p.blank = blank;
  p.tile_lo = tile_lo; p.tile_hi = tile_hi;
  const int ntile = tile_hi - tile_lo;
  int grid = sm_count() * 4;  // four CTAs of 128 threads per SM
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  if (ntile < grid) grid = ntile;
  patch_scale_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(p, row_scale, (__nv_bfloat16*)h_scaled);
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}

extern "C" int rnnt_b200_joint_bwd_ds(const void* h_stash, const void* w2_bf16, const float* b2_log2,
                                      const int32_t* labels, const int32_t* act_lens,
                                      const int32_t* label_lens, const int32_t* tile_utt,
                                      const int32_t* tile_cell0, const int32_t* n_tiles, const float* lse,
                                      const float* lp_blank, const float* lp_label,
                                      const float* c_blank, const float* c_label, const float* scale, int B,
                                      int T, int U1, int V, int H, int blank, int tile_cap, int tile_lo,
                                      int tile_hi, void* ds, void* stream) {
  if (!h_stash || !w2_bf16 || !b2_log2 || !act_lens || !label_lens || !tile_utt || !tile_cell0 || !n_tiles ||
      !lse || !lp_blank || !lp_label || !c_blank || !c_label || !scale || !ds)
    return RNNT_B200_E_NULL;
  if (U1 > 1 && !labels) return RNNT_B200_E_NULL;
  int rc = check_common(B, T, U1, V, H, blank);
  if (rc) return rc;
  if (tile_cap < 1 || tile_cap > rnnt_b200_max_tiles(B, T, U1)) return RNNT_B200_E_SHAPE;
  const int max_tiles = tile_cap;
  if (tile_lo < 0 || tile_hi < tile_lo || tile_hi > max_tiles) return RNNT_B200_E_SHAPE;
  if (((uintptr_t)h_stash & 127) || ((uintptr_t)w2_bf16 & 127) || ((uintptr_t)b2_log2 & 15) ||
      ((uintptr_t)ds & 15))
    return RNNT_B200_E_ALIGN;
  if (tile_hi == tile_lo) return RNNT_B200_OK;
  JointParams p = {};
  p.b2l = b2_log2; p.labels = labels; p.act_lens = act_lens; p.label_lens = label_lens;
  p.tile_utt = tile_utt; p.tile_cell0 = tile_cell0; p.n_tiles = n_tiles;
  p.lse = const_cast<float*>(lse); p.lpb = const_cast<float*>(lp_blank); p.lpl = const_cast<float*>(lp_label);
  p.cb = c_blank; p.cl = c_label; p.scale = scale;
  p.ds = (__nv_bfloat16*)ds;
  p.B = B; p.T = T; p.U1 = U1; p.V = V; p.Vpad = rnnt_b200_vocab_pad(V); p.Vds = rnnt_b200_ds_pitch(V); p.H = H; p.blank = blank;
  p.tile_lo = tile_lo; p.tile_hi = tile_hi;
  return dispatch_joint<1>(p, w2_bf16, h_stash, max_tiles, (cudaStream_t)stream);
}

extern "C" int rnnt_b200_selftest_gemm(const void* a_bf16, const void* b_bf16, float* c, int M, int N,
                                       int K, void* stream) {
  return rnnt_b200_selftest_gemm_ts(a_bf16, b_bf16, c, M, N, K, 0, stream);
}

extern "C" int rnnt_b200_selftest_gemm_ts(const void* a_bf16, const void* b_bf16, float* c, int M, int N,
                                          int K, int ts, void* stream) {
  if (!a_bf16 || !b_bf16 || !c) return RNNT_B200_E_NULL;
  if (M <= 0 || N <= 0 || K <= 0 || (M % 128) || (N % 64) || (K % 64)) return RNNT_B200_E_SHAPE;
  if (ts < 0 || ts > 2 || (ts && K > 512)) return RNNT_B200_E_SHAPE;
  if (((uintptr_t)a_bf16 & 15)) return RNNT_B200_E_ALIGN;
  const int BN = (N % 256 == 0) ? 256 : 64;
  CUtensorMap ma, mb;
  int rc = make_map_bf16(&ma, a_bf16, M, K, K, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_map_bf16(&mb, b_bf16, N, K, K, 32, BN, CU_TENSOR_MAP_SWIZZLE_64B);
  if (rc) return rc;
  const int smem = 1024 + 4 * (128 * 128) + 4 * 2 * BN * 64 + 256;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e;
  if (BN == 256) {
    e = cudaFuncSetAttribute(selftest_gemm_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    selftest_gemm_kernel<256><<<dim3(M / 128, N / 256), 192, smem, st>>>(ma, mb, c, N, K, ts,
                                                                          (const __nv_bfloat16*)a_bf16);
  } else {
    e = cudaFuncSetAttribute(selftest_gemm_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    selftest_gemm_kernel<64><<<dim3(M / 128, N / 64), 192, smem, st>>>(ma, mb, c, N, K, ts,
                                                                        (const __nv_bfloat16*)a_bf16);
  }
  RNNT_CUDA_RETURN_IF_ERROR();
  return RNNT_B200_OK;
}
