/* rnnt_b200 -- C ABI of the B200-native RNN-T joint + transducer-loss hot path.
 *
 * This is the drop-in boundary for the two adjacent calls the reference makes at
 *   rnnt/model.py:68   logits = self.joint(enc_state, dec_state)
 *   rnnt/model.py:70   loss   = self.crit(logits, targets.int(), inputs_length.int(), targets_length.int())
 * and their backward (train.py:45). `self.crit` is warprnnt_pytorch.RNNTLoss
 * (model.py:6,59), a third-party C++/CUDA extension that is not in the reference
 * tree; the entry points below are what that extension's FFI would bind for this
 * path. Plain pointers and sizes only: no torch types, no allocation inside, no host
 * synchronisation, no environment variable read, every launch on the caller's stream. All pointers are DEVICE
 * pointers unless stated. Re-entrant: no global mutable state (the reference's
 * nn.DataParallel, train.py:177, calls from one thread per GPU).
 *
 * Return value: 0 = ok; <0 = invalid argument (RNNT_B200_E_*); >0 = cudaError_t.
 *
 * Lattice "plane" layout (workspace owned by the caller): for utterance b the cell
 * (t,u), 0<=t<T, 0<=u<U1 (U1 = U+1), lives at
 *       b*D*U1 + (t+u)*U1 + u,          D = T + U1 - 1   (anti-diagonal major)
 * so that one wavefront step reads one contiguous row. rnnt_b200_plane_elems()
 * returns B*D*U1. Cells with t >= act_lens[b] or u > label_lens[b] are never read.
 */
#ifndef RNNT_B200_H_
#define RNNT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RNNT_B200_VERSION 100

enum {
  RNNT_B200_OK = 0,
  RNNT_B200_E_NULL = -1,      /* a required pointer is NULL                      */
  RNNT_B200_E_SHAPE = -2,     /* B,T,U1,V,H out of the supported range           */
  RNNT_B200_E_BLANK = -3,     /* blank not in [0,V)                              */
  RNNT_B200_E_ALIGN = -4,     /* pointer not aligned as documented               */
  RNNT_B200_E_WORKSPACE = -5, /* workspace too small                             */
  RNNT_B200_E_DRIVER = -6,    /* cuTensorMapEncodeTiled unavailable / failed     */
  RNNT_B200_E_ARCH = -7       /* current device is not sm_100 (compute capability 10.0) */
};

int rnnt_b200_version(void);
const char* rnnt_b200_status_string(int status);

/* Number of elements of one lattice plane (see layout above). */
size_t rnnt_b200_plane_elems(int B, int T, int U1);

/* ---- Boundary 1: strict drop-in on materialised logits ---------------------------
 * replaces warprnnt_pytorch.RNNTLoss.forward on `acts` (model.py:70).
 * logits (B,T,U1,V) float32 contiguous RAW activations (the reference applies no
 * softmax, model.py:61-72); labels (B,U1-1) int32 zero padded; act_lens, label_lens
 * (B) int32. Lengths are clamped on device to [1,T] / [0,U1-1].
 * Writes lse, lp_blank, lp_label planes (float32) for the valid cells. */
int rnnt_b200_logits_to_planes(const float* logits, const int32_t* labels,
                               const int32_t* act_lens, const int32_t* label_lens,
                               int B, int T, int U1, int V, int blank,
                               float* lse, float* lp_blank, float* lp_label,
                               void* stream);

/* alpha/beta wavefront over the two log-prob planes (float64 accumulators), then the
 * occupation coefficients
 *   c_blank[t,u] = exp(alpha[t,u] + lp_blank[t,u] + beta[t+1,u] - ll)
 *   c_label[t,u] = exp(alpha[t,u] + lp_label[t,u] + beta[t,u+1] - ll)
 * and loss[b] = -ll_b (float32) plus loss64[b] (float64, nullable).
 * alpha, beta: float64 planes (workspace). U1 <= 1024. */
int rnnt_b200_lattice(const float* lp_blank, const float* lp_label,
                      const int32_t* act_lens, const int32_t* label_lens,
                      int B, int T, int U1,
                      double* alpha, double* beta,
                      float* c_blank, float* c_label,
                      float* loss, double* loss64, void* stream);

/* rnnt_b200_lattice without a caller-supplied workspace, for callers that cannot allocate (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE.
 * Same inputs, outputs and results as rnnt_b200_lattice; the float64 alpha / beta
 * planes live in a file-scope device array inside the library (2 x (RNNT_B200_STATIC_PLANE_ELEMS +
 * RNNT_B200_STATIC_MAX_B) doubles = 32 MiB + 16 KiB, held by every GPU that loads the library until exit).
 * Limits: B <= RNNT_B200_STATIC_MAX_B and rnnt_b200_plane_elems(B,T,U1) <= RNNT_B200_STATIC_PLANE_ELEMS (covers
 * B=32,T=500,U=100); beyond them RNNT_B200_E_WORKSPACE is returned and nothing is launched. Calls no cudaMalloc /
 * cudaFree or any allocating routine. The first call on a device resolves the array's address
 * (cudaGetSymbolAddress; make one call before starting a capture). All calls on a device share the array, which
 * keeps the previous call's contents (every cell is written before it is read: nothing needs clearing); calls
 * that overlap on the GPU (other streams / threads) RACE on it -- serialise them. */
#define RNNT_B200_STATIC_PLANE_ELEMS (1 << 21)
#define RNNT_B200_STATIC_MAX_B 1024
int rnnt_b200_lattice_static(const float* lp_blank, const float* lp_label,
                             const int32_t* act_lens, const int32_t* label_lens,
                             int B, int T, int U1,
                             float* c_blank, float* c_label,
                             float* loss, double* loss64, void* stream);

/* d loss / d logits for boundary 1 (the autograd backward of model.py:70):
 *   g[b,t,u,v] = scale[b] * ( exp(logit - lse)*(c_blank+c_label) - [v=blank] c_blank
 *                             - [v=labels[b,u]] c_label ),   0 outside the valid window.
 * scale (B) float32 on device = upstream grad * reduction factor. dlogits is fully
 * written (zeros outside the valid window). */
int rnnt_b200_planes_to_dlogits(const float* logits, const int32_t* labels,
                                const int32_t* act_lens, const int32_t* label_lens,
                                int B, int T, int U1, int V, int blank,
                                const float* lse, const float* c_blank, const float* c_label,
                                const float* scale, float* dlogits, void* stream);

/* ---- Boundary 2: fused joint + loss, logits never materialised --------------------
 * replaces model.py:68 + :70 together. The first joint layer is decomposed as
 *   W1.cat(e,d)+b1 = P_enc[b,t,:] + P_dec[b,u,:]   (model.py:31-32; encoder columns first)
 * and the caller supplies P_enc (B,T,H), P_dec (B,U1,H) float32 (two small library
 * GEMMs). H multiple of 64, 64 <= H <= 512. */

/* Lattice cells are flattened per utterance, row-major over the VALID window
 * (cell = t*(U_b+1)+u), and cut into tiles of 128 consecutive cells. Upper bound of
 * the tile count for given padded dims (what workspaces must be sized for): */
int rnnt_b200_max_tiles(int B, int T, int U1);

/* Builds the device tile table from the lengths (no host read-back):
 *   tile_utt[i], tile_cell0[i] for i < n_tiles[0].
 * tile_cap = rows the two tables hold (rnnt_b200_max_tiles(B,T,U1), or the exact count when the caller
 * also has the lengths on the host); writes are bounded by it. n_tiles is TWO int32:
 *   n_tiles[0] = min(tiles the lengths ask for, tile_cap)   (what every later kernel processes)
 *   n_tiles[1] = tiles the lengths ask for                  (> tile_cap: the capacity was too small) */
int rnnt_b200_build_tiles(const int32_t* act_lens, const int32_t* label_lens,
                          int B, int T, int U1, int tile_cap,
                          int32_t* tile_utt, int32_t* tile_cell0, int32_t* n_tiles,
                          void* stream);

/* W2 (V,H) float32 -> bf16 (Vpad,H) zero-padded rows; b2 (V) -> (Vpad) float32 scaled by
 * log2(e) with -1e30 in the padding (masks padded vocabulary columns).
 * Vpad = rnnt_b200_vocab_pad(V). w2t_bf16 (nullable): the transposed pack (H,Vpad) bf16, zero columns in the
 * padding -- the K-major B operand of rnnt_b200_dh_gemm (backward only). */
int rnnt_b200_vocab_pad(int V);
int rnnt_b200_vocab_tile(int V);
int rnnt_b200_pack_vocab(const float* w2, const float* b2, int V, int H,
                         void* w2_bf16, float* b2_log2, void* w2t_bf16, void* stream);

/* Row pitch (elements) of the stashed hidden activations h. H = 128, 256, 512 (the shapes
 * rnnt_b200_dw2_gemm covers; it sums d b2 itself): H. Other H: H + 32, column H holds 1.0 for valid
 * cells (columns H+1..H+7 zero, H+8.. unwritten) so that the library GEMM h_ext^T . dS yields d b2
 * in row H of its output (pitch H+32 is the fastest of the widths the library was tried with:
 * profiles/r01_notes.md). */
int rnnt_b200_h_pitch(int H);

/* Row pitch (elements) of the staged bf16 logit gradient dS: V rounded up to 64 (128-byte rows;
 * capped at Vpad), e.g. 4288 for V=4232. The last vocabulary tile of the GEMM is issued that
 * narrow (UMMA N = 192 instead of 256), so no tensor-core work is spent on the rest of the
 * Vpad padding. Columns V..pitch-1 are written as zeros. */
int rnnt_b200_ds_pitch(int V);

/* Hidden activations of the joint: h = bf16(tanh(P_enc[b,t,:] + P_dec[b,u,:])) for every valid
 * lattice cell, tile-packed (model.py:26-34 without the repeat/cat): row tile*128 + r of h_stash is
 * cell tile_cell0[tile] + r of utterance tile_utt[tile]; rows past an utterance's last cell and
 * tiles past the table are zeros; where the pitch exceeds H, column H is 1.0 for valid cells.
 * h_stash: bf16 workspace (tile_cap*128, rnnt_b200_h_pitch(H)), 128-byte aligned. It is the A
 * operand of rnnt_b200_joint_fwd / _bwd_ds (TMA) and an operand of the backward library GEMMs. */
int rnnt_b200_joint_hidden(const float* p_enc, const float* p_dec,
                           const int32_t* act_lens, const int32_t* label_lens,
                           const int32_t* tile_utt, const int32_t* tile_cell0, const int32_t* n_tiles,
                           int B, int T, int U1, int H, int tile_cap, void* h_stash, void* stream);

/* Fused forward: tcgen05/TMEM GEMM  h . W2^T  with an online log-sum-exp epilogue; writes
 * lse / lp_blank / lp_label planes only. h_stash as written by rnnt_b200_joint_hidden.
 * tile_cap: number of tiles the tile-indexed workspaces (h_stash, tile table) are sized for;
 * rnnt_b200_max_tiles(B,T,U1) when the lengths are only known on the device, or the exact
 * sum_b ceil(T_b*(U_b+1)/128) when the caller also has them on the host (skips all padded work).
 * It must not be smaller than the number of tiles actually present. */
int rnnt_b200_joint_fwd(const void* h_stash,
                        const void* w2_bf16, const float* b2_log2,
                        const int32_t* labels, const int32_t* act_lens, const int32_t* label_lens,
                        const int32_t* tile_utt, const int32_t* tile_cell0, const int32_t* n_tiles,
                        int B, int T, int U1, int V, int H, int blank, int tile_cap,
                        float* lse, float* lp_blank, float* lp_label, void* stream);

/* Fused backward, stage 1: recompute logits tiles from h_stash on the tensor cores and
 * emit dS = scale_b*(softmax*(c_blank+c_label) - onehots) as bf16 for tiles
 * [tile_lo, tile_hi) into ds (rows (tile_hi-tile_lo)*128, pitch rnnt_b200_ds_pitch(V)). Rows of
 * tiles >= n_tiles[0] and rows past an utterance's last cell are written as zeros.
 * ds must be 128-byte aligned. lse, lp_blank, lp_label are the planes rnnt_b200_joint_fwd wrote
 * (the two one-hot entries of each row are formed from lp_blank / lp_label).
 */
int rnnt_b200_joint_bwd_ds(const void* h_stash, const void* w2_bf16, const float* b2_log2,
                           const int32_t* labels, const int32_t* act_lens, const int32_t* label_lens,
                           const int32_t* tile_utt, const int32_t* tile_cell0, const int32_t* n_tiles,
                           const float* lse, const float* lp_blank, const float* lp_label,
                           const float* c_blank, const float* c_label, const float* scale,
                           int B, int T, int U1, int V, int H, int blank, int tile_cap,
                           int tile_lo, int tile_hi, void* ds, void* stream);

/* Fused backward, stage 2 (the backward of project_layer followed by the backward of tanh, rnnt/model.py:32-35
 * in reverse) for one staged dS chunk, tiles [tile_lo, tile_hi):
 *   dpre[row, j] = bf16( (sum_v dS[row, v] * W2_bf16[v, j]) * (1 - h[row, j]^2) )
 * w2t_bf16: the TRANSPOSED pack (H, Vpad) of rnnt_b200_pack_vocab.
 * tcgen05 GEMM (CTA pairs, M = 256 rows, N = H: both column halves against one dS box per K block,
 * K = ds_pitch(V)) with the tanh derivative in the epilogue. ds: as written by
 * rnnt_b200_joint_bwd_ds for the same [tile_lo, tile_hi); h_stash and
 * dpre: (tile_cap*128, H) bf16, 128-byte aligned, rows tile*128 + r (dpre rows outside [tile_lo, tile_hi) are not
 * touched). H in {128, 256, 512} (rnnt_b200_dh_gemm_supported); other H: the caller forms dpre with library GEMMs. */
/* sm_pairs (rnnt_b200_dh_gemm, rnnt_b200_dw2_gemm, rnnt_b200_dw2_splits): how many CTA pairs (= 2 SMs each) the
 * launch may occupy; <= 0 or more than the device has: all of them. Both consumers of a dS chunk walk its rows front
 * to back at the same rate per SM, so launched SIDE BY SIDE on two streams with half of the pairs each (37 + 37 on a
 * B200) one HBM read of the chunk serves both through L2 (measured: the pair of kernels 9.7 instead of 10.1 ms per
 * step at B=32,T=500,U=100). */
int rnnt_b200_dh_gemm_supported(int H);
int rnnt_b200_dh_gemm(const void* ds_bf16, const void* w2t_bf16, const void* h_stash, int V, int H,
                      int tile_cap, int tile_lo, int tile_hi, int sm_pairs, void* dpre_bf16, void* stream);

/* Fused backward, stage 3: the backward of the broadcast add that replaces the reference's repeat + cat
 * (rnnt/model.py:26-31), once per step over the whole dpre buffer:
 *   dP_enc[b,t,:] = sum_u dpre[row(b,t,u), :],   dP_dec[b,u,:] = sum_t dpre[row(b,t,u), :]     (float32)
 * DETERMINISTIC: no atomics, every sum in a fixed order (two calls give bit-identical results).
 * dp_enc (B,T,H): rows t >= act_lens[b] are NOT written (zero them once); dp_dec (B,U1,H) is fully written.
 * scratch: rnnt_b200_dpre_reduce_scratch_elems(B,T,U1,H) floats (per-8-row partial sums of dP_dec). */
size_t rnnt_b200_dpre_reduce_scratch_elems(int B, int T, int U1, int H);
int rnnt_b200_dpre_reduce(const void* dpre_bf16, const int32_t* act_lens, const int32_t* label_lens,
                          int B, int T, int U1, int H, float* scratch, float* dp_enc, float* dp_dec,
                          void* stream);

/* Fused backward, weight/bias gradient of project_layer for one staged dS chunk (rnnt/model.py:16,35):
 *   dw2_parts[s][v, j] = sum_{r in K split s} dS[r, v] * h[r, j]      (v < ds_pitch(V), j < H)
 *   db2_parts[s][v]   += sum_{r in K split s} dS[r, v]
 * for the s < rnnt_b200_dw2_splits(rows, V, H, sm_pairs) K splits of the chunk (H = 512: split s takes the 64-row
 * blocks s, s + splits, ...; H = 128, 256: contiguous ranges of blocks). dw2_parts (splits, ds_pitch(V), H) float32 is
 * stored (every element exactly once, no zeroing needed); db2_parts (splits, ds_pitch(V)) float32 must be zeroed
 * by the caller. The caller adds the partials of all chunks in slot order with rnnt_b200_sum_partials: the
 * result does not depend on the order in which CTAs finish (DETERMINISTIC).
 * ds: bf16, rows x rnnt_b200_ds_pitch(V), 128-byte aligned; h_rows: the chunk's first row of the h stash
 * (bf16, pitch rnnt_b200_h_pitch(H)); rows a multiple of 64; H in {128, 256, 512}. tcgen05 GEMM with both
 * operands MN-major; other H use the library GEMM against the ones column of the stash. */
int rnnt_b200_dw2_splits(long long rows, int V, int H, int sm_pairs);
int rnnt_b200_dw2_gemm(const void* ds_bf16, const void* h_rows, long long rows, int V, int H, int sm_pairs,
                       float* dw2_parts, float* db2_parts, void* stream);

/* ---- The backward WITHOUT the second vocabulary GEMM: the forward stores its exponentials -------------------------
 * (the backward of rnnt/model.py:35 + :70 entered from train.py:45; same results as joint_bwd_ds -> dh_gemm / dw2_gemm
 * to bf16 rounding, one of the four vocabulary-sized GEMMs of a training step gone.)
 *
 * rnnt_b200_joint_fwd_store: rnnt_b200_joint_fwd that also writes
 *   e_out[row, v] = bf16( exp2(z[row, v] log2 e - kref[cell]) )    (tile_cap*128 rows, pitch rnnt_b200_ds_pitch(V),
 *                                                                   16-byte aligned; rows without a cell: zeros)
 * and the plane kref (an integer-valued float per cell, laid out like lse: the exponent reference, 86 above the
 * largest of the first 16 logits the cell's row sees, so that later logits may exceed those by 213 log2 units
 * (147 nats) before fp32 overflows; an overflow shows as an infinite lse / loss, never as a silently wrong one --
 * a model whose logits spread wider than that takes rnnt_b200_joint_fwd and the recomputing backward, which use
 * a running maximum). lse is formed from the fp32 sum of the same
 * exponentials (fixed reference instead of a running maximum). range_flag (optional, may be NULL): an int32 the kernel
 * sets to 1 when that happens -- device memory, or pinned host memory mapped into the device's address space, which
 * the host can then read without synchronising (_api.py switches to the recomputing path for the following calls).
 *
 * rnnt_b200_patch_scale (after rnnt_b200_lattice): with g[row] = scale_b (c_blank + c_label) exp2(kref - lse log2 e),
 *   dS[row, v] = g[row] * e[row, v]  in every column but the row's two one-hot columns: those two entries of e are
 *   overwritten with (their dS, formed from lp_blank / lp_label as in joint_bwd_ds) / g[row], so that afterwards
 *   dS = diag(g) . e everywhere. Writes row_scale[tile*128 + r] = g (float32, tile_cap*128; 0 for rows without a cell
 *   and for cells with scale (c_blank + c_label) below 1e-30) and h_scaled = bf16(g * h) (same shape as h_stash).
 *   H in {128, 256, 512} (no tail columns in the stash). scale may have either sign. Works on the tiles
 *   [tile_lo, tile_hi) (pointers are those of tile 0); max_ctas > 0 bounds the grid (a caller that runs it beside
 *   other kernels; measured beside the two GEMMs of the previous chunk it did not pay: profiles/r02_notes.md).
 *
 * rnnt_b200_dh_gemm_scaled: rnnt_b200_dh_gemm with A = e (chunk rows [tile_lo*128, tile_hi*128) at e_bf16) and
 *   dpre[row, j] = bf16( row_scale[row] * (sum_v e[row, v] W2[v, j]) * (1 - h[row, j]^2) ); row_scale is indexed by
 *   the row of ALL tiles (like h_stash / dpre).
 * rnnt_b200_dw2_gemm_weighted: rnnt_b200_dw2_gemm with A = e, B = the chunk's rows of h_scaled, and
 *   db2_parts[s][v] = sum_r row_weight[r] e[r, v]; row_weight points at the chunk's first row. H = 512 only. */
int rnnt_b200_joint_fwd_store(const void* h_stash, const void* w2_bf16, const float* b2_log2,
                              const int32_t* labels, const int32_t* act_lens, const int32_t* label_lens,
                              const int32_t* tile_utt, const int32_t* tile_cell0, const int32_t* n_tiles,
                              int B, int T, int U1, int V, int H, int blank, int tile_cap,
                              float* lse, float* lp_blank, float* lp_label, float* kref, void* e_out,
                              int32_t* range_flag, void* stream);
int rnnt_b200_patch_scale(const void* h_stash, const int32_t* labels, const int32_t* act_lens,
                          const int32_t* label_lens, const int32_t* tile_utt, const int32_t* tile_cell0,
                          const int32_t* n_tiles, const float* lse, const float* lp_blank, const float* lp_label,
                          const float* kref, const float* c_blank, const float* c_label, const float* scale,
                          int B, int T, int U1, int V, int H, int blank, int tile_cap,
                          int tile_lo, int tile_hi, int max_ctas,
                          void* e_inout, float* row_scale, void* h_scaled, void* stream);
int rnnt_b200_dh_gemm_scaled(const void* e_bf16, const void* w2t_bf16, const void* h_stash, const float* row_scale,
                             int V, int H, int tile_cap, int tile_lo, int tile_hi, int sm_pairs, void* dpre_bf16,
                             void* stream);
int rnnt_b200_dw2_gemm_weighted(const void* e_bf16, const void* h_scaled_rows, const float* row_weight,
                                long long rows, int V, int H, int sm_pairs, float* dw2_parts, float* db2_parts,
                                void* stream);

/* out[i] = parts[0][i] + parts[1][i] + ... + parts[nparts-1][i], i < n, in that order. */
int rnnt_b200_sum_partials(const float* parts, int nparts, long long n, float* out, void* stream);

/* Greedy decoding, one frame for B <= 64 utterances at once (replaces the per-utterance, per-frame
 * joint + softmax + argmax + .item() of rnnt/model.py:90-93 / rnnt/search.py):
 *   pred[b] = argmax_v ( W2 . tanh(pe_t[b,:] + pd[b,:]) + b2 )[v]       (lowest index on ties)
 * pe_t: row b at pe_t + b*pe_stride (float32, e.g. the slice [:,t,:] of the (B,T,H) first-layer
 * product: pe_stride = T*H); pd (B,H), w2 (V,H), b2 (V) float32 contiguous; H a multiple of 64 in
 * [64,512]. work: B+1 64-bit words, zeroed once by the caller before the first step (every call
 * leaves them zeroed again). pred: (B) int64. fp32 arithmetic; no logits are written. */
int rnnt_b200_joint_step_argmax(const float* pe_t, long long pe_stride, const float* pd,
                                const float* w2, const float* b2, int B, int H, int V,
                                unsigned long long* work, long long* pred, void* stream);

/* Feature front end of one batch (replaces the per-utterance numpy
 * work of rnnt/dataset.py:72-109,143-155 and the Kaldi matrix decode of rnnt/kaldi_io.py:467-554):
 * Kaldi matrix payloads as they lie in the ark file -> decompress -> CMVN -> splice context
 * frames -> subsample -> zero-pad, one launch for B utterances, results identical to the numpy
 * path bit for bit.
 *   payload     device bytes; utterance b's matrix data starts at payload + desc[b].payload_off
 *               (16-byte aligned): rows x F float32 / float64 row-major, or the CM byte matrix
 *               (F x rows, column-major as Kaldi stores it).
 *   desc        device array of B descriptors (below).
 *   col_params  device floats, CM only: six per column {p0, (p25-p0)/64, p25, (p75-p25)/128,
 *               p75, (p100-p75)/63} at col_params + desc[b].col_off + 6*f (kaldi_io.py:531-549).
 *   cmvn_mean / cmvn_sd   (n_speakers, F) tables of mean and sqrt(variance) (dataset.py:72-75),
 *               float64 when cmvn_f64 != 0 else float32; row desc[b].spk, spk < 0 = no CMVN.
 *               out = float32((x - mean) / sd), evaluated in float64 when cmvn_f64 (numpy's
 *               promotion for float64 statistics or float64 features), else in float32.
 *   blk_primary / blk_override   HOST arrays of nblk <= 32 frame shifts: output block j of frame t
 *               holds raw frame t*interval + shift. The override shift wins where its source
 *               frame exists, else the primary, else zeros (dataset.py:77-98 writes the centre,
 *               then the left, then the right context blocks over one zeroed array).
 *               RNNT_B200_FEAT_NO_SHIFT marks "no writer".
 *   out         (B, t_pad, nblk*F) float32; rows at and beyond desc[b].out_rows are zeroed.
 * No allocation, no host read-back; everything on `stream`. */
#define RNNT_B200_FEAT_F32 0
#define RNNT_B200_FEAT_F64 1
#define RNNT_B200_FEAT_CM8 2
#define RNNT_B200_FEAT_NO_SHIFT (-2147483647 - 1)
typedef struct {
  long long payload_off; /* byte offset of the matrix data in `payload`, multiple of 16 */
  long long col_off;     /* CM8: offset in floats of the 6*F column parameters; else unused */
  int rows;              /* raw frames in the ark matrix */
  int format;            /* RNNT_B200_FEAT_* */
  int out_rows;          /* frames after subsampling = ceil(rows / interval) */
  int spk;               /* CMVN table row, -1 = none */
} rnnt_b200_utt_desc;
int rnnt_b200_features_batch(const void* payload, const rnnt_b200_utt_desc* desc, const float* col_params,
                             const void* cmvn_mean, const void* cmvn_sd, int cmvn_f64, int B, int F,
                             int nblk, const int* blk_primary, const int* blk_override, int interval,
                             int t_pad, float* out, void* stream);

/* Plain bf16 tensor-core GEMM used to validate the tcgen05/TMA plumbing on its own:
 * C (M,N) float32 = A (M,K) bf16 . B (N,K)^T bf16; M%128==0, N%64==0, K%64==0. */
int rnnt_b200_selftest_gemm(const void* a_bf16, const void* b_bf16, float* c,
                            int M, int N, int K, void* stream);
/* The same product with the A operand taken from tensor memory (the TS form of tcgen05.mma):
 * ts = 1 stages A shared -> TMEM with tcgen05.cp, ts = 2 writes it from registers (tcgen05.st),
 * ts = 0 is rnnt_b200_selftest_gemm. K <= 512 for ts != 0. */
int rnnt_b200_selftest_gemm_ts(const void* a_bf16, const void* b_bf16, float* c,
                               int M, int N, int K, int ts, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RNNT_B200_H_ */
