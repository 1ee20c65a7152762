"""Host-side mirror of the reference interface for the hot path.

* ``RNNTLoss`` / ``rnnt_loss``   -- boundary 1: drop-in for ``warprnnt_pytorch.RNNTLoss`` as
  constructed and called at /root/reference rnnt/model.py:59,70 (materialised logits).
* ``JointNet``                   -- same parameters / state-dict keys as rnnt/model.py:9-37.
* ``JointRNNTLoss`` / ``joint_rnnt_loss`` -- boundary 2: replaces model.py:68 + :70 in one
  call; the (B,T,U+1,V) logits are never materialised.

PyTorch is plumbing here (device memory, streams, autograd glue, two small library GEMMs
for the first joint layer); all lattice / softmax / vocabulary-GEMM work runs in the
hand-written sm_100a kernels behind include/rnnt_b200.h. There is no CPU fallback.
"""
from __future__ import annotations

import os
import threading
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib
from ._shard import (allreduce_grads_and_loss, bind_flat_grads, global_mean_loss, shard_batch,
                     shard_batch_weighted)

__all__ = ["RNNTLoss", "rnnt_loss", "JointNet", "JointRNNTLoss", "GraphedJointRNNTLoss", "joint_rnnt_loss",
           "fold_output_projections", "ProjectedJointRNNTLoss", "greedy_decode", "GreedyDecoder",
           "build_extension", "extension_path", "selftest_gemm", "KernelTimer",
           "shard_batch", "shard_batch_weighted", "global_mean_loss", "allreduce_grads_and_loss", "bind_flat_grads"]


def build_extension(verbose: bool = False) -> str:
    return _lib.build(verbose)


def extension_path() -> str:
    return _lib.SO_PATH


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


class KernelTimer:
    """Optional per-launch CUDA-event timing of this package's own kernels (bench.py's
    roofline numbers). ``with KernelTimer() as kt: ...; kt.summary()`` -> {name: [ms,...]}.
    Events are recorded on the launching (current) stream; nothing synchronises until
    ``summary()``."""
    active = None
    # kernels launched per C-ABI call
    KERNELS = {"logits_to_planes": 1, "lattice": 2, "planes_to_dlogits": 1, "build_tiles": 1,
               "pack_vocab": 2, "joint_hidden": 1, "joint_fwd": 1, "joint_bwd_ds": 1, "dh_gemm": 1,
               "dpre_reduce": 2, "sum_partials": 1, "joint_step_argmax": 1, "dw2_gemm": 1,
               "joint_fwd_store": 1, "patch_scale": 1, "dh_gemm_scaled": 1, "dw2_gemm_weighted": 1}

    def __init__(self, events: bool = True):
        self.events = events
        self.records = []
        self.launches = 0
        self.quiet = 0

    def __enter__(self):
        KernelTimer.active = self
        return self

    def __exit__(self, *exc):
        KernelTimer.active = None

    def summary(self):
        torch.cuda.synchronize()
        out = {}
        for name, e0, e1 in self.records:
            out.setdefault(name, []).append(e0.elapsed_time(e1))
        return out


def _call(name, fn, *args):
    kt = KernelTimer.active
    if kt is None:
        _lib.check(fn(*args), "rnnt_b200_" + name)
        return
    kt.launches += KernelTimer.KERNELS[name]
    if kt.events and not kt.quiet:
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(fn(*args), "rnnt_b200_" + name)
        e1.record()
        kt.records.append((name, e0, e1))
    else:
        _lib.check(fn(*args), "rnnt_b200_" + name)


class _region:
    """Times a region of the current stream under an active KernelTimer (library GEMMs; kernels that run side by
    side on two streams). ``quiet``: the own kernels launched inside are counted but not timed one by one."""

    def __init__(self, name, quiet=False):
        self.name = name
        self.quiet = quiet

    def __enter__(self):
        kt = KernelTimer.active
        self.e0 = None
        if kt is not None and kt.events:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
            if self.quiet:
                kt.quiet += 1

    def __exit__(self, *exc):
        kt = KernelTimer.active
        if self.e0 is not None and kt is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            kt.records.append((self.name, self.e0, e1))
            if self.quiet:
                kt.quiet -= 1


_side_streams = {}


def _side_stream(dev, which=0):
    """Side streams (one per device and role) for the kernels that run beside the current stream (created once: a
    new stream per backward call would cost a cudaStreamCreate each)."""
    key = (dev.type, dev.index if dev.index is not None else torch.cuda.current_device(), which)
    s = _side_streams.get(key)
    if s is None:
        s = _side_streams[key] = torch.cuda.Stream(device=dev)
    return s


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


# --------------------------------------------------------------------------- validation
def _check_loss_args(acts_shape, acts_dtype, acts_device, labels, act_lens, label_lens, what):
    """Argument checks of the loss boundary (the reference does none at model.py:70;
    upstream warp-transducer checks dtype/rank/contiguity). Raises, never falls back."""
    if len(acts_shape) != 4 and what == "acts":
        raise ValueError(f"acts must be (B,T,U+1,V), got rank {len(acts_shape)}")
    if acts_dtype != torch.float32:
        raise ValueError(f"{what} must be float32, got {acts_dtype}")
    for name, t in (("labels", labels), ("act_lens", act_lens), ("label_lens", label_lens)):
        if t.dtype != torch.int32:
            raise ValueError(f"{name} must be int32 (the reference casts with .int(), model.py:70), got {t.dtype}")
        if not t.is_contiguous():
            raise ValueError(f"{name} must be contiguous")
        if t.device != acts_device:
            raise ValueError(f"{name} is on {t.device}, expected {acts_device}")
    if acts_device.type != "cuda":
        raise RuntimeError("rnnt_b200 has no CPU path: tensors must live on a CUDA device (sm_100)")


def _validate() -> bool:
    return os.environ.get("RNNT_B200_VALIDATE", "0") == "1"


def _check_lens(B, T, U1, labels, act_lens, label_lens):
    if labels.dim() != 2 or labels.shape[0] != B or labels.shape[1] != U1 - 1:
        raise ValueError(f"labels must be (B,U)=({B},{U1 - 1}), got {tuple(labels.shape)}")
    if act_lens.shape != (B,) or label_lens.shape != (B,):
        raise ValueError("act_lens and label_lens must have shape (B,)")
    if _validate():  # opt-in: costs a device sync
        if int(act_lens.max()) > T or int(act_lens.min()) < 1:
            raise ValueError("act_lens out of range [1,T]")
        if int(label_lens.max()) > U1 - 1 or int(label_lens.min()) < 0:
            raise ValueError("label_lens out of range [0,U]")


def _reduce(loss_b: torch.Tensor, reduction: str) -> torch.Tensor:
    if reduction == "none":
        return loss_b
    if reduction == "sum":
        return loss_b.sum().reshape(1)
    if reduction == "mean":  # sum over the batch / batch size, one element (train.py:43,47)
        return (loss_b.sum() / loss_b.shape[0]).reshape(1)
    raise ValueError(f"unknown reduction {reduction!r}")


def _lattice(lpb, lpl, act_lens, label_lens, B, T, U1):
    """alpha/beta wavefront + coefficients on the caller's stream. Returns (loss_b, c_blank, c_label)."""
    L = _lib.lib()
    dev = lpb.device
    n = L.rnnt_b200_plane_elems(B, T, U1)
    alpha = torch.empty(n + B, dtype=torch.float64, device=dev)
    beta = torch.empty(n + B, dtype=torch.float64, device=dev)
    cb = torch.empty(n, dtype=torch.float32, device=dev)
    cl = torch.empty(n, dtype=torch.float32, device=dev)
    loss = torch.empty(B, dtype=torch.float32, device=dev)
    _call("lattice", L.rnnt_b200_lattice, _ptr(lpb), _ptr(lpl), _ptr(act_lens), _ptr(label_lens), B, T, U1,
                                   _ptr(alpha), _ptr(beta), _ptr(cb), _ptr(cl), _ptr(loss), None,
                                   _stream())
    return loss, cb, cl


# --------------------------------------------------------------------------- boundary 1
class _RNNTLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, acts, labels, act_lens, label_lens, blank):
        L = _lib.lib()
        B, T, U1, V = acts.shape
        dev = acts.device
        with torch.cuda.device(dev):
            n = L.rnnt_b200_plane_elems(B, T, U1)
            lse = torch.empty(n, dtype=torch.float32, device=dev)
            lpb = torch.empty(n, dtype=torch.float32, device=dev)
            lpl = torch.empty(n, dtype=torch.float32, device=dev)
            _call("logits_to_planes", L.rnnt_b200_logits_to_planes, _ptr(acts), _ptr(labels), _ptr(act_lens),
                                                    _ptr(label_lens), B, T, U1, V, blank, _ptr(lse),
                                                    _ptr(lpb), _ptr(lpl), _stream())
            loss, cb, cl = _lattice(lpb, lpl, act_lens, label_lens, B, T, U1)
        ctx.save_for_backward(acts, labels, act_lens, label_lens, lse, cb, cl)
        ctx.blank = blank
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        acts, labels, act_lens, label_lens, lse, cb, cl = ctx.saved_tensors
        L = _lib.lib()
        B, T, U1, V = acts.shape
        with torch.cuda.device(acts.device):
            scale = grad_out.to(torch.float32).contiguous()
            dlogits = torch.empty_like(acts)
            _call("planes_to_dlogits", L.rnnt_b200_planes_to_dlogits, _ptr(acts), _ptr(labels), _ptr(act_lens),
                                                     _ptr(label_lens), B, T, U1, V, ctx.blank,
                                                     _ptr(lse), _ptr(cb), _ptr(cl), _ptr(scale),
                                                     _ptr(dlogits), _stream())
        return dlogits, None, None, None, None


def rnnt_loss(acts, labels, act_lens, label_lens, blank: int = 0, reduction: str = "mean"):
    """Functional form of ``RNNTLoss``. acts (B,T,U+1,V) float32 CUDA raw logits."""
    _check_loss_args(tuple(acts.shape), acts.dtype, acts.device, labels, act_lens, label_lens, "acts")
    if not acts.is_contiguous():
        raise ValueError("acts must be contiguous")
    B, T, U1, V = acts.shape
    _check_lens(B, T, U1, labels, act_lens, label_lens)
    if not 0 <= blank < V:
        raise ValueError(f"blank {blank} not in [0,{V})")
    if U1 > 1024:
        raise ValueError("U+1 > 1024 is not supported by the lattice kernel")
    return _reduce(_RNNTLossFn.apply(acts, labels, act_lens, label_lens, int(blank)), reduction)


class RNNTLoss(nn.Module):
    """Drop-in for ``warprnnt_pytorch.RNNTLoss`` (rnnt/model.py:6,59,70).

    ``RNNTLoss()`` is no-arg constructible; ``crit(acts, labels, act_lens, label_lens)``
    returns a 1-element float32 tensor on ``acts.device`` (``reduction='mean'`` = sum over the
    batch / B; upstream's default could not be consulted offline -- SURVEY.md 9.1) carrying
    autograd to ``acts``. blank = 0 is pinned by the reference (model.py:64, decoder.py:9).
    """

    def __init__(self, blank: int = 0, reduction: str = "mean"):
        super().__init__()
        if reduction not in ("mean", "sum", "none"):
            raise ValueError(f"unknown reduction {reduction!r}")
        self.blank = blank
        self.reduction = reduction

    def forward(self, acts, labels, act_lens, label_lens):
        return rnnt_loss(acts, labels, act_lens, label_lens, self.blank, self.reduction)


# --------------------------------------------------------------------------- boundary 2
_DS_BUDGET_BYTES = int(os.environ.get("RNNT_B200_DS_BYTES", str(4 << 30)))
# Backward without the second vocabulary GEMM: the forward keeps E = bf16(exp2(z - kref)) for every cell
# (cells x ds_pitch(V) x 2 bytes, alive from forward to backward) when that fits this budget and the joint has the
# shape the stored path's kernels cover (H = 512); else the backward recomputes the logits chunk by chunk.
# RNNT_B200_BWD = "stored" | "recompute" forces one (stored: still only where the kernels cover the shape).
_E_BUDGET_BYTES = int(os.environ.get("RNNT_B200_E_BYTES", str(48 << 30)))


_range_flags = {}     # per device: a pinned int32 the storing forward sets when a cell's logits leave its range


def _range_flag(dev) -> torch.Tensor:
    key = dev.index if dev.index is not None else torch.cuda.current_device()
    f = _range_flags.get(key)
    if f is None:
        f = _range_flags[key] = torch.zeros(1, dtype=torch.int32).pin_memory()
    return f


def _store_exponentials(H: int, max_tiles: int, Vds: int, dev=None) -> bool:
    mode = os.environ.get("RNNT_B200_BWD", "auto")
    if mode == "recompute" or H != 512 or os.environ.get("RNNT_B200_DH_LIB", "0") == "1":
        return False
    if mode == "stored":
        return True
    if dev is not None and int(_range_flag(dev)[0]) != 0:
        # an earlier call on this device produced logits beyond the fixed exponent reference's range (its loss came
        # out infinite): from then on the recomputing backward, whose running-maximum log-sum-exp has no such limit.
        # A plain read of pinned host memory the kernel wrote: no synchronisation.
        if not _range_flags.get("warned"):
            _range_flags["warned"] = True
            import warnings
            warnings.warn("rnnt_b200: logits spread beyond the stored-exponentials range in an earlier step; "
                          "using the recomputing backward from now on (RNNT_B200_BWD=recompute makes this explicit)")
        return False
    return max_tiles * 128 * Vds * 2 <= _E_BUDGET_BYTES


_SPLIT_K = int(os.environ.get("RNNT_B200_SPLIT_K", "4"))   # of the dW2 library GEMM (rows are a multiple of 256)


class _JointLossFn(torch.autograd.Function):
    """(P_enc, P_dec, W2, b2) -> per-utterance loss. The fp32 logits are never materialised; where H = 512 and the
    memory budget allows, the forward keeps bf16 exponentials of them (E, 2 bytes per cell and vocabulary entry) and the
    backward needs no second vocabulary GEMM; otherwise the backward recomputes the logits
    chunk by chunk."""

    @staticmethod
    def forward(ctx, pe, pd, w2, b2, labels, act_lens, label_lens, blank, tile_cap):
        L = _lib.lib()
        B, T, H = pe.shape
        U1 = pd.shape[1]
        V = w2.shape[0]
        dev = pe.device
        need_grad = any(ctx.needs_input_grad[:4])
        with torch.cuda.device(dev):
            st = _stream()
            max_tiles = L.rnnt_b200_max_tiles(B, T, U1)
            if tile_cap is not None:   # exact tile count from host-side lengths: no padded work
                max_tiles = max(1, min(max_tiles, int(tile_cap)))
            Vpad = L.rnnt_b200_vocab_pad(V)
            tile_utt = torch.empty(max_tiles, dtype=torch.int32, device=dev)
            tile_cell0 = torch.empty(max_tiles, dtype=torch.int32, device=dev)
            n_tiles = torch.empty(2, dtype=torch.int32, device=dev)   # [tiles in the table, tiles the lengths ask for]
            _call("build_tiles", L.rnnt_b200_build_tiles, _ptr(act_lens), _ptr(label_lens), B, T, U1, max_tiles,
                  _ptr(tile_utt), _ptr(tile_cell0), _ptr(n_tiles), st)
            if _validate():   # opt-in (costs a device sync): a host_lens hint that undercounts the device lengths
                have, need = (int(v) for v in n_tiles.tolist())
                if need > have:
                    raise ValueError(f"host_lens gives {max_tiles} tiles but the device lengths need {need}")
            w2q = torch.empty(Vpad, H, dtype=torch.bfloat16, device=dev)
            b2l = torch.empty(Vpad, dtype=torch.float32, device=dev)
            # the backward's dh GEMM contracts over v: it takes W2 transposed (H, Vpad), packed in the same call
            w2t = torch.empty(H, Vpad, dtype=torch.bfloat16, device=dev) if (need_grad and H in (128, 256, 512)) else None
            _call("pack_vocab", L.rnnt_b200_pack_vocab, _ptr(w2), _ptr(b2), V, H, _ptr(w2q), _ptr(b2l), _ptr(w2t), st)
            n = L.rnnt_b200_plane_elems(B, T, U1)
            lse = torch.empty(n, dtype=torch.float32, device=dev)
            lpb = torch.empty(n, dtype=torch.float32, device=dev)
            lpl = torch.empty(n, dtype=torch.float32, device=dev)
            # bf16 hidden activations, tile-packed (M x h_pitch, never M x V): the GEMM's A operand
            h_stash = torch.empty(max_tiles * 128, L.rnnt_b200_h_pitch(H), dtype=torch.bfloat16, device=dev)
            _call("joint_hidden", L.rnnt_b200_joint_hidden, _ptr(pe), _ptr(pd), _ptr(act_lens), _ptr(label_lens),
                  _ptr(tile_utt), _ptr(tile_cell0), _ptr(n_tiles), B, T, U1, H, max_tiles, _ptr(h_stash), st)
            Vds = L.rnnt_b200_ds_pitch(V)
            e_st = kref = None
            if need_grad and _store_exponentials(H, max_tiles, Vds, dev):
                try:
                    e_st = torch.empty(max_tiles * 128, Vds, dtype=torch.bfloat16, device=dev)
                except torch.cuda.OutOfMemoryError:
                    # within the budget but not free on this device right now: the recomputing backward needs
                    # 4 GiB of staging instead (still this package's CUDA path, nothing leaves the GPU)
                    if os.environ.get("RNNT_B200_BWD", "auto") == "stored":
                        raise
                    e_st = None
            if e_st is not None:
                kref = torch.empty(n, dtype=torch.float32, device=dev)
                _call("joint_fwd_store", L.rnnt_b200_joint_fwd_store, _ptr(h_stash), _ptr(w2q), _ptr(b2l),
                      _ptr(labels), _ptr(act_lens), _ptr(label_lens), _ptr(tile_utt), _ptr(tile_cell0),
                      _ptr(n_tiles), B, T, U1, V, H, blank, max_tiles, _ptr(lse), _ptr(lpb), _ptr(lpl),
                      _ptr(kref), _ptr(e_st), _ptr(_range_flag(dev)), st)
            else:
                _call("joint_fwd", L.rnnt_b200_joint_fwd, _ptr(h_stash), _ptr(w2q), _ptr(b2l), _ptr(labels),
                                                 _ptr(act_lens), _ptr(label_lens), _ptr(tile_utt),
                                                 _ptr(tile_cell0), _ptr(n_tiles), B, T, U1, V, H, blank, max_tiles,
                                                 _ptr(lse), _ptr(lpb), _ptr(lpl), st)
            loss, cb, cl = _lattice(lpb, lpl, act_lens, label_lens, B, T, U1)
            if e_st is not None and _validate() and not bool(torch.isfinite(loss).all()):   # opt-in: a device sync
                raise FloatingPointError(
                    "rnnt_b200: non-finite loss from the forward that stores its exponentials -- the logits of a "
                    "lattice cell spread over more than ~147 nats (or the inputs hold NaN / inf); set "
                    "RNNT_B200_BWD=recompute for such a model (running-maximum log-sum-exp, no range limit)")
        if need_grad:
            ctx.save_for_backward(labels, act_lens, label_lens, tile_utt, tile_cell0, n_tiles, w2q, b2l,
                                  lse, lpb, lpl, cb, cl, h_stash, w2t, e_st, kref)
        ctx.dims = (B, T, U1, V, H, L.rnnt_b200_ds_pitch(V), max_tiles, blank)
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        (labels, act_lens, label_lens, tile_utt, tile_cell0, n_tiles, w2q, b2l, lse, lpb, lpl, cb, cl,
         h_stash, w2t, e_st, kref) = ctx.saved_tensors
        stored = e_st is not None
        B, T, U1, V, H, Vds, max_tiles, blank = ctx.dims
        L = _lib.lib()
        dev = lse.device
        with torch.cuda.device(dev):
            st = _stream()
            scale = grad_out.to(torch.float32).contiguous()
            dpe = torch.zeros(B, T, H, dtype=torch.float32, device=dev)   # rows t >= T_b stay zero
            dpd = torch.empty(B, U1, H, dtype=torch.float32, device=dev)
            hp = h_stash.shape[1]
            # dh / dW2 / db2: own tcgen05 GEMMs (csrc/dh_gemm.cu, csrc/dw2_gemm.cu) where they cover the shape,
            # else library GEMMs (dW2/db2 against the ones column of the stash)
            own_gemm = H in (128, 256, 512)          # the stash has no ones column at these H (pitch == H)
            lib_dh = (not own_gemm) or os.environ.get("RNNT_B200_DH_LIB", "0") == "1"   # "1": A/B timing only
            # dh and dW2 of a chunk run side by side (RNNT_B200_BWD_OVERLAP=0: one after the other, A/B timing only)
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            overlap = own_gemm and not lib_dh and os.environ.get("RNNT_B200_BWD_OVERLAP", "1") != "0"
            side = _side_stream(dev) if overlap else None
            half = max(1, sms // 4) if overlap else 0      # CTA pairs per consumer kernel (37 of 74 on a B200)
            chunk = max(2, min(max_tiles, _DS_BUDGET_BYTES // (128 * Vds * 2)))
            nchunk = -(-max_tiles // chunk)
            chunk = -(-max_tiles // nchunk)          # balanced chunks,
            chunk += chunk & 1                       # whole CTA-pair tiles
            # whole waves of the persistent backward kernels (one tile per SM and wave): with balanced chunks
            # every launch ends in a partial wave (AISHELL shape: 4 x 21.3 -> 4 x 22 waves); chunks of
            # k * SMs tiles leave one partial wave per step (22 + 22 + 22 + 19.3 -> 86 instead of 88)
            if chunk >= 4 * sms and os.environ.get("RNNT_B200_WAVE_CHUNK", "1") != "0":
                up = -(-chunk // sms) * sms
                if up * 128 * Vds * 2 <= max(_DS_BUDGET_BYTES, chunk * 128 * Vds * 2):
                    chunk = up
            bounds = [(lo, min(max_tiles, lo + chunk)) for lo in range(0, max_tiles, chunk)]
            if stored:
                # dS = diag(row_scale) . E once the two one-hot entries of every row are patched (csrc/joint_tc.cu,
                # patch_scale_kernel): no second vocabulary GEMM, the consumers read E chunk by chunk
                row_scale = torch.empty(max_tiles * 128, dtype=torch.float32, device=dev)
                h_scaled = torch.empty_like(h_stash)

                def patch(lo, hi, max_ctas, stream):
                    _call("patch_scale", L.rnnt_b200_patch_scale, _ptr(h_stash), _ptr(labels), _ptr(act_lens),
                          _ptr(label_lens), _ptr(tile_utt), _ptr(tile_cell0), _ptr(n_tiles), _ptr(lse), _ptr(lpb),
                          _ptr(lpl), _ptr(kref), _ptr(cb), _ptr(cl), _ptr(scale), B, T, U1, V, H, blank, max_tiles,
                          lo, hi, max_ctas, _ptr(e_st), _ptr(row_scale), _ptr(h_scaled), stream)
            else:
                ds = torch.empty(chunk * 128, Vds, dtype=torch.bfloat16, device=dev)
            # dpre = dh * (1 - h^2), bf16, all M rows (M x H, never M x V): written chunk by chunk, reduced once
            dpre = torch.empty(max_tiles * 128, H, dtype=torch.bfloat16, device=dev)
            if own_gemm:
                # split-K / per-chunk partial results, added in slot order at the end (deterministic)
                nsplit = [L.rnnt_b200_dw2_splits((hi - lo) * 128, V, H, half) for lo, hi in bounds]
                dw2_parts = torch.empty(sum(nsplit), Vds, H, dtype=torch.float32, device=dev)
                db2_parts = torch.zeros(sum(nsplit), Vds, dtype=torch.float32, device=dev)
            else:
                dw2x = torch.zeros(hp, Vds, dtype=torch.float32, device=dev)  # [dW2 | db2 | unused rows]^T
            w2v = w2q[:Vds]                          # rows Vds..Vpad of the pack are padding only
            slot = 0
            if stored:
                patch(0, max_tiles, 0, st)
            for ci, (lo, hi) in enumerate(bounds):
                rows = (hi - lo) * 128
                if stored:
                    dsc = e_st[lo * 128: hi * 128]
                else:
                    dsc = ds[:rows]
                    _call("joint_bwd_ds", L.rnnt_b200_joint_bwd_ds, _ptr(h_stash), _ptr(w2q), _ptr(b2l), _ptr(labels),
                                                        _ptr(act_lens), _ptr(label_lens), _ptr(tile_utt),
                                                        _ptr(tile_cell0), _ptr(n_tiles), _ptr(lse), _ptr(lpb),
                                                        _ptr(lpl), _ptr(cb), _ptr(cl), _ptr(scale), B, T, U1, V, H, blank, max_tiles, lo, hi,
                                                        _ptr(dsc), st)
                if lib_dh:
                    with _region("lib_gemm_dh"):
                        dh = torch.mm(dsc, w2v)                                        # (rows, H) bf16
                    hf = h_stash[lo * 128: hi * 128, :H].float()
                    dpre[lo * 128: hi * 128] = (dh.float() * (1.0 - hf * hf)).to(torch.bfloat16)
                else:
                    if overlap:
                        # The two consumers of the dS chunk side by side, each on half of the CTA pairs: both walk
                        # the chunk's rows front to back at the same rate, so one HBM read of dS serves both through
                        # L2 (the pair 9.7 instead of 10.1 ms per step at the AISHELL shape: profiles/r02_notes.md).
                        pair_region = _region("dh_dw2_pair", quiet=True)
                        pair_region.__enter__()
                        fork = torch.cuda.Event()
                        fork.record()
                        side.wait_event(fork)
                    if stored:
                        _call("dh_gemm_scaled", L.rnnt_b200_dh_gemm_scaled, _ptr(dsc), _ptr(w2t), _ptr(h_stash),
                              _ptr(row_scale), V, H, max_tiles, lo, hi, half, _ptr(dpre), st)
                    else:
                        _call("dh_gemm", L.rnnt_b200_dh_gemm, _ptr(dsc), _ptr(w2t), _ptr(h_stash), V, H, max_tiles,
                              lo, hi, half, _ptr(dpre), st)
                if own_gemm:
                    if stored:
                        _call("dw2_gemm_weighted", L.rnnt_b200_dw2_gemm_weighted, _ptr(dsc),
                              h_scaled.data_ptr() + lo * 128 * hp * 2, row_scale.data_ptr() + lo * 128 * 4,
                              rows, V, H, half, dw2_parts.data_ptr() + slot * Vds * H * 4,
                              db2_parts.data_ptr() + slot * Vds * 4, side.cuda_stream if overlap else st)
                    else:
                        _call("dw2_gemm", L.rnnt_b200_dw2_gemm, _ptr(dsc), h_stash.data_ptr() + lo * 128 * hp * 2,
                              rows, V, H, half, dw2_parts.data_ptr() + slot * Vds * H * 4,
                              db2_parts.data_ptr() + slot * Vds * 4, side.cuda_stream if overlap else st)
                    slot += nsplit[ci]
                    if overlap:   # the next chunk's dS overwrites the staging buffer: join first
                        join = torch.cuda.Event()
                        join.record(side)
                        torch.cuda.current_stream().wait_event(join)
                        pair_region.__exit__(None, None, None)
                else:
                    hc = h_stash[lo * 128: hi * 128]
                    with _region("lib_gemm_dw2"):
                        # K = rows is huge and the output (pitch x Vds) small: split K (batched GEMM;
                        # measured best, profiles/r01_notes.md). Column H of the stash is 1 for valid
                        # cells, so row H of the product is db2; the stash pitch (H+32) and this
                        # orientation (h^T . dS) are the fastest of the shapes the library was tried
                        # with (tools/dw2_gemm_bench.py). Rows H+8.. of the product are never read.
                        dw2x += torch.bmm(hc.view(_SPLIT_K, rows // _SPLIT_K, hp).transpose(1, 2),
                                          dsc.view(_SPLIT_K, rows // _SPLIT_K, Vds),
                                          out_dtype=torch.float32).sum(0)
            scratch = torch.empty(L.rnnt_b200_dpre_reduce_scratch_elems(B, T, U1, H), dtype=torch.float32, device=dev)
            _call("dpre_reduce", L.rnnt_b200_dpre_reduce, _ptr(dpre), _ptr(act_lens), _ptr(label_lens), B, T, U1, H,
                  _ptr(scratch), _ptr(dpe), _ptr(dpd), st)
            if own_gemm:
                dw2f = torch.empty(Vds, H, dtype=torch.float32, device=dev)
                db2f = torch.empty(Vds, dtype=torch.float32, device=dev)
                _call("sum_partials", L.rnnt_b200_sum_partials, _ptr(dw2_parts), sum(nsplit), Vds * H, _ptr(dw2f), st)
                _call("sum_partials", L.rnnt_b200_sum_partials, _ptr(db2_parts), sum(nsplit), Vds, _ptr(db2f), st)
                dw2, db2 = dw2f[:V], db2f[:V]
            else:
                dw2 = dw2x[:H, :V].t().contiguous()
                db2 = dw2x[H, :V].contiguous()
        return dpe, dpd, dw2, db2, None, None, None, None, None


def _tile_count(host_lens, T: int, U1: int) -> Optional[int]:
    """Exact tile count from host-side lengths, with the clamps the device applies (t in [1,T], u in [0,U1-1])."""
    if host_lens is None:
        return None
    tl, ul = host_lens
    return int(sum((min(max(int(t), 1), T) * (min(max(int(u), 0), U1 - 1) + 1) + 127) // 128 for t, u in zip(tl, ul)))


_tf32_lock = threading.Lock()


class _FirstLayer(torch.autograd.Function):
    """``x . w^T (+ b)`` of the decomposed first joint layer (model.py:13,32). Forward in full fp32 (its
    result is rounded to bf16 behind a tanh: the oracle comparison of the quantised path allows nothing
    less). Backward on the TF32 tensor cores: the incoming gradient was formed from bf16 dS / dh (relative
    error ~3e-3, stated tolerance 1e-2), 10-bit operand mantissas with fp32 accumulation add ~5e-4 to it,
    and the library's fp32 SIMT kernels for these four GEMMs were 0.4 ms of a 23 ms step."""

    @staticmethod
    def forward(ctx, x, w, b):
        ctx.save_for_backward(x, w)
        ctx.has_bias = b is not None
        return F.linear(x, w, b)

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        g2 = g.reshape(-1, g.shape[-1])
        gx = gw = gb = None
        # the precision switch is process-global in torch: held under a lock so that two backward passes (autograd
        # runs one worker thread per device) cannot interleave their save/restore ("Numerics")
        with _tf32_lock:
            prev = torch.backends.cuda.matmul.allow_tf32
            torch.backends.cuda.matmul.allow_tf32 = os.environ.get("RNNT_B200_TF32_BWD", "1") != "0"
            try:
                if ctx.needs_input_grad[0]:
                    gx = torch.matmul(g2, w).reshape(x.shape)
                if ctx.needs_input_grad[1]:
                    gw = torch.matmul(g2.t(), x.reshape(-1, x.shape[-1]))
            finally:
                torch.backends.cuda.matmul.allow_tf32 = prev
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = g2.sum(0)
        return gx, gw, gb


def joint_rnnt_loss(enc_state, dec_state, w1, b1, w2, b2, labels, act_lens, label_lens,
                    blank: int = 0, reduction: str = "mean", host_lens=None):
    """Fused replacement of rnnt/model.py:68 + :70.

    enc_state (B,T,E), dec_state (B,U+1,D) float32 CUDA; w1 (H,E+D) = forward_layer.weight
    (encoder columns first, model.py:31), b1 (H), w2 (V,H) = project_layer.weight, b2 (V).
    Autograd reaches enc_state, dec_state and all four joint parameters.
    ``host_lens=(act_lens_cpu, label_lens_cpu)`` (optional, e.g. straight from the DataLoader as in
    train.py:24-31) lets the op size its tile loops exactly instead of for the padded (B,T,U+1)
    bound -- no device read-back is ever made to obtain it.
    """
    if enc_state.dim() != 3 or dec_state.dim() != 3:
        raise ValueError("enc_state must be (B,T,E) and dec_state (B,U+1,D)")
    _check_loss_args((), enc_state.dtype, enc_state.device, labels, act_lens, label_lens, "enc_state")
    B, T, E = enc_state.shape
    if dec_state.shape[0] != B:
        raise ValueError("enc_state and dec_state disagree on the batch size")
    U1, D = dec_state.shape[1], dec_state.shape[2]
    H, V = w1.shape[0], w2.shape[0]
    if w1.shape[1] != E + D:
        raise ValueError(f"forward_layer.weight is (H,{w1.shape[1]}) but E+D = {E + D}")
    if w2.shape[1] != H or b1.shape != (H,) or b2.shape != (V,):
        raise ValueError("joint parameter shapes are inconsistent")
    if H % 64 or not 64 <= H <= 512:
        raise ValueError(f"joint inner size {H} unsupported: need a multiple of 64 in [64,512]")
    _check_lens(B, T, U1, labels, act_lens, label_lens)
    if not 0 <= blank < V:
        raise ValueError(f"blank {blank} not in [0,{V})")
    if U1 > 1024:
        raise ValueError("U+1 > 1024 is not supported by the lattice kernel")
    # first joint layer, decomposed (two small library GEMMs instead of one over B*T*(U+1) rows)
    pe = _FirstLayer.apply(enc_state, w1[:, :E], b1).contiguous()
    pd = _FirstLayer.apply(dec_state, w1[:, E:], None).contiguous()
    loss_b = _JointLossFn.apply(pe, pd, w2.contiguous(), b2.contiguous(), labels, act_lens, label_lens,
                                int(blank), _tile_count(host_lens, T, U1))
    return _reduce(loss_b, reduction)


def fold_output_projections(enc_proj_w, enc_proj_b, dec_proj_w, dec_proj_b, w1, b1):
    """SURVEY.md section 8(f) rank 1. The encoder's and the decoder's ``output_proj``
    (rnnt/encoder.py:18-20,38; rnnt/decoder.py:19,42) feed the joint's ``forward_layer``
    (rnnt/model.py:13,32) with no nonlinearity in between, so

        forward_layer([enc_proj(x_t); dec_proj(y_u)]) = (W1e Weo) x_t + (W1d Wdo) y_u
                                                        + (b1 + W1e beo + W1d bdo)

    Returns ``(w_enc (H,He), w_dec (H,Hd), bias (H))``: one GEMM per side straight from the LSTM
    outputs instead of two, and no (B,T,E) / (B,U+1,D) round trip through HBM. Plain torch ops:
    autograd carries the gradients back to all six parameter tensors."""
    E = enc_proj_w.shape[0]
    if w1.shape[1] != E + dec_proj_w.shape[0]:
        raise ValueError("forward_layer.weight does not match the two output projections")
    w1e, w1d = w1[:, :E], w1[:, E:]
    return w1e @ enc_proj_w, w1d @ dec_proj_w, b1 + w1e @ enc_proj_b + w1d @ dec_proj_b


class ProjectedJointRNNTLoss(nn.Module):
    """``ProjectedJointRNNTLoss(joint, enc_proj, dec_proj)(enc_rnn_out, dec_rnn_out, labels,
    act_lens, label_lens) -> loss`` where ``enc_rnn_out`` (B,T,He) / ``dec_rnn_out`` (B,U+1,Hd)
    are the LSTM outputs BEFORE ``output_proj`` and ``enc_proj`` / ``dec_proj`` are those
    ``nn.Linear`` modules (``model.encoder.output_proj``, ``model.decoder.output_proj`` of the
    reference). Same value as ``JointRNNTLoss(joint)(enc_proj(enc_rnn_out),
    dec_proj(dec_rnn_out), ...)`` up to fp32 reassociation of the first layer."""

    def __init__(self, joint: nn.Module, enc_proj: nn.Linear, dec_proj: nn.Linear, blank: int = 0,
                 reduction: str = "mean"):
        super().__init__()
        self.joint, self.enc_proj, self.dec_proj = joint, enc_proj, dec_proj
        self.blank, self.reduction = blank, reduction

    def forward(self, enc_rnn_out, dec_rnn_out, labels, act_lens, label_lens, host_lens=None):
        j = self.joint
        w_enc, w_dec, bias = fold_output_projections(
            self.enc_proj.weight, self.enc_proj.bias, self.dec_proj.weight, self.dec_proj.bias,
            j.forward_layer.weight, j.forward_layer.bias)
        # [w_enc | w_dec] plays forward_layer.weight for inputs of widths He and Hd
        return joint_rnnt_loss(enc_rnn_out, dec_rnn_out, torch.cat((w_enc, w_dec), dim=1), bias,
                               j.project_layer.weight, j.project_layer.bias, labels, act_lens,
                               label_lens, self.blank, self.reduction, host_lens)


class JointNet(nn.Module):
    """Same parameters, names and forward semantics as rnnt/model.py:9-37, so reference
    checkpoints (utils.py:74-85, key 'joint') load unchanged. ``forward`` materialises logits
    (needed by greedy decoding, model.py:90); training should use ``JointRNNTLoss``."""

    def __init__(self, input_size: int, inner_dim: int, vocab_size: int):
        super().__init__()
        self.forward_layer = nn.Linear(input_size, inner_dim, bias=True)
        self.tanh = nn.Tanh()
        self.project_layer = nn.Linear(inner_dim, vocab_size, bias=True)

    def forward(self, enc_state, dec_state):
        if enc_state.dim() == 3 and dec_state.dim() == 3:
            E = enc_state.shape[-1]
            w1 = self.forward_layer.weight
            pe = F.linear(enc_state, w1[:, :E], self.forward_layer.bias)
            pd = F.linear(dec_state, w1[:, E:])
            hidden = self.tanh(pe.unsqueeze(2) + pd.unsqueeze(1))
        else:
            assert enc_state.dim() == dec_state.dim()
            hidden = self.tanh(self.forward_layer(torch.cat((enc_state, dec_state), dim=-1)))
        return self.project_layer(hidden)


class JointRNNTLoss(nn.Module):
    """``JointRNNTLoss(joint)(enc_state, dec_state, labels, act_lens, label_lens) -> loss``.

    ``joint`` is any module with ``forward_layer`` / ``project_layer`` Linear children (this
    package's ``JointNet`` or the reference's). Weight tying with the decoder embedding
    (model.py:55-57) works because gradients accumulate into whatever tensor
    ``project_layer.weight`` is."""

    def __init__(self, joint: nn.Module, blank: int = 0, reduction: str = "mean"):
        super().__init__()
        self.joint = joint
        self.blank = blank
        self.reduction = reduction

    def forward(self, enc_state, dec_state, labels, act_lens, label_lens, host_lens=None):
        j = self.joint
        return joint_rnnt_loss(enc_state, dec_state, j.forward_layer.weight, j.forward_layer.bias,
                               j.project_layer.weight, j.project_layer.bias, labels, act_lens,
                               label_lens, self.blank, self.reduction, host_lens)


class _ReplayFn(torch.autograd.Function):
    """Autograd glue of GraphedJointRNNTLoss: the graph has already produced the loss and the gradients of
    sum_b loss_b; backward hands them out scaled by the incoming (scalar) gradient."""

    @staticmethod
    def forward(ctx, loss_sum, grads, *leaves):
        ctx.grads = grads
        return loss_sum.clone()

    @staticmethod
    def backward(ctx, g):
        scale = g.reshape(()).to(torch.float32)
        return (None, None) + tuple(gr * scale for gr in ctx.grads)


class GraphedJointRNNTLoss(nn.Module):
    """``JointRNNTLoss`` with forward AND backward of the fused path captured once per input shape in a CUDA graph
    and replayed (the launch overhead around the path). The fused step is ~50 launches; at
    the small BASELINE shapes (configs[0], configs[1]: 0.6-0.8 ms of kernels) the eager path is launch bound, at the
    AISHELL sweep shape it is not. Lengths are read on the device (the tile loops are sized for the padded
    (B,T,U+1) bound), so one graph serves every batch of the same padded shape. The joint's parameters are read at
    replay time. ``reduction`` is 'mean' or 'sum' (the gradients of the parameters are summed over the batch inside
    the graph, so a per-utterance ``grad_output`` cannot be applied afterwards; the scalar one multiplies the finished
    gradients, whereas the eager path feeds it into the dS epilogue ahead of the bf16 rounding: identical bits for
    powers of two, bf16-rounding-level differences otherwise). The returned loss and gradients are fresh tensors."""

    def __init__(self, joint: nn.Module, blank: int = 0, reduction: str = "mean"):
        super().__init__()
        if reduction not in ("mean", "sum"):
            raise ValueError("GraphedJointRNNTLoss supports reduction 'mean' or 'sum'")
        self.joint, self.blank, self.reduction = joint, blank, reduction
        self._graphs = {}

    def _params(self):
        j = self.joint
        return [j.forward_layer.weight, j.forward_layer.bias, j.project_layer.weight, j.project_layer.bias]

    def _run(self, ent):
        enc_s, dec_s, y_s, tl_s, ul_s = ent["inputs"]
        # fresh leaves created on the current (capturing) stream: aliases of the static inputs and of the
        # parameters' storage (a replay reads the parameters' current values). Differentiating with respect to the
        # parameters themselves would tie the capture to the stream their grad accumulators were created on.
        leaves = [t.detach().requires_grad_(True) for t in [enc_s, dec_s] + self._params()]
        loss_b = joint_rnnt_loss(*leaves, y_s, tl_s, ul_s, self.blank, "none")
        total = loss_b.sum()
        grads = torch.autograd.grad(total, leaves)
        return total.detach(), grads

    def forward(self, enc_state, dec_state, labels, act_lens, label_lens):
        dev = enc_state.device
        key = (tuple(enc_state.shape), tuple(dec_state.shape), dev)
        ent = self._graphs.get(key)
        if ent is None:
            ent = {"inputs": (enc_state.detach().clone(), dec_state.detach().clone(),
                              labels.clone(), act_lens.clone(), label_lens.clone())}
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):          # warm-up off the capture (lazy library initialisation)
                self._run(ent)
            torch.cuda.current_stream(dev).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                ent["total"], ent["grads"] = self._run(ent)
            ent["graph"] = graph
            self._graphs[key] = ent
        for dst, src in zip(ent["inputs"], (enc_state, dec_state, labels, act_lens, label_lens)):
            dst.copy_(src.detach())
        ent["graph"].replay()
        coeff = 1.0 / enc_state.shape[0] if self.reduction == "mean" else 1.0
        out = _ReplayFn.apply(ent["total"], ent["grads"], enc_state, dec_state, *self._params())
        return (out * coeff).reshape(1)


@torch.no_grad()
def greedy_decode(enc_state, inputs_length, decoder: nn.Module, joint: nn.Module, blank: int = 0):
    """Batched, on-device greedy search with the semantics of ``Transducer.recognize``
    (rnnt/model.py:74-111; the same loop is duplicated in rnnt/search.py:5-40): per utterance, walk
    the frames t = 0..len-1; at each frame take argmax of the joint on (enc_state[t], dec_state);
    a non-blank prediction is emitted and fed to the decoder (at most ONE symbol per frame).

    The reference runs this per utterance with a ``.item()`` sync per frame. Here all B utterances
    advance together: the first joint layer's encoder half is one GEMM for the whole batch, each
    frame is one ``rnnt_b200_joint_step_argmax`` launch (fp32, no logits written) plus one batched
    decoder step whose result is kept only where a symbol was emitted; nothing is read back until
    the end. ``decoder(tokens (B,1), hidden=hidden) -> (out (B,1,D), hidden)`` is the reference's
    ``BaseDecoder`` interface (rnnt/decoder.py:24-44).

    Returns ``(tokens (B,T) int64, zero padded; counts (B) int64)`` on ``enc_state.device``;
    ``[tokens[i, :counts[i]].tolist() for i in range(B)]`` is ``recognize``'s return value."""
    if enc_state.dim() != 3 or not enc_state.is_cuda or enc_state.dtype != torch.float32:
        raise ValueError("enc_state must be a float32 CUDA tensor (B,T,E)")
    B, T, E = enc_state.shape
    w1, b1 = joint.forward_layer.weight, joint.forward_layer.bias
    w2 = joint.project_layer.weight.detach().float().contiguous()
    b2 = joint.project_layer.bias.detach().float().contiguous()
    H, V = w1.shape[0], w2.shape[0]
    if H % 64 or not 64 <= H <= 512:
        raise ValueError(f"joint inner size {H} unsupported: need a multiple of 64 in [64,512]")
    if not 0 <= blank < V:
        raise ValueError(f"blank {blank} not in [0,{V})")
    dev = enc_state.device
    L = _lib.lib()
    lengths = inputs_length.to(dev).long()
    pe = F.linear(enc_state, w1[:, :E], b1).contiguous()          # (B,T,H), once
    w1d = w1[:, E:]
    tokens = torch.zeros(B, max(T, 1), dtype=torch.long, device=dev)
    counts = torch.zeros(B, dtype=torch.long, device=dev)
    pred = torch.empty(B, dtype=torch.long, device=dev)
    work = torch.zeros(B + (B + 63) // 64, dtype=torch.long, device=dev)  # one B+1 area per 64-row launch
    dec_out, hidden = decoder(torch.full((B, 1), blank, dtype=torch.long, device=dev))
    with torch.cuda.device(dev):
        st = _stream()
        for t in range(T):
            pd = F.linear(dec_out[:, 0], w1d).contiguous()          # (B,H)
            for c, lo in enumerate(range(0, B, 64)):
                n = min(64, B - lo)
                _call("joint_step_argmax", L.rnnt_b200_joint_step_argmax,
                      pe.data_ptr() + (lo * T + t) * H * 4, T * H, _ptr(pd[lo:]), _ptr(w2), _ptr(b2),
                      n, H, V, work.data_ptr() + (lo + c) * 8, pred.data_ptr() + lo * 8, st)
            emit = (pred != blank) & (lengths > t)
            new_out, new_hidden = decoder(pred.unsqueeze(1), hidden=hidden)
            dec_out = torch.where(emit[:, None, None], new_out, dec_out)
            hidden = tuple(torch.where(emit[None, :, None], n_, o_) for n_, o_ in zip(new_hidden, hidden))
            # a row that emits nothing rewrites a zero at its next free slot
            tokens.scatter_(1, counts.unsqueeze(1), torch.where(emit, pred, torch.zeros_like(pred)).unsqueeze(1))
            counts += emit
    return tokens, counts


class GreedyDecoder:
    """``GreedyDecoder(decoder, joint)(enc_state, inputs_length) -> (tokens, counts)``:
    ``greedy_decode`` with the whole frame loop captured once per input shape in a CUDA graph
    (the loop is launch bound: ~15 small launches per frame) and
    replayed on static buffers. Weights are read at replay time, so parameter updates between
    calls are seen; a change of shape captures a new graph. ``tokens``/``counts`` are views of
    static buffers: clone them if they must survive the next call."""

    def __init__(self, decoder: nn.Module, joint: nn.Module, blank: int = 0):
        self.decoder, self.joint, self.blank = decoder, joint, blank
        self._graphs = {}

    @torch.no_grad()
    def __call__(self, enc_state, inputs_length):
        key = (tuple(enc_state.shape), enc_state.device)
        ent = self._graphs.get(key)
        if ent is None:
            enc_s = enc_state.detach().clone()
            len_s = inputs_length.to(enc_state.device).long().clone()
            side = torch.cuda.Stream(device=enc_state.device)
            side.wait_stream(torch.cuda.current_stream(enc_state.device))
            with torch.cuda.stream(side):  # warm-up off the capture: lazy init of cuDNN / cuBLAS
                greedy_decode(enc_s, len_s, self.decoder, self.joint, self.blank)
            torch.cuda.current_stream(enc_state.device).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = greedy_decode(enc_s, len_s, self.decoder, self.joint, self.blank)
            ent = self._graphs[key] = (graph, enc_s, len_s, out)
        graph, enc_s, len_s, out = ent
        enc_s.copy_(enc_state)
        len_s.copy_(inputs_length.to(enc_s.device))
        graph.replay()
        return out


def selftest_gemm(a: torch.Tensor, b: torch.Tensor, ts: int = 0) -> torch.Tensor:
    """C = A . B^T through the bare tcgen05/TMA plumbing (bf16 in, fp32 out); ts = 1 / 2 takes the A
    operand from tensor memory (staged by tcgen05.cp / written by tcgen05.st)."""
    M, K = a.shape
    N = b.shape[0]
    c = torch.empty(M, N, dtype=torch.float32, device=a.device)
    with torch.cuda.device(a.device):
        _lib.check(_lib.lib().rnnt_b200_selftest_gemm_ts(_ptr(a), _ptr(b), _ptr(c), M, N, K, int(ts), _stream()),
                   "rnnt_b200_selftest_gemm_ts")
    return c
