"""Parity measurements of the CUDA path against the oracle (functions return metric dicts;
tests/test_*_gpu.py assert on them, tools/gpu_check.py prints them). Everything here calls
the product through its public API -> C ABI; the oracle (oracle/) is only the checker."""
import numpy as np
import torch

import rnnt_transducer_b200 as R
from oracle import c_oracle as C
from oracle import rnnt_oracle as O
from tests.helpers import load_golden, max_abs, rel_l2, synth_labels, synth_lengths

DEV = "cuda:0"


def selftest_gemm(M, N, K, seed=0, ts=0):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(M, K, generator=g).to(torch.bfloat16)
    b = torch.randn(N, K, generator=g).to(torch.bfloat16)
    c = R.selftest_gemm(a.to(DEV), b.to(DEV), ts).cpu()
    ref = a.double() @ b.double().t()
    return {"max_abs": max_abs(c, ref), "rel_l2": rel_l2(c, ref), "ref_absmax": float(ref.abs().max())}


def _dropin_run(x, y, tl, ul, reduction="none"):
    xg = x.to(DEV).requires_grad_(True)
    crit = R.RNNTLoss(reduction=reduction)
    loss = crit(xg, y.to(DEV), tl.to(DEV), ul.to(DEV))
    loss.sum().backward()
    torch.cuda.synchronize()
    return loss.detach().cpu(), xg.grad.detach().cpu()


def dropin_vs_oracle(x, y, tl, ul):
    """Boundary 1 on materialised fp32 logits vs the float64 C oracle."""
    loss, grad = _dropin_run(x, y, tl, ul)
    lo, go = C.rnnt_loss(x.numpy(), y.numpy(), tl.numpy(), ul.numpy())
    lo = torch.from_numpy(lo)
    out = {
        "loss_rel": float(((loss.double() - lo).abs() / lo.abs()).max()),
        "grad_max_abs": max_abs(grad, go),
        "grad_rel_l2": rel_l2(grad, go),
        "grad_outside_window_max": 0.0,
        "loss_mean": float(lo.mean()),
    }
    B, T, U1, V = x.shape
    for b in range(B):
        tb, ub = int(tl[b]), int(ul[b])
        outside = torch.ones(T, U1, dtype=torch.bool)
        outside[:tb, :ub + 1] = False
        if outside.any():
            out["grad_outside_window_max"] = max(out["grad_outside_window_max"],
                                                 float(grad[b][outside].abs().max()))
    return out


def dropin_golden(name):
    g = load_golden(name)
    t = lambda k: torch.from_numpy(g[k])
    m = dropin_vs_oracle(t("ref_joint_logits"), t("labels"), t("act_lens"), t("label_lens"))
    loss, grad = _dropin_run(t("ref_joint_logits"), t("labels"), t("act_lens"), t("label_lens"))
    m["loss_rel_vs_torchaudio"] = float(((loss - t("torchaudio_loss")).abs() / t("torchaudio_loss").abs()).max())
    m["grad_max_abs_vs_torchaudio"] = max_abs(grad, g["torchaudio_dlogits"])
    return m


def dropin_synth(B, T, U, V, mode="ragged", seed=2019, trained_like=False):
    gen = torch.Generator().manual_seed(seed)
    tl, ul = synth_lengths(B, T, U, mode, gen)
    y = synth_labels(B, U, V, ul, gen)
    x = torch.randn(B, T, U + 1, V, generator=gen)
    if trained_like:  # +8 on blank and on the next label
        x[..., 0] += 8.0
        for b in range(B):
            for u in range(int(ul[b])):
                x[b, :, u, int(y[b, u])] += 8.0
    return dropin_vs_oracle(x, y, tl, ul)


def _joint_inputs(B, T, U, E, D, H, V, mode, seed):
    gen = torch.Generator().manual_seed(seed)
    tl, ul = synth_lengths(B, T, U, mode, gen)
    y = synth_labels(B, U, V, ul, gen)
    enc = torch.randn(B, T, E, generator=gen)
    dec = torch.randn(B, U + 1, D, generator=gen)
    k1, k2 = 1.0 / (E + D) ** 0.5, 1.0 / H ** 0.5   # nn.Linear default init range
    w1 = (torch.rand(H, E + D, generator=gen) * 2 - 1) * k1
    b1 = (torch.rand(H, generator=gen) * 2 - 1) * k1
    w2 = (torch.rand(V, H, generator=gen) * 2 - 1) * k2
    b2 = (torch.rand(V, generator=gen) * 2 - 1) * k2
    return enc, dec, w1, b1, w2, b2, y, tl, ul


def fused_vs_oracle(enc, dec, w1, b1, w2, b2, y, tl, ul, grads=True, host_lens=False):
    """Boundary 2 vs (i) the float64 oracle fed bf16-rounded h and W2 (kernel error only) and
    (ii) the full-precision float64 oracle (kernel + quantisation error)."""
    names = ["enc", "dec", "w1", "b1", "w2", "b2"]
    leaves = [t.to(DEV).requires_grad_(grads) for t in (enc, dec, w1, b1, w2, b2)]
    loss = R.joint_rnnt_loss(*leaves, y.to(DEV), tl.to(DEV), ul.to(DEV), reduction="none",
                             host_lens=(tl, ul) if host_lens else None)
    if grads:
        loss.sum().backward()
    torch.cuda.synchronize()
    out = {}
    for tag, q in (("q", True), ("full", False)):
        ol = [t.double().requires_grad_(grads) for t in (enc, dec, w1, b1, w2, b2)]
        lo = O.joint_rnnt_loss(*ol, y, tl, ul, reduction="none", quantize_bf16=q)
        out[f"loss_rel_{tag}"] = float(((loss.detach().cpu().double() - lo.detach()).abs() / lo.detach().abs()).max())
        if grads:
            lo.sum().backward()
            for n, a, b in zip(names, leaves, ol):
                out[f"d{n}_rel_l2_{tag}"] = rel_l2(a.grad, b.grad)
                out[f"d{n}_max_abs_{tag}"] = max_abs(a.grad, b.grad)
    out["loss_mean"] = float(loss.detach().mean())
    return out


def fused_golden(name, grads=True):
    g = load_golden(name)
    t = lambda k: torch.from_numpy(g[k])
    return fused_vs_oracle(t("enc"), t("dec"), t("w1"), t("b1"), t("w2"), t("b2"), t("labels"),
                           t("act_lens"), t("label_lens"), grads)


def fused_synth(B, T, U, E, D, H, V, mode="ragged", seed=2019, grads=True, host_lens=False):
    return fused_vs_oracle(*_joint_inputs(B, T, U, E, D, H, V, mode, seed), grads=grads, host_lens=host_lens)


def fused_matches_dropin(B, T, U, E, D, H, V, mode="ragged", seed=7):
    """Size-independent consistency: the fused path and the drop-in path fed with the bf16-
    consistent logits agree on loss; runs at sizes where the CPU oracle would be slow."""
    enc, dec, w1, b1, w2, b2, y, tl, ul = [t.to(DEV) for t in _joint_inputs(B, T, U, E, D, H, V, mode, seed)]
    loss_f = R.joint_rnnt_loss(enc, dec, w1, b1, w2, b2, y, tl, ul, reduction="none")
    pe = torch.nn.functional.linear(enc, w1[:, :E], b1)
    pd = torch.nn.functional.linear(dec, w1[:, E:])
    h = torch.tanh(pe.unsqueeze(2) + pd.unsqueeze(1)).to(torch.bfloat16).float()
    logits = (h @ w2.to(torch.bfloat16).float().t() + b2).contiguous()
    loss_d = R.rnnt_loss(logits, y, tl, ul, reduction="none")
    torch.cuda.synchronize()
    return {"loss_rel": float(((loss_f - loss_d).abs() / loss_d.abs()).max()), "loss_mean": float(loss_d.mean())}


def projected_matches_two_step(B, T, U, He, Hd, E, D, H, V, mode="ragged", seed=11):
    """ProjectedJointRNNTLoss on the LSTM outputs against output_proj followed by
    JointRNNTLoss -- loss and the gradients of every tensor both paths share."""
    g = torch.Generator().manual_seed(seed)
    joint = R.JointNet(E + D, H, V).to(DEV)
    enc_proj = torch.nn.Linear(He, E).to(DEV)
    dec_proj = torch.nn.Linear(Hd, D).to(DEV)
    xe = torch.randn(B, T, He, generator=g).to(DEV).requires_grad_(True)
    xd = torch.randn(B, U + 1, Hd, generator=g).to(DEV).requires_grad_(True)
    from tests.helpers import synth_labels, synth_lengths
    tl, ul = synth_lengths(B, T, U, mode, g)
    y = synth_labels(B, U, V, ul, g).to(DEV)
    tl, ul = tl.to(DEV), ul.to(DEV)
    params = [xe, xd] + list(enc_proj.parameters()) + list(dec_proj.parameters()) + list(joint.parameters())

    def run(fn):
        for q in params:
            q.grad = None
        loss = fn()
        loss.sum().backward()
        return loss.detach().clone(), [q.grad.detach().clone() for q in params]

    two = R.JointRNNTLoss(joint, reduction="none")
    one = R.ProjectedJointRNNTLoss(joint, enc_proj, dec_proj, reduction="none")
    l2, g2 = run(lambda: two(enc_proj(xe), dec_proj(xd), y, tl, ul))
    l1, g1 = run(lambda: one(xe, xd, y, tl, ul))
    torch.cuda.synchronize()
    names = ["enc_rnn", "dec_rnn", "enc_proj.w", "enc_proj.b", "dec_proj.w", "dec_proj.b", "w1", "b1", "w2", "b2"]
    out = {"loss_rel": float(((l1 - l2).abs() / l2.abs()).max())}
    for n, a, b in zip(names, g1, g2):
        out["d" + n] = float((a - b).norm() / b.norm().clamp_min(1e-30))
    return out


def greedy_matches_oracle(B, T, E, D, H, V, hidden=48, layers=1, seed=5, scale=3.0):
    """batched on-device greedy search against the per-utterance restatement of
    Transducer.recognize (oracle/greedy_oracle.py), same modules, same device."""
    from oracle.greedy_oracle import greedy_search
    from tests.helpers import make_decoder
    g = torch.Generator().manual_seed(seed)
    torch.manual_seed(seed)
    joint = R.JointNet(E + D, H, V).to(DEV)
    with torch.no_grad():  # spread the logits so that blanks and symbols both occur
        joint.project_layer.weight.mul_(scale * 4)
        joint.project_layer.bias.mul_(scale)
    dec = make_decoder(hidden, V, D, layers, seed).to(DEV).eval()
    enc = (torch.randn(B, T, E, generator=g) * scale).to(DEV)
    with torch.no_grad():  # calibrate the blank bias so that about half of the frames emit nothing
        d0, _ = dec(torch.zeros(1, 1, dtype=torch.long, device=DEV))
        lg = joint.project_layer(torch.tanh(joint.forward_layer(
            torch.cat((enc.reshape(-1, E), d0.view(1, -1).expand(B * T, -1)), dim=-1))))
        joint.project_layer.bias[0] += (lg[:, 1:].max(dim=1).values - lg[:, 0]).median()
    lens = torch.randint(max(1, T // 2), T + 1, (B,), generator=g)
    lens[0] = T
    tokens, counts = R.greedy_decode(enc, lens, dec, joint)
    got = [tokens[i, :int(counts[i])].tolist() for i in range(B)]
    want = greedy_search(enc, lens, dec, joint)
    gd = R.GreedyDecoder(dec, joint)
    for _ in range(2):  # capture, then a pure replay
        tg, cg = gd(enc, lens)
    graph_equal = [tg[i, :int(cg[i])].tolist() for i in range(B)] == want
    return {"equal": got == want, "graph_equal": graph_equal, "n_symbols": sum(len(w) for w in want),
            "n_frames": int(lens.sum()), "got": got[:2], "want": want[:2]}


def full_size_properties(B=32, T=500, U=100, E=320, D=320, H=512, V=4232, seed=2019):
    """BASELINE.json's headline shape, where no CPU oracle finishes: size-independent properties.
    (1) every row of dS sums to zero (softmax*c0 minus the two occupation terms), hence sum(db2) = 0;
    (2) an utterance's loss does not depend on what else is in the batch;
    (3) the backward pass is linear in grad_output (a power-of-two factor is exact in bf16/fp32
        up to the order of the fp32 reductions);
    (4) full T and U against the drop-in boundary (materialised logits) on a 2-utterance slice."""
    names = ["enc", "dec", "w1", "b1", "w2", "b2"]
    ins = _joint_inputs(B, T, U, E, D, H, V, "ragged", seed)
    y, tl, ul = (t.to(DEV) for t in ins[6:])

    def run(scale, sl=slice(None)):
        leaves = [t.to(DEV).clone().requires_grad_(True) for t in ins[:6]]
        enc, dec = leaves[0][sl], leaves[1][sl]
        loss = R.joint_rnnt_loss(enc, dec, *leaves[2:], y[sl], tl[sl], ul[sl], reduction="none")
        (loss.sum() * scale).backward()
        return loss.detach(), [t.grad for t in leaves]

    l1, g1 = run(1.0)
    l1b, g1b = run(1.0)
    l2, g2 = run(2.0)
    ls, _ = run(1.0, slice(0, 3))
    out = {"loss_mean": float(l1.mean()), "finite": bool(all(torch.isfinite(g).all() for g in g1)),
           "db2_sum_over_abs": float(g1[5].sum().abs() / g1[5].abs().sum()),
           "loss_subbatch_rel": float(((ls - l1[:3]).abs() / l1[:3].abs()).max()),
           "deterministic": bool(torch.equal(l1, l1b) and all(torch.equal(a, b) for a, b in zip(g1, g1b)))}
    for n, a, b in zip(names, g1, g2):
        out[f"lin_{n}"] = float((b - 2 * a).norm() / (2 * a).norm())
    # (4) drop-in boundary on two utterances at full T, U, fed the bf16-consistent logits
    enc, dec, w1, b1, w2, b2 = (t.to(DEV) for t in ins[:6])
    sl = slice(0, 2)
    pe = torch.nn.functional.linear(enc[sl], w1[:, :E], b1)
    pd = torch.nn.functional.linear(dec[sl], w1[:, E:])
    h = torch.tanh(pe.unsqueeze(2) + pd.unsqueeze(1)).to(torch.bfloat16).float()
    logits = (h @ w2.to(torch.bfloat16).float().t() + b2).contiguous()
    ld = R.rnnt_loss(logits, y[sl], tl[sl], ul[sl], reduction="none")
    torch.cuda.synchronize()
    out["loss_vs_dropin_rel"] = float(((l1[sl] - ld).abs() / ld.abs()).max())
    return out


def _dw2_run(L, ds, h, rows, V, H, Vds, sm_pairs=0, weight=None):
    """One chunk through rnnt_b200_dw2_gemm (or _weighted) + the ordered sum of its K-split partials."""
    ns = L.rnnt_b200_dw2_splits(rows, V, H, sm_pairs)
    parts = torch.empty(ns, Vds, H, device=DEV)
    bparts = torch.zeros(ns, Vds, device=DEV)
    dw2 = torch.empty(Vds, H, device=DEV)
    db2 = torch.empty(Vds, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    if weight is None:
        rc = L.rnnt_b200_dw2_gemm(ds.data_ptr(), h.data_ptr(), rows, V, H, sm_pairs, parts.data_ptr(), bparts.data_ptr(), st)
    else:
        rc = L.rnnt_b200_dw2_gemm_weighted(ds.data_ptr(), h.data_ptr(), weight.data_ptr(), rows, V, H, sm_pairs,
                                           parts.data_ptr(), bparts.data_ptr(), st)
    assert rc == 0, rc
    assert L.rnnt_b200_sum_partials(parts.data_ptr(), ns, Vds * H, dw2.data_ptr(), st) == 0
    assert L.rnnt_b200_sum_partials(bparts.data_ptr(), ns, Vds, db2.data_ptr(), st) == 0
    torch.cuda.synchronize()
    return dw2, db2, ns


def dw2_gemm_vs_torch(rows, V, H, seed=0, sm_pairs=0, weighted=False):
    """rnnt_b200_dw2_gemm (tcgen05, MN-major operands, K-split partials added in slot order) against float64 torch
    products of the same bf16 operands; run twice: the two results must be bit-identical (deterministic).
    ``weighted``: rnnt_b200_dw2_gemm_weighted, db2 = sum_r w[r] dS[r, :] (the stored-E backward)."""
    from rnnt_transducer_b200 import _lib
    L = _lib.lib()
    Vds, hp = L.rnnt_b200_ds_pitch(V), L.rnnt_b200_h_pitch(H)
    g = torch.Generator().manual_seed(seed)
    ds = (torch.randn(rows, Vds, generator=g) * 0.05).to(torch.bfloat16).to(DEV)
    h = torch.tanh(torch.randn(rows, hp, generator=g)).to(torch.bfloat16).to(DEV)
    w = None
    if weighted:
        w = torch.randn(rows, generator=g)
        w[::7] = 0.0
        w = w.to(DEV)
    dw2, db2, ns = _dw2_run(L, ds, h, rows, V, H, Vds, sm_pairs, w)
    dw2b, db2b, _ = _dw2_run(L, ds, h, rows, V, H, Vds, sm_pairs, w)
    ref = ds.double().t() @ h[:, :H].double()
    refb = ds.double().sum(0) if w is None else (ds.double() * w.double()[:, None]).sum(0)
    return {"dw2_rel_l2": rel_l2(dw2.cpu(), ref.cpu()), "db2_rel_l2": rel_l2(db2.cpu(), refb.cpu()), "splits": ns,
            "bit_identical": bool(torch.equal(dw2, dw2b) and torch.equal(db2, db2b))}


def dh_gemm_vs_torch(tiles, V, H, lo=0, hi=None, seed=0, sm_pairs=0, scaled=False):
    """rnnt_b200_dh_gemm on its own: dpre = (dS . W2q) * (1 - h^2) for tiles [lo, hi) of a stash of ``tiles`` tiles,
    against the float64 product of the same bf16 operands; rows outside [lo, hi) must stay untouched.
    ``scaled``: rnnt_b200_dh_gemm_scaled, every row times row_scale[row] (the stored-E backward)."""
    from rnnt_transducer_b200 import _lib
    L = _lib.lib()
    hi = tiles if hi is None else hi
    Vds, Vpad = L.rnnt_b200_ds_pitch(V), L.rnnt_b200_vocab_pad(V)
    g = torch.Generator().manual_seed(seed)
    rows = (hi - lo) * 128
    ds = (torch.randn(rows, Vds, generator=g) * 0.05).to(torch.bfloat16)
    ds[:, V:] = 0
    w2q = torch.zeros(Vpad, H, dtype=torch.bfloat16)
    w2q[:V] = (torch.randn(V, H, generator=g) * 0.05).to(torch.bfloat16)
    h = torch.tanh(torch.randn(tiles * 128, H, generator=g)).to(torch.bfloat16)
    dsd, w2d, hd = ds.to(DEV), w2q.t().contiguous().to(DEV), h.to(DEV)   # the kernel takes W2 transposed (H, Vpad)
    dpre = torch.full((tiles * 128, H), 7.0, dtype=torch.bfloat16, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    if scaled:
        rs = torch.randn(tiles * 128, generator=g) * 3.0
        rs[::5] = 0.0
        rsd = rs.to(DEV)
        rc = L.rnnt_b200_dh_gemm_scaled(dsd.data_ptr(), w2d.data_ptr(), hd.data_ptr(), rsd.data_ptr(), V, H, tiles,
                                        lo, hi, sm_pairs, dpre.data_ptr(), st)
    else:
        rc = L.rnnt_b200_dh_gemm(dsd.data_ptr(), w2d.data_ptr(), hd.data_ptr(), V, H, tiles, lo, hi, sm_pairs, dpre.data_ptr(), st)
    torch.cuda.synchronize()
    assert rc == 0, rc
    hs = h[lo * 128: hi * 128].double()
    ref = (ds.double() @ w2q[:Vds].double()) * (1.0 - hs * hs)
    if scaled:
        ref = ref * rs[lo * 128: hi * 128].double()[:, None]
    got = dpre.cpu()
    outside = torch.cat((got[:lo * 128], got[hi * 128:]))
    return {"rel_l2": rel_l2(got[lo * 128: hi * 128], ref), "max_abs": max_abs(got[lo * 128: hi * 128], ref),
            "ref_absmax": float(ref.abs().max()), "outside_untouched": bool((outside == 7.0).all())}


def dpre_reduce_vs_torch(B, T, U, H, mode="ragged", seed=0):
    """rnnt_b200_dpre_reduce against float64 sums of the same bf16 dpre rows (tile-packed), twice (bit-identical)."""
    from rnnt_transducer_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(seed)
    tl, ul = synth_lengths(B, T, U, mode, g)
    U1 = U + 1
    ntile = [(int(t) * (int(u) + 1) + 127) // 128 for t, u in zip(tl, ul)]
    tiles = sum(ntile)
    dpre = (torch.randn(tiles * 128, H, generator=g) * 0.01).to(torch.bfloat16)
    want_e = torch.zeros(B, T, H, dtype=torch.float64)
    want_d = torch.zeros(B, U1, H, dtype=torch.float64)
    first = 0
    for b in range(B):
        tb, w = int(tl[b]), int(ul[b]) + 1
        blk = dpre[first * 128: first * 128 + tb * w].double().view(tb, w, H)
        want_e[b, :tb] = blk.sum(1)
        want_d[b, :w] = blk.sum(0)
        first += ntile[b]
    dd, tld, uld = dpre.to(DEV), tl.to(DEV), ul.to(DEV)
    outs = []
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        scratch = torch.empty(L.rnnt_b200_dpre_reduce_scratch_elems(B, T, U1, H), device=DEV)
        dpe = torch.zeros(B, T, H, device=DEV)
        dpd = torch.full((B, U1, H), 7.0, device=DEV)
        rc = L.rnnt_b200_dpre_reduce(dd.data_ptr(), tld.data_ptr(), uld.data_ptr(), B, T, U1, H, scratch.data_ptr(),
                                     dpe.data_ptr(), dpd.data_ptr(), st)
        assert rc == 0, rc
        torch.cuda.synchronize()
        outs.append((dpe, dpd))
    return {"dpe_rel_l2": rel_l2(outs[0][0].cpu(), want_e), "dpd_rel_l2": rel_l2(outs[0][1].cpu(), want_d),
            "bit_identical": bool(torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1]))}


def fused_vs_large_oracle(B, T, U, E, D, H, V, mode="ragged", seed=2019, full_grads=True, lens=None):
    """Boundary 2 at a BASELINE shape against the per-utterance, t-chunked float64 oracle (oracle/large_oracle.py):
    per-utterance loss and all six gradients of sum_b loss_b, (i) with the same bf16-rounded h / W2 and (ii) in
    full precision (``full_grads=False``: loss only for (ii))."""
    from oracle import large_oracle as LO
    ins = _joint_inputs(B, T, U, E, D, H, V, mode, seed)
    enc, dec, w1, b1, w2, b2, y, tl, ul = ins
    if lens is not None:
        tl, ul = (torch.tensor(v, dtype=torch.int32) for v in lens)
        y = synth_labels(B, U, V, ul, torch.Generator().manual_seed(seed + 1))
    names = ["enc", "dec", "w1", "b1", "w2", "b2"]
    leaves = [t.to(DEV).requires_grad_(True) for t in (enc, dec, w1, b1, w2, b2)]
    loss = R.joint_rnnt_loss(*leaves, y.to(DEV), tl.to(DEV), ul.to(DEV), reduction="none", host_lens=(tl, ul))
    loss.sum().backward()
    torch.cuda.synchronize()
    out = {"loss_mean": float(loss.detach().mean()), "cells": int((tl.long() * (ul.long() + 1)).sum())}
    lg = loss.detach().cpu().double()
    for tag, q, want_g in (("q", True, True), ("full", False, full_grads)):
        lo, g = LO.joint_rnnt_loss_and_grads(enc, dec, w1, b1, w2, b2, y, tl, ul, quantize_bf16=q, grads=want_g)
        out[f"loss_rel_{tag}"] = float(((lg - lo).abs() / lo.abs()).max())
        if want_g:
            for n, a in zip(names, leaves):
                out[f"d{n}_rel_l2_{tag}"] = rel_l2(a.grad, g[n])
                out[f"d{n}_max_abs_{tag}"] = max_abs(a.grad, g[n])
    return out


def _fused_grads(inputs, y, tl, ul, mode, weight=None):
    """loss and the six gradients of boundary 2 with the backward forced to ``mode`` (RNNT_B200_BWD)."""
    import os
    prev = os.environ.get("RNNT_B200_BWD")
    os.environ["RNNT_B200_BWD"] = mode
    try:
        leaves = [t.to(DEV).requires_grad_(True) for t in inputs]
        loss = R.joint_rnnt_loss(*leaves, y.to(DEV), tl.to(DEV), ul.to(DEV), reduction="none")
        (loss if weight is None else loss * weight.to(DEV)).sum().backward()
        torch.cuda.synchronize()
    finally:
        if prev is None:
            os.environ.pop("RNNT_B200_BWD", None)
        else:
            os.environ["RNNT_B200_BWD"] = prev
    return loss.detach(), [t.grad for t in leaves]


def stored_matches_recompute(B, T, U, E, D, H, V, mode="ragged", seed=8):
    """The two backward paths on the same inputs: E kept by the forward (no second vocabulary GEMM) against the
    recomputing one. The loss differs only in the LSE's reference (fixed against running maximum), the gradients in
    where the bf16 roundings fall (E and g h instead of dS)."""
    enc, dec, w1, b1, w2, b2, y, tl, ul = _joint_inputs(B, T, U, E, D, H, V, mode, seed)
    la, ga = _fused_grads((enc, dec, w1, b1, w2, b2), y, tl, ul, "stored")
    lb, gb = _fused_grads((enc, dec, w1, b1, w2, b2), y, tl, ul, "recompute")
    out = {"loss_rel": float(((la - lb).abs() / lb.abs()).max())}
    for n, a, b in zip(["enc", "dec", "w1", "b1", "w2", "b2"], ga, gb):
        out[f"d{n}_rel_l2"] = rel_l2(a, b)
    return out


def negative_upstream_gradient(B, T, U, E, D, H, V, bwd, seed=9):
    """Per-utterance weights of both signs on the loss (grad_output < 0 for some utterances): the gradients against
    the float64 oracle's for the same weighted sum."""
    enc, dec, w1, b1, w2, b2, y, tl, ul = _joint_inputs(B, T, U, E, D, H, V, "ragged", seed)
    wgt = torch.tensor([(-1.0) ** i * (1.0 + 0.5 * i) for i in range(B)])
    _, ga = _fused_grads((enc, dec, w1, b1, w2, b2), y, tl, ul, bwd, weight=wgt)
    ol = [t.double().requires_grad_(True) for t in (enc, dec, w1, b1, w2, b2)]
    lo = O.joint_rnnt_loss(*ol, y, tl, ul, reduction="none", quantize_bf16=True)
    (lo * wgt.double()).sum().backward()
    return {f"d{n}_rel_l2_q": rel_l2(a, b.grad) for n, a, b in zip(["enc", "dec", "w1", "b1", "w2", "b2"], ga, ol)}


def logit_spread_beyond_stored_range(seed=12):
    """One vocabulary entry's bias 400 nats above the rest (outside the first columns a row sees): the forward that
    stores exponentials against a fixed reference must come out non-finite (never finite and wrong), raise under
    RNNT_B200_VALIDATE=1, and the recomputing path (running maximum) must agree with the float64 oracle."""
    import os
    B, T, U, E, D, H, V = 2, 9, 4, 64, 64, 512, 600
    enc, dec, w1, b1, w2, b2, y, tl, ul = _joint_inputs(B, T, U, E, D, H, V, "full", seed)
    b2 = b2.clone()
    b2[300] = 400.0
    ins = (enc, dec, w1, b1, w2, b2)
    out = {}
    la, _ = _fused_grads(ins, y, tl, ul, "stored")
    out["stored_finite"] = bool(torch.isfinite(la).all())
    prev = os.environ.get("RNNT_B200_VALIDATE")
    os.environ["RNNT_B200_VALIDATE"] = "1"
    try:
        _fused_grads(ins, y, tl, ul, "stored")
        out["validate_raises"] = False
    except FloatingPointError:
        out["validate_raises"] = True
    finally:
        if prev is None:
            del os.environ["RNNT_B200_VALIDATE"]
        else:
            os.environ["RNNT_B200_VALIDATE"] = prev
    # the kernel has set this device's range flag (pinned host memory): without a forced mode the next call takes
    # the recomputing path by itself
    from rnnt_transducer_b200 import _api
    flag = _api._range_flag(torch.device(DEV))
    out["flag_set"] = int(flag[0]) == 1
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        lc, _ = _fused_grads(ins, y, tl, ul, "auto")
    out["auto_after_flag_finite"] = bool(torch.isfinite(lc).all())
    flag.zero_()
    _api._range_flags.pop("warned", None)
    lb, gb = _fused_grads(ins, y, tl, ul, "recompute")
    out["auto_equals_recompute"] = bool(torch.equal(lc, lb))
    ol = [t.double().requires_grad_(True) for t in ins]
    lo = O.joint_rnnt_loss(*ol, y, tl, ul, reduction="none", quantize_bf16=True)
    lo.sum().backward()
    out["recompute_loss_rel"] = float(((lb.cpu().double() - lo.detach()).abs() / lo.detach().abs()).max())
    out["recompute_dw2_rel_l2"] = rel_l2(gb[4], ol[4].grad)
    return out


def fused_is_deterministic(B, T, U, E, D, H, V, mode="ragged", seed=3, repeats=3):
    """Identical calls must give BIT-identical losses and gradients (no order-dependent atomics in the backward)."""
    ins = _joint_inputs(B, T, U, E, D, H, V, mode, seed)
    y, tl, ul = (t.to(DEV) for t in ins[6:])
    runs = []
    for _ in range(repeats):
        leaves = [t.to(DEV).clone().requires_grad_(True) for t in ins[:6]]
        loss = R.joint_rnnt_loss(*leaves, y, tl, ul, reduction="none")
        loss.sum().backward()
        torch.cuda.synchronize()
        runs.append([loss.detach()] + [t.grad for t in leaves])
    names = ["loss", "enc", "dec", "w1", "b1", "w2", "b2"]
    return {n: all(torch.equal(runs[0][i], r[i]) for r in runs[1:]) for i, n in enumerate(names)}


def graphed_matches_eager(B, T, U, E, D, H, V, seed=6):
    """GraphedJointRNNTLoss (fused fwd+bwd captured in a CUDA graph, replayed) against the eager JointRNNTLoss on the
    same module: capture + two replays with different inputs and lengths (one graph per padded shape; lengths are
    read on the device). reduction 'sum' with grad_output 1: the same kernels see the same numbers, and they are
    deterministic, so loss and gradients must be bit-identical. reduction 'mean' with grad_output 2: the eager path
    feeds grad_output / B into the dS epilogue (through a log2f) ahead of the bf16 rounding, the graph scales the
    finished gradients, so single dS entries round differently: rel-L2 is reported."""
    import time
    torch.manual_seed(seed)
    joint = R.JointNet(E + D, H, V).to(DEV)
    params = list(joint.parameters())
    out = {"equal_sum": True, "mean_rel_l2": 0.0}
    forms = {"sum": (R.JointRNNTLoss(joint, reduction="sum"), R.GraphedJointRNNTLoss(joint, reduction="sum"), 1.0),
             "mean": (R.JointRNNTLoss(joint), R.GraphedJointRNNTLoss(joint), 2.0)}
    for rep, mode in enumerate(["full", "ragged", "edge"]):
        g = torch.Generator().manual_seed(seed + rep)
        tl, ul = synth_lengths(B, T, U, mode, g)
        y = synth_labels(B, U, V, ul, g).to(DEV)
        tl, ul = tl.to(DEV), ul.to(DEV)
        enc = torch.randn(B, T, E, generator=g).to(DEV)
        dec = torch.randn(B, U + 1, D, generator=g).to(DEV)
        for red, (eager, graphed, factor) in forms.items():
            res = []
            for fn in (eager, graphed):
                e, d = enc.clone().requires_grad_(True), dec.clone().requires_grad_(True)
                for p_ in params:
                    p_.grad = None
                loss = fn(e, d, y, tl, ul)
                (loss * factor).backward()
                torch.cuda.synchronize()
                res.append([loss.detach().clone(), e.grad.clone(), d.grad.clone()] + [p_.grad.clone() for p_ in params])
            for a, b in zip(*res):
                if red == "sum":
                    out["equal_sum"] = out["equal_sum"] and bool(torch.equal(a, b))
                else:
                    out["mean_rel_l2"] = max(out["mean_rel_l2"], rel_l2(a, b))
    out["n_graphs"] = len(forms["sum"][1]._graphs)
    # timing: ms per fwd+bwd of both forms at this shape (host-launch bound at small shapes)
    for name, fn in (("eager_ms", forms["mean"][0]), ("graphed_ms", forms["mean"][1])):
        e, d = enc.clone().requires_grad_(True), dec.clone().requires_grad_(True)
        for _ in range(3):
            fn(e, d, y, tl, ul).backward()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(20):
            fn(e, d, y, tl, ul).backward()
        torch.cuda.synchronize()
        out[name] = (time.perf_counter() - t0) / 20 * 1e3
    return out


def shared_embedding_grads(seed=4):
    """rnnt/model.py:55-57 (share_embedding): decoder.embedding.weight IS joint.project_layer.weight. The fused
    loss's dW2 and the embedding's own gradient must both land in the one shared tensor: compared with the float64
    oracle's d loss / d W2 plus the embedding gradient of an identical model without the tying."""
    import copy
    cfg = dict(enc=dict(type="lstm", hidden_size=16, output_size=12, n_layers=1, bidirectional=True),
               dec=dict(type="lstm", hidden_size=128, output_size=12, n_layers=1),
               joint=dict(input_size=24, inner_size=128), vocab_size=30, share_weight=False,
               share_embedding=True, feature_dim=20, dropout=0.0)
    torch.manual_seed(seed)
    tied = R.Transducer(cfg).to(DEV)
    assert tied.decoder.embedding.weight is tied.joint.project_layer.weight
    cfg2 = dict(cfg, share_embedding=False)
    free = R.Transducer(cfg2).to(DEV)
    sd = {k: v for k, v in tied.state_dict().items()}
    free.load_state_dict(sd, strict=True)   # same values, two separate tensors
    g = torch.Generator().manual_seed(seed)
    B, T, U = 3, 14, 5
    x = torch.randn(B, T, 20, generator=g).to(DEV)
    xl = torch.tensor([14, 11, 9], dtype=torch.int32, device=DEV)
    yl = torch.tensor([5, 3, 4], dtype=torch.int32, device=DEV)
    yy = torch.randint(1, 30, (B, U), generator=g).to(DEV)
    for b in range(B):
        yy[b, int(yl[b]):] = 0
    lt = tied(x, xl, yy, yl)
    lt.backward()
    lf = free(x, xl, yy, yl)
    lf.backward()
    torch.cuda.synchronize()
    want = free.joint.project_layer.weight.grad + free.decoder.embedding.weight.grad
    got = tied.joint.project_layer.weight.grad
    return {"loss_equal": abs(float(lt) - float(lf)) / abs(float(lf)), "shared_grad_rel_l2": rel_l2(got, want),
            "emb_part_norm": float(free.decoder.embedding.weight.grad.norm()),
            "w2_part_norm": float(free.joint.project_layer.weight.grad.norm())}


def transducer_vs_reference_fixture():
    """the assembled Transducer with the reference's weights (fixture from the unmodified
    reference): recognize() must return the reference's greedy tokens; the training forward (fused loss,
    folded projections) is checked against the float64 oracle fed the reference's encoder/decoder
    outputs, and against the unfolded variant."""
    from tests.helpers import load_ref_model_tiny
    model, g = load_ref_model_tiny()
    model = model.to(DEV)
    t = lambda k: torch.from_numpy(g[k])
    tokens = model.recognize(t("inputs").to(DEV), t("inputs_length"))
    want = [g["tokens"][i, :int(g["token_counts"][i])].tolist() for i in range(len(tokens))]
    args = (t("inputs").to(DEV), t("inputs_length").to(DEV), t("targets").to(DEV), t("targets_length").to(DEV))
    model.train()
    loss = model(*args)
    loss.backward()
    grads_finite = all(p.grad is not None and bool(torch.isfinite(p.grad).all()) for p in model.parameters())
    import copy
    m2 = copy.deepcopy(model)
    m2.fold_projections = False
    loss2 = m2(*args)
    loss3 = model(args[0], t("inputs_length"), args[2], t("targets_length"))   # lengths on the host
    sd = {k: v.detach().cpu().double() for k, v in model.joint.state_dict().items()}
    Umax = int(g["targets_length"].max())
    lo = O.joint_rnnt_loss(t("enc_state").double(), t("dec_state").double(), sd["forward_layer.weight"],
                           sd["forward_layer.bias"], sd["project_layer.weight"], sd["project_layer.bias"],
                           t("targets")[:, :Umax].int(), t("inputs_length").int(), t("targets_length").int(),
                           reduction="mean")
    torch.cuda.synchronize()
    return {"tokens_equal": tokens == want, "tokens": tokens, "want": want, "grads_finite": grads_finite,
            "loss_rel_oracle": float(abs(float(loss) - float(lo)) / abs(float(lo))),
            "loss_rel_unfolded": float(abs(float(loss) - float(loss2)) / abs(float(loss2))),
            "loss_rel_host_lens": float(abs(float(loss) - float(loss3)) / abs(float(loss)))}


def dw2_gemm_guard_bands(rows=192, V=300, H=128):
    """compute-sanitizer is not available on the GPU pool: run the GEMM with its outputs embedded
    in larger poisoned buffers and check that nothing outside the partial slots was touched."""
    from rnnt_transducer_b200 import _lib
    L = _lib.lib()
    Vds, hp = L.rnnt_b200_ds_pitch(V), L.rnnt_b200_h_pitch(H)
    g = torch.Generator().manual_seed(1)
    ds = (torch.randn(rows, Vds, generator=g) * 0.05).to(torch.bfloat16).to(DEV)
    h = torch.tanh(torch.randn(rows, hp, generator=g)).to(torch.bfloat16).to(DEV)
    ns = L.rnnt_b200_dw2_splits(rows, V, H, 0)
    pad = 4096
    big = torch.full((pad + ns * Vds * H + pad,), 7.0, device=DEV)
    bigb = torch.full((pad + ns * Vds + pad,), 7.0, device=DEV)
    bigb[pad:pad + ns * Vds] = 0
    st = torch.cuda.current_stream().cuda_stream
    rc = L.rnnt_b200_dw2_gemm(ds.data_ptr(), h.data_ptr(), rows, V, H, 0, big.data_ptr() + pad * 4,
                              bigb.data_ptr() + pad * 4, st)
    torch.cuda.synchronize()
    ok = bool((big[:pad] == 7).all() and (big[pad + ns * Vds * H:] == 7).all() and
              (bigb[:pad] == 7).all() and (bigb[pad + ns * Vds:] == 7).all())
    ref = ds.double().t() @ h[:, :H].double()
    got = big[pad:pad + ns * Vds * H].view(ns, Vds, H).double().sum(0)
    return {"rc": rc, "guards_intact": ok, "dw2_rel_l2": rel_l2(got.cpu(), ref.cpu())}
