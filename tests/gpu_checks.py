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


def selftest_gemm(M, N, K, seed=0):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(M, K, generator=g).to(torch.bfloat16)
    b = torch.randn(N, K, generator=g).to(torch.bfloat16)
    c = R.selftest_gemm(a.to(DEV), b.to(DEV)).cpu()
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
    out["loss_mean"] = float(loss.mean())
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
