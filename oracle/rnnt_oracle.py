"""CPU oracle for the RNN-T joint + transducer loss (TEST INFRASTRUCTURE ONLY).

This file is the parity checker for the CUDA path. It is NOT part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import it. The product path never routes through it.

What it restates (citations into /root/reference):

* ``joint_logits``      -- ``rnnt/model.py:18-37`` (``JointNet.forward``, 3-D branch):
  ``logits[b,t,u,:] = W2 . tanh(W1 . cat(enc[b,t], dec[b,u]) + b1) + b2`` with the
  encoder half of ``forward_layer.weight`` first (``model.py:31``).
* ``rnnt_loss``         -- the call ``rnnt/model.py:70``
  ``RNNTLoss()(logits, targets.int(), inputs_length.int(), targets_length.int())``.
  The arithmetic of that call lives in third-party ``warprnnt_pytorch``
  (HawkAaron/warp-transducer), which is NOT in /root/reference, has no pinned
  version (README.md:4-6 only names it) and cannot be installed offline. We
  therefore restate the published algorithm (Graves 2012, eq. 16-20; the
  log-space alpha/beta recursion warp-transducer implements) in float64.
  blank = 0 is pinned by the reference itself (``model.py:64``, ``decoder.py:9``).
  The default reduction (``'mean'`` = sum_b loss_b / B, one element) is what the
  caller needs (``train.py:43,47``: ``.item()`` on the result).

PARITY STATUS: **parity unpinned** for the loss half -- the reference ships no test,
golden vector or known-answer value for ``RNNTLoss`` and its dependency is absent.
The oracle is instead pinned three ways (tests/test_oracle.py, tests/golden/):
  1. the reference's own ``JointNet`` class, imported from /root/reference under a
     ``warprnnt_pytorch`` stub, generated the golden joint logits;
  2. ``torchaudio.functional.rnnt_loss`` (an independent compiled CPU
     implementation, NOT the reference's dependency) generated golden losses/grads;
  3. closed-form cases (T=1,U=0; uniform logits -> (T+U) log V - log C(T+U-1,U)).
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch


def joint_logits(enc_state: torch.Tensor, dec_state: torch.Tensor,
                 w1: torch.Tensor, b1: torch.Tensor,
                 w2: torch.Tensor, b2: torch.Tensor) -> torch.Tensor:
    """Restates JointNet.forward 3-D branch (rnnt/model.py:18-37) without repeat/cat.

    enc_state (B,T,E), dec_state (B,U1,D); w1 (H,E+D) encoder columns first
    (model.py:31), b1 (H), w2 (V,H), b2 (V). Returns (B,T,U1,V).
    """
    E = enc_state.shape[-1]
    pe = enc_state @ w1[:, :E].t() + b1            # (B,T,H)
    pd = dec_state @ w1[:, E:].t()                 # (B,U1,H)
    h = torch.tanh(pe.unsqueeze(2) + pd.unsqueeze(1))
    return h @ w2.t() + b2


def log_probs_planes(logits: torch.Tensor, labels: torch.Tensor, blank: int = 0
                     ) -> Tuple[torch.Tensor, torch.Tensor]:
    """lp_blank[b,t,u] = log_softmax(logits)[b,t,u,blank];
    lp_label[b,t,u] = log_softmax(logits)[b,t,u,labels[b,u]] for u < U (else -inf)."""
    B, T, U1, V = logits.shape
    lp = torch.log_softmax(logits, dim=-1)
    lpb = lp[..., blank]
    idx = torch.zeros(B, U1, dtype=torch.long, device=logits.device)
    if U1 > 1:
        idx[:, :U1 - 1] = labels.long()
    lpl = lp.gather(3, idx.view(B, 1, U1, 1).expand(B, T, U1, 1)).squeeze(3)
    mask = torch.zeros(B, 1, U1, dtype=torch.bool, device=logits.device)
    mask[:, :, U1 - 1] = True
    lpl = lpl.masked_fill(mask, float("-inf"))
    return lpb, lpl


def alpha_from_planes(lpb: torch.Tensor, lpl: torch.Tensor) -> torch.Tensor:
    """Forward variable on the padded (T,U1) grid, all utterances at once.

    alpha[0,0] = 0; alpha[t,u] = logaddexp(alpha[t-1,u] + lpb[t-1,u],
                                           alpha[t,u-1] + lpl[t,u-1]).
    Row scan: with L_u = sum_{j<u} lpl[t,j],
        alpha[t,u] = L_u + logcumsumexp_u(a_u - L_u),  a_u = alpha[t-1,u]+lpb[t-1,u].
    Cells outside an utterance's (T_b, U_b+1) window never feed cells inside it.
    """
    B, T, U1 = lpb.shape
    neg_inf = float("-inf")
    rows = []
    lpl_f = lpl.clone()
    lpl_f[:, :, U1 - 1] = 0.0  # unused column; keep the cumsum finite
    Lc = torch.cumsum(lpl_f, dim=2)
    L = torch.cat([torch.zeros(B, T, 1, dtype=lpb.dtype, device=lpb.device), Lc[:, :, :-1]], dim=2)
    for t in range(T):
        if t == 0:
            a = torch.full((B, U1), neg_inf, dtype=lpb.dtype, device=lpb.device)
            a[:, 0] = 0.0
        else:
            a = rows[-1] + lpb[:, t - 1, :]
        rows.append(L[:, t] + torch.logcumsumexp(a - L[:, t], dim=1))
    return torch.stack(rows, dim=1)


def rnnt_loss(logits: torch.Tensor, labels: torch.Tensor, act_lens: torch.Tensor,
              label_lens: torch.Tensor, blank: int = 0, reduction: str = "none"
              ) -> torch.Tensor:
    """Float64 log-space DP restating the call at rnnt/model.py:70.

    logits (B,T,U1,V) raw activations (no softmax applied, model.py:61-72);
    labels (B,U1-1) int, zero padded; act_lens (B); label_lens (B).
    Gradients come from autograd. reduction: 'none' -> (B,), 'sum', 'mean' (= sum/B).
    """
    x = logits.double()
    B, T, U1, V = x.shape
    lpb, lpl = log_probs_planes(x, labels, blank)
    alpha = alpha_from_planes(lpb, lpl)
    bi = torch.arange(B, device=x.device)
    tl = act_lens.long() - 1
    ul = label_lens.long()
    ll = alpha[bi, tl, ul] + lpb[bi, tl, ul]
    loss = -ll
    if reduction == "none":
        return loss
    if reduction == "sum":
        return loss.sum().reshape(1)
    if reduction == "mean":
        return (loss.sum() / B).reshape(1)
    raise ValueError(reduction)


def rnnt_loss_and_grad_explicit(logits: torch.Tensor, labels, act_lens, label_lens,
                                blank: int = 0):
    """Explicit alpha/beta + analytic gradient (no autograd), float64, loops in
    Python -- small cases only. This is the form the CUDA kernels implement:

      c_blank[t,u] = exp(alpha[t,u] + lpb[t,u] + beta[t+1,u] - ll)   (beta[T,U]:=0 at the exit cell)
      c_label[t,u] = exp(alpha[t,u] + lpl[t,u] + beta[t,u+1] - ll)
      dL/dlogit[t,u,v] = softmax[t,u,v] * (c_blank+c_label) - [v=blank] c_blank - [v=y_u+1] c_label
    Returns loss (B,), grad (B,T,U1,V), c_blank, c_label (B,T,U1).
    """
    x = logits.double()
    B, T, U1, V = x.shape
    lp = torch.log_softmax(x, dim=-1)
    loss = torch.zeros(B, dtype=torch.float64)
    grad = torch.zeros_like(x)
    cb_all = torch.zeros(B, T, U1, dtype=torch.float64)
    cl_all = torch.zeros(B, T, U1, dtype=torch.float64)
    NEG = float("-inf")
    for b in range(B):
        Tb, Ub = int(act_lens[b]), int(label_lens[b])
        y = [int(v) for v in labels[b][:Ub]]
        al = [[NEG] * (Ub + 1) for _ in range(Tb)]
        be = [[NEG] * (Ub + 1) for _ in range(Tb)]
        lpb = lp[b, :, :, blank].tolist()
        lpl = [[lp[b, t, u, y[u]].item() if u < Ub else NEG for u in range(Ub + 1)] for t in range(Tb)]

        def lae(a, c):
            if a == NEG:
                return c
            if c == NEG:
                return a
            m = max(a, c)
            return m + math.log1p(math.exp(-abs(a - c)))

        for t in range(Tb):
            for u in range(Ub + 1):
                if t == 0 and u == 0:
                    al[t][u] = 0.0
                    continue
                a = al[t - 1][u] + lpb[t - 1][u] if t > 0 else NEG
                c = al[t][u - 1] + lpl[t][u - 1] if u > 0 else NEG
                al[t][u] = lae(a, c)
        for t in range(Tb - 1, -1, -1):
            for u in range(Ub, -1, -1):
                if t == Tb - 1 and u == Ub:
                    be[t][u] = lpb[t][u]
                    continue
                a = be[t + 1][u] + lpb[t][u] if t + 1 < Tb else NEG
                c = be[t][u + 1] + lpl[t][u] if u + 1 <= Ub else NEG
                be[t][u] = lae(a, c)
        ll = be[0][0]
        loss[b] = -ll
        for t in range(Tb):
            for u in range(Ub + 1):
                if t + 1 < Tb:
                    cb = math.exp(al[t][u] + lpb[t][u] + be[t + 1][u] - ll)
                elif u == Ub:
                    cb = math.exp(al[t][u] + lpb[t][u] - ll)
                else:
                    cb = 0.0
                cl = math.exp(al[t][u] + lpl[t][u] + be[t][u + 1] - ll) if u < Ub else 0.0
                cb_all[b, t, u] = cb
                cl_all[b, t, u] = cl
                g = torch.exp(lp[b, t, u]) * (cb + cl)
                g[blank] -= cb
                if u < Ub:
                    g[y[u]] -= cl
                grad[b, t, u] = g
    return loss, grad, cb_all, cl_all


def joint_rnnt_loss(enc_state, dec_state, w1, b1, w2, b2, labels, act_lens, label_lens,
                    blank: int = 0, reduction: str = "mean", dtype=torch.float64,
                    quantize_bf16: bool = False) -> torch.Tensor:
    """model.py:68-70 end to end in ``dtype``. With ``quantize_bf16`` the hidden
    activations h and W2 are rounded to bf16 before the vocab contraction (the
    operand precision of the tensor-core path) while everything else stays in
    ``dtype`` -- isolates kernel error from quantisation error."""
    enc_state, dec_state = enc_state.to(dtype), dec_state.to(dtype)
    w1, b1, w2, b2 = w1.to(dtype), b1.to(dtype), w2.to(dtype), b2.to(dtype)
    E = enc_state.shape[-1]
    pe = enc_state @ w1[:, :E].t() + b1
    pd = dec_state @ w1[:, E:].t()
    if quantize_bf16:
        # the CUDA path adds the two fp32 projections in fp32, then rounds tanh to bf16
        pre = (pe.float().unsqueeze(2) + pd.float().unsqueeze(1))
        h = _ste_round_bf16(torch.tanh(pre.to(dtype)))
        w2q = _ste_round_bf16(w2)
        logits = h @ w2q.t() + b2
    else:
        h = torch.tanh(pe.unsqueeze(2) + pd.unsqueeze(1))
        logits = h @ w2.t() + b2
    return rnnt_loss(logits, labels, act_lens, label_lens, blank, reduction)


class _RoundBF16(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).to(x.dtype)

    @staticmethod
    def backward(ctx, g):
        return g


def _ste_round_bf16(x):
    return _RoundBF16.apply(x)


def uniform_closed_form(T: int, U: int, V: int) -> float:
    """Loss for all-equal logits: every path has T+U steps of prob 1/V and there are
    C(T+U-1, U) monotone paths ending with the final blank."""
    return (T + U) * math.log(V) - math.log(math.comb(T + U - 1, U))
