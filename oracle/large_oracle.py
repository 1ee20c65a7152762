"""Per-utterance, t-chunked float64 oracle for the fused joint + transducer loss at the BASELINE
shapes (TEST INFRASTRUCTURE ONLY -- SURVEY.md 8(c) "Large shapes"; never imported by the product).

``oracle/rnnt_oracle.py::joint_rnnt_loss`` materialises the (B,T,U+1,V) float64 logits and their
autograd graph: 1.7 GB per utterance at BASELINE configs[2] (T=500, U=100, V=4232), 10 GB at configs[4].
This file restates the same computation (rnnt/model.py:18-37 joint, model.py:70 loss, train.py:45 backward)
in bounded memory, exactly (no approximation):

  pass 1, per chunk of lattice rows t: logits of the chunk -> log-softmax -> the two planes the lattice
          reads, lp_blank[t,u] and lp_label[t,u] (float64);
  lattice: alpha / beta over anti-diagonals (numpy float64), log-likelihood ll, and the occupation
          coefficients c_blank = exp(alpha + lp_blank + beta[t+1,u] - ll), c_label = exp(alpha + lp_label +
          beta[t,u+1] - ll) -- which are minus the partial derivatives of loss = -ll wrt the two planes;
  pass 2, per chunk: the planes recomputed WITH autograd from enc/dec/W1/b1/W2/b2 and the surrogate
          -(sum c_blank*lp_blank + sum c_label*lp_label) back-propagated: by the chain rule its gradient is the
          chunk's contribution to d loss / d (enc, dec, W1, b1, W2, b2). Autograd differentiates the same
          operations ``rnnt_oracle.joint_rnnt_loss`` uses (same bf16 straight-through rounding when
          ``quantize_bf16``), so the two oracles agree to float64 rounding; tests/test_oracle.py pins that.

Loss reduction: gradients are those of sum_b loss_b (reduction 'sum'); per-utterance losses are returned.
"""
from __future__ import annotations

import numpy as np
import torch

from .rnnt_oracle import _ste_round_bf16


def lattice_coeffs(lpb: np.ndarray, lpl: np.ndarray):
    """lpb, lpl: float64 (T,U1) planes of ONE utterance's valid window (lpl[:, U1-1] unused).
    Returns (ll, c_blank (T,U1), c_label (T,U1)); same recursion as rnnt_oracle.c, walked by anti-diagonals."""
    T, U1 = lpb.shape
    NEG = -np.inf
    alpha = np.full((T, U1), NEG)
    beta = np.full((T, U1), NEG)
    alpha[0, 0] = 0.0
    for d in range(1, T + U1 - 1):
        u = np.arange(max(0, d - (T - 1)), min(d, U1 - 1) + 1)
        t = d - u
        top = np.full(u.shape, NEG)
        m = t >= 1
        top[m] = alpha[t[m] - 1, u[m]] + lpb[t[m] - 1, u[m]]
        left = np.full(u.shape, NEG)
        m = u >= 1
        left[m] = alpha[t[m], u[m] - 1] + lpl[t[m], u[m] - 1]
        alpha[t, u] = np.logaddexp(top, left)
    beta[T - 1, U1 - 1] = lpb[T - 1, U1 - 1]
    for d in range(T + U1 - 3, -1, -1):
        u = np.arange(max(0, d - (T - 1)), min(d, U1 - 1) + 1)
        t = d - u
        down = np.full(u.shape, NEG)
        m = t + 1 < T
        down[m] = beta[t[m] + 1, u[m]] + lpb[t[m], u[m]]
        right = np.full(u.shape, NEG)
        m = u + 1 < U1
        right[m] = beta[t[m], u[m] + 1] + lpl[t[m], u[m]]
        beta[t, u] = np.logaddexp(down, right)
    ll = beta[0, 0]
    cb = np.zeros((T, U1))
    cl = np.zeros((T, U1))
    cb[:-1] = np.exp(alpha[:-1] + lpb[:-1] + beta[1:] - ll)
    cb[T - 1, U1 - 1] = np.exp(alpha[T - 1, U1 - 1] + lpb[T - 1, U1 - 1] - ll)
    if U1 > 1:
        cl[:, :-1] = np.exp(alpha[:, :-1] + lpl[:, :-1] + beta[:, 1:] - ll)
    return float(ll), cb, cl


def _planes_chunk(pe, pd, w2, b2, y_idx, quantize_bf16):
    """pe (n,H), pd (U1,H) float64 -> (lp_blank (n,U1), lp_label (n,U1)) of the chunk; label column U1-1 unused."""
    if quantize_bf16:
        # the CUDA path adds the two fp32 projections in fp32, then rounds tanh to bf16
        pre = pe.float().unsqueeze(1) + pd.float().unsqueeze(0)
        h = _ste_round_bf16(torch.tanh(pre.double()))
        w2u = _ste_round_bf16(w2)
    else:
        h = torch.tanh(pe.unsqueeze(1) + pd.unsqueeze(0))
        w2u = w2
    lp = torch.log_softmax(h @ w2u.t() + b2, dim=-1)          # (n,U1,V)
    n, U1 = lp.shape[0], lp.shape[1]
    lpb = lp[..., 0]
    lpl = lp.gather(2, y_idx.view(1, U1, 1).expand(n, U1, 1)).squeeze(2)
    return lpb, lpl


def utterance_loss_and_grads(enc, dec, w1, b1, w2, b2, y, quantize_bf16=False, cells_per_chunk=2048, grads=True):
    """One utterance, already trimmed to its valid window: enc (T_b,E), dec (U_b+1,D), y (U_b) labels.
    Returns (loss, dict of float64 gradients of that loss: enc, dec, w1, b1, w2, b2; None if not ``grads``)."""
    enc, dec = enc.double(), dec.double()
    leaves = [t.detach().double().clone().requires_grad_(True) for t in (enc, dec, w1, b1, w2, b2)]
    e, d, W1, B1, W2, B2 = leaves
    T, E = e.shape
    U1 = d.shape[0]
    y_idx = torch.zeros(U1, dtype=torch.long)
    y_idx[:U1 - 1] = torch.as_tensor(y).long()[:U1 - 1]
    rows = max(1, cells_per_chunk // U1)
    lpb = np.empty((T, U1))
    lpl = np.empty((T, U1))
    with torch.no_grad():
        pe = e @ W1[:, :E].t() + B1
        pd = d @ W1[:, E:].t()
        for t0 in range(0, T, rows):
            a, b = _planes_chunk(pe[t0:t0 + rows], pd, W2, B2, y_idx, quantize_bf16)
            lpb[t0:t0 + rows] = a.numpy()
            lpl[t0:t0 + rows] = b.numpy()
    ll, cb, cl = lattice_coeffs(lpb, lpl)
    if not grads:
        return -ll, None
    cb_t, cl_t = torch.from_numpy(cb), torch.from_numpy(cl)
    cl_t[:, U1 - 1] = 0.0
    for t0 in range(0, T, rows):
        pe = e[t0:t0 + rows] @ W1[:, :E].t() + B1
        pd = d @ W1[:, E:].t()
        a, b = _planes_chunk(pe, pd, W2, B2, y_idx, quantize_bf16)
        s = -((cb_t[t0:t0 + rows] * a).sum() + (cl_t[t0:t0 + rows, :U1 - 1] * b[:, :U1 - 1]).sum())
        s.backward()
    names = ["enc", "dec", "w1", "b1", "w2", "b2"]
    return -ll, {n: t.grad for n, t in zip(names, leaves)}


def joint_rnnt_loss_and_grads(enc_state, dec_state, w1, b1, w2, b2, labels, act_lens, label_lens,
                              quantize_bf16=False, utterances=None, cells_per_chunk=2048, grads=True):
    """enc_state (B,T,E), dec_state (B,U1,D), labels (B,U1-1) zero padded. Returns (loss (B,) float64, grads)
    with grads = d (sum over the chosen utterances of loss_b) / d {enc (B,T,E), dec (B,U1,D), w1, b1, w2, b2},
    zero outside the valid windows and for utterances not in ``utterances`` (default: all)."""
    B = enc_state.shape[0]
    utts = list(range(B)) if utterances is None else list(utterances)
    loss = torch.zeros(B, dtype=torch.float64)
    g = {"enc": torch.zeros(enc_state.shape, dtype=torch.float64),
         "dec": torch.zeros(dec_state.shape, dtype=torch.float64),
         "w1": torch.zeros(w1.shape, dtype=torch.float64), "b1": torch.zeros(b1.shape, dtype=torch.float64),
         "w2": torch.zeros(w2.shape, dtype=torch.float64), "b2": torch.zeros(b2.shape, dtype=torch.float64)}
    for b in utts:
        Tb, Ub = int(act_lens[b]), int(label_lens[b])
        l, gb = utterance_loss_and_grads(enc_state[b, :Tb], dec_state[b, :Ub + 1], w1, b1, w2, b2, labels[b, :Ub],
                                         quantize_bf16, cells_per_chunk, grads)
        loss[b] = l
        if not grads:
            continue
        g["enc"][b, :Tb] = gb["enc"]
        g["dec"][b, :Ub + 1] = gb["dec"]
        for n in ("w1", "b1", "w2", "b2"):
            g[n] += gb[n]
    return loss, (g if grads else None)
