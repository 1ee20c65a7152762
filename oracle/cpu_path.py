"""CPU restatement of the reference's hot path, for TIMING (bench.py's cpu_baseline and
``--impl reference`` legs) and as a checker. TEST/BENCH INFRASTRUCTURE ONLY -- the product
never imports this.

What is timed follows the reference's own CPU code path step by step:
  * joint  -- rnnt/model.py:18-37: unsqueeze, ``repeat`` both states to (B,T,U+1,.), ``cat``,
    ``forward_layer``, tanh, ``project_layer`` -> materialised (B,T,U+1,V) float32 logits,
    executed by torch's CPU kernels (MKL GEMM) exactly as the reference would on CPU;
  * loss   -- rnnt/model.py:70. The reference's implementation is warprnnt_pytorch's CPU
    path, which is absent (un-vendored, un-pinned, not installable offline); the port is
    oracle/rnnt_oracle.c (float64 DP, OpenMP over lattice rows) -> ``kind: "port"``;
  * backward -- train.py:45: gradient wrt logits from the loss, then torch autograd back
    through the joint to enc_state, dec_state and the four joint parameters.
/root/reference does not exist on the GPU box, so nothing here reads it.
"""
from __future__ import annotations

import os
import time

import numpy as np
import torch

from . import c_oracle


class RefJointPort(torch.nn.Module):
    """Port of the reference JointNet's materialising forward (model.py:18-37)."""

    def __init__(self, input_size, inner_dim, vocab_size):
        super().__init__()
        self.forward_layer = torch.nn.Linear(input_size, inner_dim, bias=True)
        self.project_layer = torch.nn.Linear(inner_dim, vocab_size, bias=True)

    def forward(self, enc_state, dec_state):
        T, U1 = enc_state.size(1), dec_state.size(1)
        e = enc_state.unsqueeze(2).repeat([1, 1, U1, 1])     # materialised, as the reference does
        d = dec_state.unsqueeze(1).repeat([1, T, 1, 1])
        z = torch.cat((e, d), dim=-1)
        return self.project_layer(torch.tanh(self.forward_layer(z)))


class _CLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, act_lens, label_lens):
        loss, grad = c_oracle.rnnt_loss(logits.detach().numpy(), labels.numpy(), act_lens.numpy(),
                                        label_lens.numpy(), blank=0, want_grad=True)
        ctx.save_for_backward(torch.from_numpy(grad))
        return torch.from_numpy(loss)

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g.view(-1, 1, 1, 1).to(grad.dtype), None, None, None


def cpu_step(joint: RefJointPort, enc, dec, labels, act_lens, label_lens):
    """One fwd+bwd of model.py:68,70 + train.py:45 on the CPU. Returns per-utterance loss (float64)."""
    for p in joint.parameters():
        p.grad = None
    enc = enc.detach().requires_grad_(True)
    dec = dec.detach().requires_grad_(True)
    logits = joint(enc, dec)
    loss_b = _CLoss.apply(logits, labels, act_lens, label_lens)
    (loss_b.sum() / loss_b.shape[0]).backward()
    return loss_b.detach(), enc.grad, dec.grad


def time_cpu_path(E, D, H, V, T, U, n_utts=1, steps=1, warmup=0, seed=2019):
    """Times ``steps`` CPU fwd+bwd passes, each over ``n_utts`` full-length utterances processed
    one at a time (bounds memory: one utterance's logits at AISHELL cfg3 are 855 MB).
    Returns dict(value=cells/s, cells=..., seconds=[...], cores=...)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    gen = torch.Generator().manual_seed(seed)
    joint = RefJointPort(E + D, H, V)
    enc = torch.randn(n_utts, T, E, generator=gen)
    dec = torch.randn(n_utts, U + 1, D, generator=gen)
    labels = torch.randint(1, V, (n_utts, U), generator=gen, dtype=torch.int32)
    tl = torch.full((1,), T, dtype=torch.int32)
    ul = torch.full((1,), U, dtype=torch.int32)
    cells = n_utts * T * (U + 1)
    secs = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        for b in range(n_utts):
            cpu_step(joint, enc[b:b + 1], dec[b:b + 1], labels[b:b + 1], tl, ul)
        dt = time.perf_counter() - t0
        if it >= warmup:
            secs.append(dt)
    med = float(np.median(secs))
    return {"value": cells / med, "cells": cells, "seconds": secs, "cores": cores,
            "omp_threads": c_oracle.num_threads()}
