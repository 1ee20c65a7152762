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


def make_joint_sample(wl, T_s, cores, seed=2019):
    """Returns a zero-argument callable: ONE CPU fwd+bwd of the joint + loss path over one utterance of the
    workload's shape cut to its first ``T_s`` frames (full label length). Used by bench.py's cpu_baseline and
    ``--impl reference`` legs."""
    torch.set_num_threads(cores)
    E, D, H, V, U = wl["E"], wl["D"], wl["H"], wl["V"], wl["U"]
    gen = torch.Generator().manual_seed(seed)
    joint = RefJointPort(E + D, H, V)
    enc = torch.randn(1, T_s, E, generator=gen)
    dec = torch.randn(1, U + 1, D, generator=gen)
    labels = torch.randint(1, V, (1, U), generator=gen, dtype=torch.int32)
    tl = torch.full((1,), T_s, dtype=torch.int32)
    ul = torch.full((1,), U, dtype=torch.int32)
    return lambda: cpu_step(joint, enc, dec, labels, tl, ul)


class RefTransducerPort(torch.nn.Module):
    """Port of the reference's assembled model for CPU timing (BASELINE.json configs[3]): encoder
    rnnt/encoder.py:5-40 (LSTM + output_proj), decoder rnnt/decoder.py:5-44 (embedding, LSTM, output_proj),
    joint rnnt/model.py:9-37 (materialised logits), loss rnnt/model.py:70 -> oracle/rnnt_oracle.c."""

    def __init__(self, cfg):
        super().__init__()
        enc, dec = cfg["enc"], cfg["dec"]
        self.enc_lstm = torch.nn.LSTM(cfg["feature_dim"], enc["hidden_size"], enc["n_layers"], batch_first=True,
                                      dropout=cfg["dropout"], bidirectional=enc["bidirectional"])
        self.enc_proj = torch.nn.Linear((2 if enc["bidirectional"] else 1) * enc["hidden_size"], enc["output_size"])
        self.embedding = torch.nn.Embedding(cfg["vocab_size"], dec["hidden_size"], padding_idx=0)
        self.dec_lstm = torch.nn.LSTM(dec["hidden_size"], dec["hidden_size"], dec["n_layers"], batch_first=True)
        self.dec_proj = torch.nn.Linear(dec["hidden_size"], dec["output_size"])
        self.joint = RefJointPort(cfg["joint"]["input_size"], cfg["joint"]["inner_size"], cfg["vocab_size"])

    def forward(self, x, y, tl, ul):
        enc = self.enc_proj(self.enc_lstm(x)[0])
        ypad = torch.nn.functional.pad(y, pad=(1, 0, 0, 0), value=0)          # model.py:64
        dec = self.dec_proj(self.dec_lstm(self.embedding(ypad.long()))[0])
        logits = self.joint(enc, dec)
        return _CLoss.apply(logits, y.int(), tl, ul).sum() / x.shape[0]


def make_train_step_sample(wl, T_s, cores, seed=2019):
    """As make_joint_sample for the full training step (train.py:24-52: forward, backward, grad-norm clip, SGD)."""
    torch.set_num_threads(cores)
    cfg = dict(enc=dict(hidden_size=320, output_size=wl["E"], n_layers=4, bidirectional=True),
               dec=dict(hidden_size=512, output_size=wl["D"], n_layers=1),
               joint=dict(input_size=wl["E"] + wl["D"], inner_size=wl["H"]), vocab_size=wl["V"], feature_dim=160,
               dropout=0.3)                                                   # config/aishell.yaml:21-39
    torch.manual_seed(seed)
    model = RefTransducerPort(cfg).train()
    opt = torch.optim.SGD(model.parameters(), lr=1e-4, momentum=0.9)
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(1, T_s, 160, generator=gen)
    y = torch.randint(1, wl["V"], (1, wl["U"]), generator=gen)
    tl = torch.full((1,), T_s, dtype=torch.int32)
    ul = torch.full((1,), wl["U"], dtype=torch.int32)

    def step():
        opt.zero_grad()
        loss = model(x, y, tl, ul)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 200.0)
        opt.step()
        return loss.detach()
    return step
