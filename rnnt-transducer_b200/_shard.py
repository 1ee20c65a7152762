"""Host logic of the multi-GPU path: the hot path shards by utterance with no
data-path exchange; the only collectives are the loss sum and the parameter-gradient
all-reduce. Replaces the reference's nn.DataParallel (train.py:42-43,175-177). Backend-agnostic
(NCCL on GPUs; the CPU tests run it under gloo)."""
from __future__ import annotations

from typing import List, Sequence

import torch
import torch.distributed as dist

__all__ = ["shard_batch", "shard_batch_weighted", "global_mean_loss", "allreduce_grads_and_loss", "bind_flat_grads"]


def shard_batch(act_lens: Sequence[int], label_lens: Sequence[int], world_size: int) -> List[List[int]]:
    """Assign utterances to ranks balancing lattice cells T_b*(U_b+1): sort by cell count
    descending and deal in snake order (0..N-1, N-1..0, ...). Returns per-rank index lists;
    every utterance appears exactly once. Ranks may differ by at most one utterance."""
    cells = [int(t) * (int(u) + 1) for t, u in zip(act_lens, label_lens)]
    order = sorted(range(len(cells)), key=lambda i: (-cells[i], i))
    shards: List[List[int]] = [[] for _ in range(world_size)]
    for j, idx in enumerate(order):
        r = j % world_size
        if (j // world_size) & 1:
            r = world_size - 1 - r
        shards[r].append(idx)
    return shards


def shard_batch_weighted(act_lens: Sequence[int], label_lens: Sequence[int], rates: Sequence[float],
                         min_spread: float = 0.015, damping: float = 1.0) -> List[List[int]]:
    """Utterances to ranks in proportion to the ranks' measured speeds (``rates[r]``: lattice cells per unit time of
    rank r, e.g. from a few untimed steps): the GPUs of one box settle at different clocks under their power caps
    (8 % apart on an 8-GPU box), and a synchronous step waits for the slowest. Longest-processing-time greedy: the
    utterances in descending cell count, each to the rank whose finishing time (load + cells) / rate stays
    smallest. With rates closer than ``min_spread`` the equal split of ``shard_batch`` is returned. ``damping`` < 1
    corrects only part of the measured difference (rates ** damping): a power-capped GPU that is given less work
    idles in the step's all-reduce and regains clock headroom, one that is given more loses it. Deterministic: every
    rank computes the same assignment from the same (all-gathered) rates."""
    world = len(rates)
    lo, hi = min(rates), max(rates)
    if world == 1 or lo <= 0 or (hi - lo) / hi < min_spread:
        return shard_batch(act_lens, label_lens, world)
    rates = [float(r) ** damping for r in rates]
    cells = [int(t) * (int(u) + 1) for t, u in zip(act_lens, label_lens)]
    order = sorted(range(len(cells)), key=lambda i: (-cells[i], i))
    load = [0.0] * world
    shards: List[List[int]] = [[] for _ in range(world)]
    for idx in order:
        r = min(range(world), key=lambda q: ((load[q] + cells[idx]) / rates[q], q))
        shards[r].append(idx)
        load[r] += cells[idx]
    if any(not s_ for s_ in shards):      # never leave a rank without work
        return shard_batch(act_lens, label_lens, world)
    return shards


def global_mean_loss(local_loss_b: torch.Tensor, global_batch: int) -> torch.Tensor:
    """sum over ALL ranks' per-utterance losses / global batch (what RNNTLoss 'mean' would give on
    the unsharded batch). The returned value is the same on every rank; gradients flow only
    through the local term (scaled by 1/global_batch), as DDP expects."""
    local = local_loss_b.sum() / global_batch
    total = local.detach().clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(total)
    return local + (total - local.detach())


def allreduce_grads_and_loss(params, loss_sum: torch.Tensor) -> torch.Tensor:
    """One flat all-reduce carrying every parameter gradient plus the loss sum in its last slot
    (one collective launch instead of two). Gradients are written back in place; returns the
    global loss sum."""
    params = [p for p in params if p.grad is not None]
    flat = torch.cat([p.grad.reshape(-1) for p in params] + [loss_sum.detach().reshape(1).to(params[0].grad.dtype)])
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(flat)
    off = 0
    for p in params:
        n = p.numel()
        p.grad.copy_(flat[off:off + n].view_as(p.grad))
        off += n
    return flat[-1]


def bind_flat_grads(params, extra: int = 1) -> torch.Tensor:
    """One flat buffer behind every parameter gradient: each ``p.grad`` becomes a view of it, ``extra`` trailing
    elements are free for scalars that ride along (the loss sum). A step then ends in ONE all-reduce of the flat
    buffer -- no concatenation before it, no copy-back after it (``allreduce_grads_and_loss`` does both). Zero the
    buffer (``flat.zero_()``) instead of setting ``p.grad = None`` between steps: autograd accumulates into the views."""
    params = list(params)
    n = sum(p.numel() for p in params)
    flat = torch.zeros(n + extra, dtype=params[0].dtype, device=params[0].device)
    off = 0
    for p in params:
        p.grad = flat[off:off + p.numel()].view_as(p)
        off += p.numel()
    return flat
