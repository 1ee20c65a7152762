"""Host logic of the multi-GPU path: the hot path shards by utterance with no
data-path exchange; the only collectives are the loss sum and the parameter-gradient
all-reduce. Replaces the reference's nn.DataParallel (train.py:42-43,175-177). Backend-agnostic
(NCCL on GPUs; the CPU tests run it under gloo)."""
from __future__ import annotations

from typing import List, Sequence

import torch
import torch.distributed as dist

__all__ = ["shard_batch", "global_mean_loss", "allreduce_grads_and_loss", "bind_flat_grads"]


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
