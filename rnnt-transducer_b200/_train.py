"""The training / validation loop around the hot path: what
train.py:15-114 does per step, with the host overheads it pays removed --

* no ``.max().item()`` pair per step (train.py:30-31): lengths arrive on the host (as a DataLoader
  or ``KaldiFeatureSet.loader`` yields them) and are used there; the batch is trimmed with them;
* no ``loss.item()`` per step (train.py:47,53,64): the running loss is accumulated on the device
  and read back once every ``show_interval`` steps together with the gradient norm, and once at
  the end of the epoch;
* gradient clipping (train.py:49-50) stays on the device (``foreach`` kernels, no read-back);
* the optimiser (rnnt/optim.py:36-62, same hyper-parameters) uses torch's fused multi-tensor
  update where the type has one.

Under one process per GPU wrap the model in DistributedDataParallel and pass it as ``model``:
the loop itself holds no collective. The numbers logged are the reference's (loss of the step,
running average, gradient norm, learning rate). ``save_checkpoint`` / ``load_checkpoint`` keep
the reference's file layout (rnnt/utils.py:74-85, train.py:150-170) so checkpoints move both ways.
"""
from __future__ import annotations

import time
from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

__all__ = ["Optimizer", "train_epoch", "eval_epoch", "edit_distance", "compute_cer", "save_checkpoint",
           "load_checkpoint", "trim_batch"]


def _get(config, name, default=None):
    if isinstance(config, dict):
        return config.get(name, default)
    v = getattr(config, name, default)
    return default if v is None else v


class Optimizer:
    """rnnt/optim.py:4-33 -- same attributes (``global_step`` starts at 1, ``current_epoch``,
    ``lr``) and methods; ``config`` is the ``optim`` section (dict or attribute object)."""

    def __init__(self, parameters, config):
        self.config = config
        params = [p for p in parameters]
        self.optimizer = build_optimizer(params, config)
        self.global_step = 1
        self.current_epoch = 0
        self.lr = _get(config, "lr")
        self.decay_ratio = _get(config, "decay_ratio", 1.0)

    def step(self):
        self.global_step += 1
        self.optimizer.step()

    def epoch(self):
        self.current_epoch += 1

    def zero_grad(self):
        self.optimizer.zero_grad(set_to_none=True)

    # implementation-only switches of torch's optimisers: kept out of checkpoints so that the reference's plain
    # optim.SGD / Adam (rnnt/optim.py) loads them unchanged, and kept as configured here when loading
    _IMPL_KEYS = ("fused", "foreach", "capturable", "differentiable")

    def state_dict(self):
        sd = self.optimizer.state_dict()
        sd["param_groups"] = [{k: v for k, v in g.items() if k not in self._IMPL_KEYS} for g in sd["param_groups"]]
        return sd

    def load_state_dict(self, state_dict):
        mine = self.optimizer.param_groups
        groups = []
        for i, g in enumerate(state_dict["param_groups"]):
            g = {k: v for k, v in g.items() if k not in self._IMPL_KEYS}
            if i < len(mine):
                g.update({k: mine[i][k] for k in self._IMPL_KEYS if k in mine[i]})
            groups.append(g)
        self.optimizer.load_state_dict({"state": state_dict["state"], "param_groups": groups})
        self.lr = self.optimizer.param_groups[0]["lr"]   # the decayed rate the checkpoint carries

    def decay_lr(self):
        self.lr *= self.decay_ratio
        for group in self.optimizer.param_groups:
            group["lr"] = self.lr


def build_optimizer(params: List[torch.Tensor], config):
    kind = _get(config, "type")
    on_gpu = bool(params) and all(p.is_cuda for p in params)
    wd = _get(config, "weight_decay", 0)
    if kind == "adam":
        return torch.optim.Adam(params, lr=_get(config, "lr"), betas=(0.9, 0.98), eps=1e-8, weight_decay=wd,
                                fused=on_gpu or None)
    if kind == "sgd":
        return torch.optim.SGD(params, lr=_get(config, "lr"), momentum=_get(config, "momentum", 0),
                               nesterov=bool(_get(config, "nesterov", False)), weight_decay=wd, fused=on_gpu or None)
    if kind == "adadelta":
        return torch.optim.Adadelta(params, lr=_get(config, "lr"), rho=_get(config, "rho", 0.9),
                                    eps=_get(config, "eps", 1e-6), weight_decay=wd, foreach=on_gpu or None)
    raise NotImplementedError(kind)


def trim_batch(inputs, inputs_length, targets, targets_length):
    """train.py:30-33 without the device read-back when the lengths are host tensors."""
    t_max, u_max = int(inputs_length.max()), int(targets_length.max())
    return inputs[:, :t_max], targets[:, :u_max]


def train_epoch(model: nn.Module, batches: Iterable, optimizer: Optimizer, max_grad_norm: float,
                show_interval: int = 0, log: Optional[Callable[[Dict], None]] = None,
                step_wise_decay: bool = False, device=None) -> float:
    """One epoch of train.py:15-72. ``batches`` yields ``(inputs, inputs_length, targets,
    targets_length)``; tensors still on the host are moved (non-blocking), length vectors are left
    where they are. Returns the epoch's average loss (one read-back)."""
    model.train()
    optimizer.epoch()
    params = [p for p in model.parameters() if p.requires_grad]
    dev = torch.device(device) if device is not None else params[0].device
    total = torch.zeros((), dtype=torch.float32, device=dev)
    steps = 0
    t0 = time.perf_counter()
    for inputs, inputs_length, targets, targets_length in batches:
        inputs, targets = inputs.to(dev, non_blocking=True), targets.to(dev, non_blocking=True)
        inputs, targets = trim_batch(inputs, inputs_length, targets, targets_length)
        if step_wise_decay and hasattr(optimizer, "step_decay_lr"):
            optimizer.step_decay_lr()
        optimizer.zero_grad()
        loss = model(inputs, inputs_length, targets, targets_length)
        if loss.dim() > 0:
            loss = loss.mean()  # train.py:42-43
        loss.backward()
        total += loss.detach()
        grad_norm = nn.utils.clip_grad_norm_(params, max_grad_norm, foreach=True if dev.type == "cuda" else None)
        optimizer.step()
        steps += 1
        if log is not None and show_interval and optimizer.global_step % show_interval == 0:
            vals = torch.stack([loss.detach().float(), total, grad_norm.float()]).tolist()  # the one read-back
            log({"epoch": optimizer.current_epoch, "step": steps, "global_step": optimizer.global_step,
                 "lr": optimizer.lr, "grad_norm": vals[2], "loss": vals[0], "avg_loss": vals[1] / steps,
                 "run_time": time.perf_counter() - t0})
    return float(total) / max(steps, 1)


def edit_distance(a: Sequence[int], b: Sequence[int]) -> int:
    """Levenshtein distance between two token lists (what ``editdistance.eval`` returns,
    rnnt/utils.py:40-43)."""
    if len(a) < len(b):
        a, b = b, a
    prev = list(range(len(b) + 1))
    for i, x in enumerate(a, 1):
        cur = [i]
        for j, y in enumerate(b, 1):
            cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (x != y)))
        prev = cur
    return prev[-1]


def compute_cer(preds: Sequence[Sequence[int]], labels: Sequence[Sequence[int]]) -> Tuple[int, int]:
    """rnnt/utils.py:40-43: summed edit distance and summed reference length."""
    return (sum(edit_distance(list(l), list(p)) for l, p in zip(labels, preds)), sum(len(l) for l in labels))


@torch.no_grad()
def eval_epoch(model: nn.Module, batches: Iterable, show_interval: int = 0,
               log: Optional[Callable[[Dict], None]] = None, device=None) -> float:
    """train.py:74-113: greedy search over the validation batches, character error rate in %.
    The transcripts are cut on the host from host lengths; ``recognize`` reads its tokens back
    once per batch."""
    model.eval()
    net = model.module if hasattr(model, "module") else model
    dev = torch.device(device) if device is not None else next(net.parameters()).device
    dist = words = 0
    for step, (inputs, inputs_length, targets, targets_length) in enumerate(batches):
        inputs = inputs.to(dev, non_blocking=True)
        inputs, targets = trim_batch(inputs, inputs_length, targets, targets_length)
        preds = net.recognize(inputs, inputs_length)
        tgt, tl = targets.cpu(), targets_length.cpu()
        d, n = compute_cer(preds, [tgt[i, :int(tl[i])].tolist() for i in range(tgt.shape[0])])
        dist, words = dist + d, words + n
        if log is not None and show_interval and step % show_interval == 0:
            log({"step": step, "cer": dist / max(words, 1) * 100})
    return dist / max(words, 1) * 100


def save_checkpoint(model: nn.Module, optimizer: Optimizer, path: str) -> None:
    """rnnt/utils.py:74-85: per-module state dicts under 'encoder' / 'decoder' / 'joint'."""
    net = model.module if hasattr(model, "module") else model
    torch.save({"encoder": net.encoder.state_dict(), "decoder": net.decoder.state_dict(),
                "joint": net.joint.state_dict(), "optimizer": optimizer.state_dict(),
                "epoch": optimizer.current_epoch, "step": optimizer.global_step}, path)


def load_checkpoint(model: nn.Module, path: str, optimizer: Optional[Optimizer] = None,
                    parts: Sequence[str] = ("encoder", "decoder", "joint")) -> int:
    """train.py:150-170,188-191: load the listed modules (``load_model`` loads all three,
    ``load_encoder`` / ``load_decoder`` one) and, for ``-mode continue``, the optimiser state.
    Returns the epoch to resume from."""
    net = model.module if hasattr(model, "module") else model
    ckpt = torch.load(path, map_location="cpu")
    for part in parts:
        getattr(net, part).load_state_dict(ckpt[part])
    if optimizer is not None:
        optimizer.load_state_dict(ckpt["optimizer"])
        optimizer.current_epoch = ckpt.get("epoch", 0)
        optimizer.global_step = ckpt.get("step", 1)
    return int(ckpt.get("epoch", 0))
