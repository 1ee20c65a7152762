"""Multi-rank host logic on CPU (gloo, world_size 2): utterance sharding, the global-mean loss
and the flat gradient+loss all-reduce reproduce the single-process result. The per-rank loss here
is the float64 oracle (the CUDA path needs a GPU); what is under test is the sharding/collective
logic around it."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import rnnt_transducer_b200 as R
from oracle import rnnt_oracle as O
from tests.helpers import synth_labels, synth_lengths


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _inputs():
    gen = torch.Generator().manual_seed(11)
    B, T, U, E, D, H, V = 6, 9, 5, 8, 8, 16, 13
    tl, ul = synth_lengths(B, T, U, "ragged", gen)
    y = synth_labels(B, U, V, ul, gen)
    enc = torch.randn(B, T, E, generator=gen, dtype=torch.float64)
    dec = torch.randn(B, U + 1, D, generator=gen, dtype=torch.float64)
    w1 = torch.randn(H, E + D, generator=gen, dtype=torch.float64) * 0.3
    b1 = torch.randn(H, generator=gen, dtype=torch.float64) * 0.1
    w2 = torch.randn(V, H, generator=gen, dtype=torch.float64) * 0.3
    b2 = torch.randn(V, generator=gen, dtype=torch.float64) * 0.1
    return enc, dec, w1, b1, w2, b2, y, tl, ul


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    enc, dec, w1, b1, w2, b2, y, tl, ul = _inputs()
    params = [t.clone().requires_grad_(True) for t in (w1, b1, w2, b2)]
    mine = R.shard_batch(tl.tolist(), ul.tolist(), world)[rank]
    idx = torch.tensor(mine)
    loss_b = O.joint_rnnt_loss(enc[idx], dec[idx], *params, y[idx], tl[idx], ul[idx], reduction="none")
    loss = R.global_mean_loss(loss_b, global_batch=enc.shape[0])
    loss.backward()
    total = R.allreduce_grads_and_loss(params, loss_b.sum())
    if rank == 0:
        torch.save({"loss": loss.detach(), "total": total, "grads": [p.grad for p in params]}, out)
    dist.destroy_process_group()


def _worker_flat(rank, world, port, out):
    """bench.py's step: gradients are views of one flat buffer whose last element carries the loss sum."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    enc, dec, w1, b1, w2, b2, y, tl, ul = _inputs()
    params = [t.clone().requires_grad_(True) for t in (w1, b1, w2, b2)]
    flat = R.bind_flat_grads(params, extra=1)
    idx = torch.tensor(R.shard_batch(tl.tolist(), ul.tolist(), world)[rank])
    for _ in range(2):                      # two steps: the buffer is zeroed, not re-bound
        flat.zero_()
        loss_b = O.joint_rnnt_loss(enc[idx], dec[idx], *params, y[idx], tl[idx], ul[idx], reduction="none")
        (loss_b.sum() / enc.shape[0]).backward()
        flat[-1] = loss_b.sum().detach()
        dist.all_reduce(flat)
    assert all(p.grad.data_ptr() >= flat.data_ptr() for p in params)   # still views of the flat buffer
    if rank == 0:
        torch.save({"total": flat[-1].clone(), "grads": [p.grad.clone() for p in params]}, out)
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_rank_flat_gradient_buffer_matches_single_process(tmp_path):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker_flat, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    enc, dec, w1, b1, w2, b2, y, tl, ul = _inputs()
    params = [t.clone().requires_grad_(True) for t in (w1, b1, w2, b2)]
    ref = O.joint_rnnt_loss(enc, dec, *params, y, tl, ul, reduction="mean")
    ref.backward()
    assert abs(float(got["total"]) - float(ref) * enc.shape[0]) < 1e-10
    for g, p in zip(got["grads"], params):
        assert torch.allclose(g, p.grad, atol=1e-12)


def test_shard_batch_is_a_balanced_partition():
    tl = [50, 10, 40, 30, 20, 60, 5]
    ul = [5, 1, 4, 3, 2, 6, 0]
    for world in (1, 2, 3, 4):
        shards = R.shard_batch(tl, ul, world)
        assert sorted(i for s in shards for i in s) == list(range(len(tl)))
        assert max(len(s) for s in shards) - min(len(s) for s in shards) <= 1
    s2 = R.shard_batch(tl, ul, 2)
    cells = [t * (u + 1) for t, u in zip(tl, ul)]
    load = [sum(cells[i] for i in s) for s in s2]
    assert abs(load[0] - load[1]) <= max(cells)


@pytest.mark.timeout(120)
def test_speed_weighted_shards_follow_the_ranks_rates():
    """shard_batch_weighted: a partition of the batch, more cells to faster ranks, the equal split when the rates
    are close, never an empty rank, the same answer from the same rates (every rank computes it by itself)."""
    g = torch.Generator().manual_seed(3)
    tl, ul = synth_lengths(64, 500, 100, "ragged", g)
    tl, ul = tl.tolist(), ul.tolist()
    cells = [t * (u + 1) for t, u in zip(tl, ul)]
    rates = [1.0, 0.9, 1.05, 0.8]
    sh = R.shard_batch_weighted(tl, ul, rates)
    assert sorted(i for s_ in sh for i in s_) == list(range(64))
    load = [sum(cells[i] for i in s_) for s_ in sh]
    finish = [l / r for l, r in zip(load, rates)]
    assert max(finish) / min(finish) < 1.06                       # finishing times level, not the loads
    assert load[3] < load[1] < load[0] < load[2]
    assert sh == R.shard_batch_weighted(tl, ul, list(rates))
    equal = R.shard_batch(tl, ul, 4)
    assert R.shard_batch_weighted(tl, ul, [1.0, 0.99, 1.0, 0.995]) == equal      # inside the 1.5 % dead band
    assert R.shard_batch_weighted(tl, ul, [1.0, 0.0, 1.0, 1.0]) == equal          # a nonsense rate: equal split
    assert R.shard_batch_weighted(tl[:2], ul[:2], [1.0, 1e-6]) == R.shard_batch(tl[:2], ul[:2], 2)   # no empty rank
    # identical utterances (bench.py's default batch): counts follow the rates to within one utterance
    sh = R.shard_batch_weighted([500] * 256, [100] * 256, [1, 1, .92, 1, .97, 1, 1.02, .95])
    assert [len(s_) for s_ in sh] == [33, 33, 30, 32, 32, 32, 33, 31]


def test_two_rank_gloo_matches_single_process(tmp_path):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    enc, dec, w1, b1, w2, b2, y, tl, ul = _inputs()
    params = [t.clone().requires_grad_(True) for t in (w1, b1, w2, b2)]
    ref = O.joint_rnnt_loss(enc, dec, *params, y, tl, ul, reduction="mean")
    ref.backward()
    assert abs(float(got["loss"]) - float(ref)) < 1e-12
    assert abs(float(got["total"]) - float(ref) * enc.shape[0]) < 1e-10
    for g, p in zip(got["grads"], params):
        assert torch.allclose(g, p.grad, atol=1e-12)
