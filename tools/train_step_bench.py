"""BASELINE.json configs[3]: full AISHELL training step (stock 4x320 biLSTM encoder, 1x512 LSTM
prediction network, fused joint + loss) batch-sharded by utterance over N GPUs with DDP/NCCL gradient
all-reduce and an SGD(momentum) update (config/aishell.yaml:21-39,59-66). One process per GPU:
  python tools/train_step_bench.py                                   # 1 GPU
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/train_step_bench.py
Synthetic features/targets of the configured maximum lengths (T=500 frames, U=50 labels)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import rnnt_transducer_b200 as R

CFG = dict(enc=dict(type="lstm", hidden_size=320, output_size=320, n_layers=4, bidirectional=True),
           dec=dict(type="lstm", hidden_size=512, output_size=320, n_layers=1),
           joint=dict(input_size=640, inner_size=512), vocab_size=4232, share_weight=False,
           feature_dim=160, dropout=0.3)
B, T, U = int(os.environ.get("BENCH_B", 32)), 500, 50
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
torch.manual_seed(2019)
model = R.Transducer(CFG).to(dev).train()
ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
LOOP = os.environ.get("LOOP", "ours")  # "reference": the per-step host work of train.py:26-50 as written there
if LOOP == "reference":
    opt = torch.optim.SGD(model.parameters(), lr=1e-4, momentum=0.9, foreach=False)
else:
    opt = R.Optimizer(model.parameters(), dict(type="sgd", lr=1e-4, momentum=0.9)).optimizer  # fused update
g = torch.Generator().manual_seed(rank)
x = torch.randn(B, T, CFG["feature_dim"], generator=g).to(dev)
xl = torch.full((B,), T, dtype=torch.long)
y = torch.randint(1, CFG["vocab_size"], (B, U), generator=g).to(dev)
yl = torch.full((B,), U, dtype=torch.long)
xp, yp = torch.zeros(B, T + 100, CFG["feature_dim"]).pin_memory(), torch.zeros(B, U + 10, dtype=torch.long).pin_memory()
xp[:, :T], yp[:, :U] = x.cpu(), y.cpu()
total = torch.zeros((), device=dev)

def step():
    global total
    if LOOP == "reference":  # train.py:26-33: everything to the device, two read-backs, trim
        xi, xli, yi, yli = xp.cuda(), xl.cuda(), yp.cuda(), yl.cuda()
        xi, yi = xi[:, :xli.max().item(), :], yi[:, :yli.max().item()]
    else:                    # lengths stay on the host; padded batch from pinned memory, trimmed there
        xli, yli = xl, yl
        xi, yi = R.trim_batch(xp.to(dev, non_blocking=True), xl, yp.to(dev, non_blocking=True), yl)
    opt.zero_grad(set_to_none=True)
    loss = ddp(xi, xli, yi, yli)
    loss.backward()
    if LOOP == "reference":
        total_host = loss.item()                                 # train.py:47
        torch.nn.utils.clip_grad_norm_(model.parameters(), 200.0, foreach=False)
    else:
        total += loss.detach().sum()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 200.0, foreach=True)   # train.py:49-50, max_grad_norm 200
    opt.step()
    return loss

for _ in range(3):
    step()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
n = 10
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n):
    loss = step()
e1.record()
torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    cells = world * B * T * (U + 1)
    print(json.dumps({"workload": f"aishell_train_step_B{B}_T{T}_U{U}_per_gpu", "n_gpus": world, "loop": LOOP,
                      "ms_per_step": round(float(ms), 3), "cells_per_s": round(cells / float(ms) * 1e3),
                      "utterances_per_s": round(world * B / float(ms) * 1e3, 1), "loss": float(loss)}))
if world > 1:
    dist.destroy_process_group()
