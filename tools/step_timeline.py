"""Timeline of one fused fwd+bwd step at the AISHELL shape from the KernelTimer's CUDA events: start offset,
duration and the gap to the previous own kernel (torch glue, launch gaps) for every C-ABI call of the step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rnnt_transducer_b200 as R
B, T, U, E, D, H, V = 32, 500, 100, 320, 320, 512, 4232
dev = "cuda:0"
g = torch.Generator().manual_seed(0)
joint = R.JointNet(E + D, H, V).to(dev)
f = R.JointRNNTLoss(joint)
enc = torch.randn(B, T, E, generator=g).to(dev).requires_grad_(True)
dec = torch.randn(B, U + 1, D, generator=g).to(dev).requires_grad_(True)
y = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
tl = torch.full((B,), T, dtype=torch.int32, device=dev)
ul = torch.full((B,), U, dtype=torch.int32, device=dev)
hl = (tl.cpu(), ul.cpu())
def step():
    for p in joint.parameters():
        p.grad = None
    enc.grad = dec.grad = None
    f(enc, dec, y, tl, ul, host_lens=hl).backward()
for _ in range(5):
    step()
torch.cuda.synchronize()
with R.KernelTimer() as kt:
    t0 = torch.cuda.Event(enable_timing=True); t0.record()
    for _ in range(3):
        step()
    t1 = torch.cuda.Event(enable_timing=True); t1.record()
torch.cuda.synchronize()
recs = kt.records
n = len(recs) // 3
last = recs[-n:]                      # the third step
base = recs[-n - 1][2] if len(recs) > n else t0
print(f"three steps: {t0.elapsed_time(t1) / 3:.3f} ms per step; the third step:")
prev_end = base
tot_k = tot_gap = 0.0
for name, e0, e1 in last:
    gap = prev_end.elapsed_time(e0)
    dur = e0.elapsed_time(e1)
    print(f"  +{base.elapsed_time(e0):8.3f} ms  {name:18s} {dur:7.3f} ms   gap before {gap:6.3f}")
    if gap > 0:
        tot_gap += gap
    tot_k += dur
    prev_end = e1 if prev_end.elapsed_time(e1) > 0 else prev_end
print(f"  kernels {tot_k:.3f} ms, gaps {tot_gap:.3f} ms")
