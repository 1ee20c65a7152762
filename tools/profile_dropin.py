"""One drop-in RNNTLoss fwd+bwd on materialised logits (for ncu of the HBM-streaming kernels)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, rnnt_transducer_b200 as R
B, T, U, V = (int(x) for x in sys.argv[1:5])
dev = "cuda:0"
g = torch.Generator().manual_seed(0)
x = torch.randn(B, T, U + 1, V, generator=g).to(dev).requires_grad_(True)
y = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
tl = torch.full((B,), T, dtype=torch.int32, device=dev); ul = torch.full((B,), U, dtype=torch.int32, device=dev)
for _ in range(2):
    x.grad = None
    loss = R.RNNTLoss()(x, y, tl, ul); loss.backward()
torch.cuda.synchronize()
with R.KernelTimer() as kt:
    x.grad = None
    loss = R.RNNTLoss()(x, y, tl, ul); loss.backward()
print("loss", float(loss), {k: [round(v, 4) for v in vs] for k, vs in kt.summary().items()})
