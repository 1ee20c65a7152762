"""One fused fwd+bwd at a given shape (for `ncu -k regex:joint_kernel`). Usage:
   python tools/profile_joint.py B T U [H V]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rnnt_transducer_b200 as R

B, T, U = (int(x) for x in sys.argv[1:4])
H = int(sys.argv[4]) if len(sys.argv) > 4 else 512
V = int(sys.argv[5]) if len(sys.argv) > 5 else 4232
E = D = 320 if H == 512 else H
dev = "cuda:0"
g = torch.Generator().manual_seed(0)
joint = R.JointNet(E + D, H, V).to(dev)
f = R.JointRNNTLoss(joint)
enc = torch.randn(B, T, E, generator=g).to(dev).requires_grad_(True)
dec = torch.randn(B, U + 1, D, generator=g).to(dev).requires_grad_(True)
y = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
tl = torch.full((B,), T, dtype=torch.int32, device=dev)
ul = torch.full((B,), U, dtype=torch.int32, device=dev)
for _ in range(2):
    loss = f(enc, dec, y, tl, ul)
    loss.backward()
torch.cuda.synchronize()
print("loss", float(loss))
