"""Two fused fwd+bwd steps at the AISHELL shape; the SECOND step's backward (argv[1] == "step": the whole second
step, forward included) is a profiler range (cudaProfilerStart/Stop) for `ncu --replay-mode app-range --metrics
dram__bytes_read.sum,dram__bytes_write.sum`: a range is measured as a whole, so kernels that run side by side on
two streams are measured side by side."""
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
STEP = len(sys.argv) > 1 and sys.argv[1] == "step"
for it in range(2):
    if it == 1 and STEP:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    loss = f(enc, dec, y, tl, ul, host_lens=hl)
    torch.cuda.synchronize()
    if it == 1 and not STEP:
        torch.cuda.profiler.start()
    loss.backward()
    torch.cuda.synchronize()
    if it == 1:
        torch.cuda.profiler.stop()
print("loss", float(loss))
