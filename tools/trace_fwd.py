"""Runs one fused forward at cfg3 with RNNT_B200_TRACE and summarises the per-stage clock stamps."""
import os, sys, struct
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["RNNT_B200_TRACE"] = "gpurun_out/trace_fwd.bin"
os.environ["RNNT_B200_DEBUG_MODE"] = "0"
import numpy as np, torch, rnnt_transducer_b200 as R
B,T,U,E,D,H,V = 32,500,100,320,320,512,4232
dev="cuda:0"; g=torch.Generator().manual_seed(0)
joint=R.JointNet(E+D,H,V).to(dev); f=R.JointRNNTLoss(joint, reduction="none")
enc=torch.randn(B,T,E,generator=g).to(dev); dec=torch.randn(B,U+1,D,generator=g).to(dev)
y=torch.randint(1,V,(B,U),generator=g,dtype=torch.int32).to(dev)
tl=torch.full((B,),T,dtype=torch.int32,device=dev); ul=torch.full((B,),U,dtype=torch.int32,device=dev)
with torch.no_grad():
    for _ in range(2): f(enc,dec,y,tl,ul)
torch.cuda.synchronize()
a = np.fromfile("gpurun_out/trace_fwd.bin", dtype=np.int64).reshape(3, 4096, 3)
for name, t in zip(["MMA thread (wait b_full | issue+commit)", "TMA thread (wait b_empty | issue)", "epilogue w0 (wait acc_full | compute)"], a):
    t = t[t[:, 0] > 0]
    n = len(t)
    if n < 10: print(name, "no data"); continue
    lo, hi = n // 4, n // 4 + 1024 if name.startswith("epi") is False else n
    s = t[lo:min(hi, n)]
    wait = (s[:, 1] - s[:, 0]); work = (s[:, 2] - s[:, 1]); period = np.diff(s[:, 0])
    print(f"{name}: n={n} period mean {period.mean():.0f} med {np.median(period):.0f} | wait mean {wait.mean():.0f} med {np.median(wait):.0f} p90 {np.percentile(wait,90):.0f} | work mean {work.mean():.0f} med {np.median(work):.0f}")
    print("   first 24 periods:", period[:24].tolist())
    print("   first 24 waits  :", wait[:24].tolist())
