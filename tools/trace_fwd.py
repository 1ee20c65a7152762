"""Runs one fused forward (or, with argv[1] == "bwd", forward+backward: the LAST dS chunk is traced) at
cfg3 with RNNT_B200_TRACE and summarises the per-stage clock stamps."""
# NOTE: the RNNT_B200_* kernel knobs and clock64 traces exist only in the tuning build
# (make -C rnnt-transducer_b200/csrc tune); this tool selects it through RNNT_B200_LIB.
import os as _os
_os.environ.setdefault("RNNT_B200_LIB", _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                                                      "rnnt-transducer_b200", "librnnt_b200_tune.so"))
import os, sys, struct
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["RNNT_B200_TRACE"] = "gpurun_out/trace_fwd.bin"
BWD = len(sys.argv) > 1 and sys.argv[1] == "bwd"
STORE = len(sys.argv) > 1 and sys.argv[1] == "store"     # the forward that stores its exponentials (MODE 2)
os.environ["RNNT_B200_DEBUG_MODE"] = "1" if BWD else "2" if STORE else "0"
if BWD:
    os.environ["RNNT_B200_BWD"] = "recompute"
import numpy as np, torch, rnnt_transducer_b200 as R
B,T,U,E,D,H,V = 32,500,100,320,320,512,4232
dev="cuda:0"; g=torch.Generator().manual_seed(0)
joint=R.JointNet(E+D,H,V).to(dev); f=R.JointRNNTLoss(joint, reduction="none")
enc=torch.randn(B,T,E,generator=g).to(dev); dec=torch.randn(B,U+1,D,generator=g).to(dev)
y=torch.randint(1,V,(B,U),generator=g,dtype=torch.int32).to(dev)
tl=torch.full((B,),T,dtype=torch.int32,device=dev); ul=torch.full((B,),U,dtype=torch.int32,device=dev)
if STORE:
    enc.requires_grad_(True); dec.requires_grad_(True)
    for _ in range(2): f(enc,dec,y,tl,ul)
elif BWD:
    enc.requires_grad_(True); dec.requires_grad_(True)
    for _ in range(2): f(enc,dec,y,tl,ul).sum().backward()
else:
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
    if name.startswith("epi"):
        # per position in the lattice tile's sweep (17 vocabulary tiles at cfg3): where do the slow ones sit?
        NVT = (V + 255) // 256
        full = t
        per, wrk = np.diff(full[:, 0]), (full[:, 2] - full[:, 1])
        print("   by sweep position (period/work):", " ".join(
            f"{k}:{int(per[np.arange(k, len(per), NVT)].mean())}/{int(wrk[np.arange(k, len(per), NVT)].mean())}" for k in range(NVT)))
    elif name.startswith("MMA"):
        P = 4 * ((V + 255) // 256)
        per = np.diff(t[:, 0])
        print("   MMA period by (vocabulary tile, K group) position, summed per vocabulary tile:",
              [int(sum(per[np.arange(k + j, len(per), P)].mean() for j in range(4))) for k in range(0, P, 4)])
