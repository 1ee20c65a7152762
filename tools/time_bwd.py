"""Times the fused backward kernels under RNNT_B200_DEBUG variants (bwd kernel only)."""
# NOTE: the RNNT_B200_* kernel knobs and clock64 traces exist only in the tuning build
# (make -C rnnt-transducer_b200/csrc tune); this tool selects it through RNNT_B200_LIB.
import os as _os
_os.environ.setdefault("RNNT_B200_LIB", _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                                                      "rnnt-transducer_b200", "librnnt_b200_tune.so"))
import os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = r'''
import sys, os
sys.path.insert(0, %r)
import torch, rnnt_transducer_b200 as R
B,T,U,E,D,H,V = 32,500,100,320,320,512,4232
dev="cuda:0"; g=torch.Generator().manual_seed(0)
joint=R.JointNet(E+D,H,V).to(dev); f=R.JointRNNTLoss(joint)
enc=torch.randn(B,T,E,generator=g).to(dev).requires_grad_(True); dec=torch.randn(B,U+1,D,generator=g).to(dev).requires_grad_(True)
y=torch.randint(1,V,(B,U),generator=g,dtype=torch.int32).to(dev)
tl=torch.full((B,),T,dtype=torch.int32,device=dev); ul=torch.full((B,),U,dtype=torch.int32,device=dev)
for _ in range(2): f(enc,dec,y,tl,ul).backward()
with R.KernelTimer() as kt:
    for _ in range(2): f(enc,dec,y,tl,ul).backward()
s=kt.summary()
print("RESULT", os.environ.get("RNNT_B200_DEBUG","0"), {k: round(sum(v)/2,3) for k,v in s.items()})
''' % ROOT
for dbg in ["0", "8", "16", "24", "1"]:
    env = dict(os.environ, RNNT_B200_DEBUG=dbg, RNNT_B200_DEBUG_MODE="1")
    p = subprocess.run([sys.executable, "-c", CODE], env=env, capture_output=True, text=True, timeout=300)
    out = [l for l in p.stdout.splitlines() if l.startswith("RESULT")]
    print(out[-1] if out else ("FAIL " + dbg + " " + p.stderr[-400:]), flush=True)
# dW2 GEMM: contiguous vs strided h operand
import torch
dev = "cuda:0"
rows, Vp, H = 3600 * 128, 4352, 512
ds = torch.randn(rows, Vp, device=dev, dtype=torch.bfloat16) * 0.01
hs = torch.randn(rows, H + 8, device=dev, dtype=torch.bfloat16)
hc = hs[:, :H].contiguous()
def timeit(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
a = ds.view(4, rows // 4, Vp).transpose(1, 2)
print("bmm contiguous h", timeit(lambda: torch.bmm(a, hc.view(4, rows // 4, H), out_dtype=torch.float32).sum(0)))
print("bmm strided h   ", timeit(lambda: torch.bmm(a, hs[:, :H].unflatten(0, (4, rows // 4)), out_dtype=torch.float32).sum(0)))
print("mm strided h    ", timeit(lambda: torch.mm(ds.t(), hs[:, :H], out_dtype=torch.float32)))
