"""A/B timing of one full fused step (per-kernel CUDA-event times) under environment variants,
interleaved in ONE process group on ONE box (boxes differ by several % under the power cap).
Usage: python tools/ab_step.py "NAME=VAL,NAME2=VAL2" "NAME=VAL3" ...   ("-" = no overrides)"""
import os, sys, subprocess, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = r'''
import sys, os, json
sys.path.insert(0, %r)
import torch, rnnt_transducer_b200 as R
B,T,U,E,D,H,V = 32,500,100,320,320,512,4232
dev="cuda:0"; g=torch.Generator().manual_seed(0)
joint=R.JointNet(E+D,H,V).to(dev); f=R.JointRNNTLoss(joint)
enc=torch.randn(B,T,E,generator=g).to(dev).requires_grad_(True); dec=torch.randn(B,U+1,D,generator=g).to(dev).requires_grad_(True)
y=torch.randint(1,V,(B,U),generator=g,dtype=torch.int32).to(dev)
tl=torch.full((B,),T,dtype=torch.int32,device=dev); ul=torch.full((B,),U,dtype=torch.int32,device=dev)
hl=(tl.cpu(),ul.cpu())
for _ in range(3): f(enc,dec,y,tl,ul,host_lens=hl).backward()
torch.cuda.synchronize()
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
N=8
e0.record()
for _ in range(N): f(enc,dec,y,tl,ul,host_lens=hl).backward()
e1.record(); torch.cuda.synchronize()
step=e0.elapsed_time(e1)/N
with R.KernelTimer() as kt:
    for _ in range(3): f(enc,dec,y,tl,ul,host_lens=hl).backward()
s=kt.summary()
print("RESULT", json.dumps({"step_ms": round(step,3), **{k: round(sum(v)/3,3) for k,v in s.items()}}))
''' % ROOT
variants = sys.argv[1:] or ["-"]
for rnd in range(2):
    for var in variants:
        env = dict(os.environ)
        if var != "-":
            for kv in var.split(","):
                k, v = kv.split("=")
                env[k] = v
        p = subprocess.run([sys.executable, "-c", CODE], env=env, capture_output=True, text=True, timeout=600)
        out = [l for l in p.stdout.splitlines() if l.startswith("RESULT")]
        print(f"[{var}] " + (out[-1] if out else "FAIL " + p.stderr[-600:]), flush=True)
