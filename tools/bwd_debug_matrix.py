"""joint_bwd_ds time (KERNEL=joint_fwd_store MODE=2 in the environment: the storing forward) under RNNT_B200_DEBUG
bits (1 no epilogue math, 4 no MMA issue, 16 no dS / E stores) for environment variants given as "NAME=VAL,..."
arguments ("-" = none): which side bounds a kernel variant."""
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
print("RESULT", round(sum(s[%r])/2,3))
''' % (ROOT, os.environ.get("KERNEL", "joint_bwd_ds"))
for var in (sys.argv[1:] or ["-"]):
    row = []
    for dbg in os.environ.get("DBGS", "0,1,16,4,5").split(","):
        env = dict(os.environ, RNNT_B200_DEBUG=dbg, RNNT_B200_DEBUG_MODE=os.environ.get("MODE", "1"))
        if env["RNNT_B200_DEBUG_MODE"] in ("0", "1"):
            env["RNNT_B200_BWD"] = "recompute"
        if var != "-":
            for kv in var.split(","):
                k, v = kv.split("="); env[k] = v
        p = subprocess.run([sys.executable, "-c", CODE], env=env, capture_output=True, text=True, timeout=300)
        out = [l for l in p.stdout.splitlines() if l.startswith("RESULT")]
        row.append(f"dbg{dbg}=" + (out[-1].split()[1] if out else "FAIL"))
    print(f"[{var}] {os.environ.get('KERNEL', 'joint_bwd_ds')} ms: " + "  ".join(row), flush=True)
