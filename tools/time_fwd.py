"""Times the fused forward kernel alone (no grad) under RNNT_B200_DEBUG variants."""
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
joint=R.JointNet(E+D,H,V).to(dev); f=R.JointRNNTLoss(joint, reduction="none")
enc=torch.randn(B,T,E,generator=g).to(dev); dec=torch.randn(B,U+1,D,generator=g).to(dev)
y=torch.randint(1,V,(B,U),generator=g,dtype=torch.int32).to(dev)
tl=torch.full((B,),T,dtype=torch.int32,device=dev); ul=torch.full((B,),U,dtype=torch.int32,device=dev)
with torch.no_grad():
    for _ in range(2): f(enc,dec,y,tl,ul)
    with R.KernelTimer() as kt:
        for _ in range(3): f(enc,dec,y,tl,ul)
    s=kt.summary()
print("RESULT", os.environ.get("RNNT_B200_DEBUG","0"), os.environ.get("RNNT_B200_STAGES","-"), os.environ.get("RNNT_B200_GRID","-"), ["%%.3f"%%x for x in s["joint_fwd"]], "lattice", ["%%.3f"%%x for x in s["lattice"]])
''' % ROOT
for dbg, stg, grd in [("0","","148"),("1","","148"),("4","","148"),("5","","148")]:
    env = dict(os.environ, RNNT_B200_DEBUG=dbg, RNNT_B200_GRID=grd)
    if stg: env["RNNT_B200_STAGES"] = stg
    p = subprocess.run([sys.executable, "-c", CODE], env=env, capture_output=True, text=True, timeout=300)
    out = [l for l in p.stdout.splitlines() if l.startswith("RESULT")]
    print(out[-1] if out else ("FAIL " + dbg + " " + p.stderr[-400:]), flush=True)
