"""One dS chunk of the AISHELL shape through the dh GEMM of the TUNING build with RNNT_B200_DH_TRACE; summarises the
clock64 stamps of CTA 0 (MMA thread per K item, dS producer per K item, epilogue warp 0 per unit).
RNNT_B200_DH_DEBUG (see csrc/dh_gemm.cu) selects knock-out variants."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["RNNT_B200_LIB"] = os.path.join(ROOT, "rnnt-transducer_b200", "librnnt_b200_tune.so")
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
os.environ["RNNT_B200_DH_TRACE"] = os.path.join(ROOT, "gpurun_out", "trace_dh.bin")
import numpy as np, torch
from rnnt_transducer_b200 import _lib
L = _lib.lib()
V, H, tiles = 4232, 512, 3256
Vds, Vpad = L.rnnt_b200_ds_pitch(V), L.rnnt_b200_vocab_pad(V)
dev = "cuda:0"; rows = tiles * 128
ds = (torch.randn(rows, Vds, device=dev) * 0.05).to(torch.bfloat16)
w2t = (torch.randn(H, Vpad, device=dev) * 0.05).to(torch.bfloat16)
h = torch.tanh(torch.randn(rows, H, device=dev)).to(torch.bfloat16)
dpre = torch.empty(rows, H, dtype=torch.bfloat16, device=dev)
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    assert L.rnnt_b200_dh_gemm(ds.data_ptr(), w2t.data_ptr(), h.data_ptr(), V, H, tiles, 0, tiles, 0, dpre.data_ptr(), st) == 0
torch.cuda.synchronize()
a = np.fromfile(os.environ["RNNT_B200_DH_TRACE"], dtype=np.int64).reshape(3, 4096, 3)
NK = Vds // 64
for name, t in zip(["MMA thread per K item (wait full | issue)", "dS producer per K item (acquire | issue)"], a[:2]):
    idx = np.nonzero(t[:, 0] > 0)[0]
    if len(idx) < 10:
        print(name, "no data"); continue
    t = t[idx]
    wait, work, per = t[:, 1] - t[:, 0], t[:, 2] - t[:, 1], np.diff(t[:, 0]) / np.maximum(np.diff(idx), 1)
    print(f"{name}: n={len(t)} period med {np.median(per):.0f} mean {per.mean():.0f} | wait med {np.median(wait):.0f} "
          f"mean {wait.mean():.0f} p90 {np.percentile(wait, 90):.0f} | work med {np.median(work):.0f} mean {work.mean():.0f}"
          f" | ideal MMA per K item 1024")
e = a[2]; e = e[e[:, 0] > 0]
print(f"epilogue warp 0 per unit: n={len(e)} wait acc_full mean {int((e[:,1]-e[:,0]).mean())} | epilogue work mean "
      f"{int((e[:,2]-e[:,1]).mean())} max {int((e[:,2]-e[:,1]).max())} | ideal MMA per unit {NK * 1024}")
