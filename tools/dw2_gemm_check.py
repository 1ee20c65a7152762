"""rnnt_b200_dw2_gemm against torch (fp32 reference of the same bf16 operands) + timing at one
chunk of the AISHELL shape. Usage: python tools/dw2_gemm_check.py [small|full]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rnnt_transducer_b200 import _lib
L = _lib.lib()
dev = "cuda:0"

def run(rows, V, H, seed=0, time_it=False):
    Vds, hp = L.rnnt_b200_ds_pitch(V), L.rnnt_b200_h_pitch(H)
    g = torch.Generator(device=dev).manual_seed(seed)
    ds = (torch.randn(rows, Vds, device=dev, generator=g) * 0.05).to(torch.bfloat16)
    h = torch.tanh(torch.randn(rows, hp, device=dev, generator=g)).to(torch.bfloat16)
    dw2 = torch.zeros(Vds, H, device=dev); db2 = torch.zeros(Vds, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    rc = L.rnnt_b200_dw2_gemm(ds.data_ptr(), h.data_ptr(), rows, V, H, dw2.data_ptr(), db2.data_ptr(), st)
    torch.cuda.synchronize()
    assert rc == 0, rc
    n = min(rows, 8192)  # reference on a row subset would not match: use all rows but in fp32 chunks
    ref = torch.zeros(Vds, H, device=dev); refb = torch.zeros(Vds, device=dev)
    for lo in range(0, rows, 65536):
        a = ds[lo:lo + 65536].float(); b = h[lo:lo + 65536, :H].float()
        ref += a.t() @ b; refb += a.sum(0)
    e1 = float((dw2 - ref).norm() / ref.norm()); e2 = float((db2 - refb).norm() / refb.norm())
    msg = f"rows {rows} V {V} (Vds {Vds}) H {H}: dW2 rel-L2 {e1:.3e}, db2 rel-L2 {e2:.3e}"
    if time_it:
        e0_, e1_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(2): L.rnnt_b200_dw2_gemm(ds.data_ptr(), h.data_ptr(), rows, V, H, dw2.data_ptr(), db2.data_ptr(), st)
        e0_.record()
        for _ in range(5): L.rnnt_b200_dw2_gemm(ds.data_ptr(), h.data_ptr(), rows, V, H, dw2.data_ptr(), db2.data_ptr(), st)
        e1_.record(); torch.cuda.synchronize()
        ms = e0_.elapsed_time(e1_) / 5
        SK = 4
        def lib():
            return torch.bmm(h.view(SK, rows // SK, hp).transpose(1, 2), ds.view(SK, rows // SK, Vds), out_dtype=torch.float32).sum(0)
        for _ in range(2): lib()
        e0_.record()
        for _ in range(5): lib()
        e1_.record(); torch.cuda.synchronize()
        msg += f" | {ms:.3f} ms ({2*rows*Vds*H/ms*1e-9:.0f} TF/s) vs library h^T.dS at N={hp}: {e0_.elapsed_time(e1_)/5:.3f} ms"
    print(msg, flush=True)
    return e1, e2

mode = sys.argv[1] if len(sys.argv) > 1 else "small"
run(64, 200, 256)
run(128, 256, 512)
run(1024, 500, 512)
run(4096, 4232, 512)
if mode == "full":
    run(3158 * 128, 4232, 512, time_it=True)
