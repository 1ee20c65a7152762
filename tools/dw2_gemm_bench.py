"""Which shape should the library GEMM [dW2 | db2] = dS^T . h_ext be given? Times torch.bmm (split-K 4,
as in _api.py) for several widths of h_ext (H + pad columns, pitch = width or wider) and for the
transposed problem. One dS chunk of the AISHELL shape."""
import torch, sys
rows, Vds, H, SK = 3158 * 128, 4288, 512, 4
dev = "cuda:0"
ds = (torch.randn(rows, Vds, device=dev) * 0.01).to(torch.bfloat16)
def t(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for pitch, N in [(512, 512), (520, 512), (520, 520), (528, 528), (544, 544), (576, 576), (640, 640), (576, 520), (640, 520), (576, 528)]:
    h = torch.randn(rows, pitch, device=dev).to(torch.bfloat16)
    a = ds.view(SK, rows // SK, Vds).transpose(1, 2)
    b = h.view(SK, rows // SK, pitch)[:, :, :N]
    ms = t(lambda: torch.bmm(a, b, out_dtype=torch.float32).sum(0))
    ms_t = t(lambda: torch.bmm(b.transpose(1, 2), a.transpose(1, 2), out_dtype=torch.float32).sum(0))
    ms1 = t(lambda: torch.mm(ds.t(), h[:, :N], out_dtype=torch.float32))
    print(f"pitch {pitch} N {N}: dS^T.h {ms:.3f} ms | h^T.dS {ms_t:.3f} ms | no split {ms1:.3f} ms  ({2*rows*Vds*N/ms*1e-9:.0f} TF/s)", flush=True)
    del h
