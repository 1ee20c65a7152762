"""Library dh GEMM (dS . W2) at one dS chunk of the AISHELL shape: operand layouts and row counts.
Usage: python tools/dh_gemm_bench.py"""
import torch
dev = "cuda:0"
Vds, H = 4288, 512
def timeit(fn, n=8):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
w = (torch.randn(Vds, H, device=dev) * 0.04).to(torch.bfloat16)      # (V, H): NN
wt = w.t().contiguous()                                               # (H, V): NT through .t()
for tiles in (3157, 3158, 3256, 3108, 3404, 3552):
    rows = tiles * 128
    ds = (torch.randn(rows, Vds, device=dev) * 0.05).to(torch.bfloat16)
    out = torch.empty(rows, H, device=dev, dtype=torch.bfloat16)
    a = timeit(lambda: torch.mm(ds, w, out=out))
    b = timeit(lambda: torch.mm(ds, wt.t(), out=out))
    fl = 2.0 * rows * Vds * H
    print(f"tiles {tiles} rows {rows}: NN {a:.3f} ms ({fl/a*1e-9:.0f} TF/s) | NT {b:.3f} ms ({fl/b*1e-9:.0f} TF/s)", flush=True)
    del ds, out
