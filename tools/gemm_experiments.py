"""cuBLAS shape experiments for the two library GEMMs of the backward (dh, dW2)."""
import torch, time
dev = "cuda:0"
rows, Vp, H = 480 * 128 * 7, 4352, 512   # ~ one 4 GB chunk of cfg3
rows = 3600 * 128
ds = torch.randn(rows, Vp, device=dev, dtype=torch.bfloat16) * 0.01
def timeit(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for N in (512, 520, 576, 640):
    h = torch.randn(rows, N, device=dev, dtype=torch.bfloat16)
    fl = 2.0 * rows * Vp * N
    t = timeit(lambda: torch.mm(ds.t(), h, out_dtype=torch.float32))
    print(f"dw2 = dS^T[{Vp}x{rows}] @ h[{rows}x{N}] f32out: {t:.3f} ms {fl/t/1e9:.0f} TF/s")
    t = timeit(lambda: torch.mm(h.t(), ds, out_dtype=torch.float32))
    print(f"dw2T = h^T[{N}x{rows}] @ dS f32out: {t:.3f} ms {fl/t/1e9:.0f} TF/s")
    t = timeit(lambda: torch.mm(ds.t(), h))
    print(f"dw2 bf16 out: {t:.3f} ms {fl/t/1e9:.0f} TF/s")
    for S in (4, 16, 36):
        if rows % S: continue
        a = ds.view(S, rows // S, Vp).transpose(1, 2)
        b = h.view(S, rows // S, N)
        t = timeit(lambda: torch.bmm(a, b, out_dtype=torch.float32).sum(0))
        print(f"  split-K bmm S={S}: {t:.3f} ms {fl/t/1e9:.0f} TF/s")
w = torch.randn(Vp, H, device=dev, dtype=torch.bfloat16)
t = timeit(lambda: torch.mm(ds, w, out_dtype=torch.float32))
print(f"dh f32 out: {t:.3f} ms {2.0*rows*Vp*H/t/1e9:.0f} TF/s")
t = timeit(lambda: torch.mm(ds, w))
print(f"dh bf16 out: {t:.3f} ms {2.0*rows*Vp*H/t/1e9:.0f} TF/s")
