import torch
dev = "cuda:0"
rows, Vp = 3160 * 128, 4352
ds = torch.randn(rows, Vp, device=dev, dtype=torch.bfloat16) * 0.01
def timeit(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for N in (512, 520, 528, 544, 576):
    h = torch.randn(rows, N, device=dev, dtype=torch.bfloat16)
    fl = 2.0 * rows * Vp * 512
    for S in (2, 4, 8, 10):
        if rows % S: continue
        a = ds.view(S, rows // S, Vp).transpose(1, 2)
        b = h.view(S, rows // S, N)
        t = timeit(lambda: torch.bmm(a, b, out_dtype=torch.float32).sum(0))
        print(f"N={N} split-K S={S}: {t:.3f} ms  useful {fl/t/1e9:.0f} TF/s")
    # transposed formulation: [N x rows] @ [rows x Vp]
    for S in (4, 8):
        a = h.view(S, rows // S, N).transpose(1, 2)
        b = ds.view(S, rows // S, Vp)
        t = timeit(lambda: torch.bmm(a, b, out_dtype=torch.float32).sum(0))
        print(f"N={N} transposed split-K S={S}: {t:.3f} ms  useful {fl/t/1e9:.0f} TF/s")
