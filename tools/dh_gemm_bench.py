"""One dS chunk of the AISHELL shape through the own dh GEMM (csrc/dh_gemm.cu: dpre = (dS.W2)*(1-h^2), bf16 out)
against the library GEMM it replaced (torch.mm, bf16 out, no epilogue). Usage: python tools/dh_gemm_bench.py [tiles]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rnnt_transducer_b200 import _lib
L = _lib.lib()
V, H = 4232, 512
Vds, Vpad = L.rnnt_b200_ds_pitch(V), L.rnnt_b200_vocab_pad(V)
dev = "cuda:0"
st = torch.cuda.current_stream().cuda_stream
for tiles in [int(a) for a in sys.argv[1:]] or [3256, 3108, 2960]:
    rows = tiles * 128
    g = torch.Generator(device=dev).manual_seed(0)
    ds = (torch.randn(rows, Vds, device=dev, generator=g) * 0.05).to(torch.bfloat16)
    w2 = torch.zeros(Vpad, H, dtype=torch.bfloat16, device=dev)
    w2[:V] = (torch.randn(V, H, device=dev, generator=g) * 0.05).to(torch.bfloat16)
    w2t = w2.t().contiguous()
    h = torch.tanh(torch.randn(rows, H, device=dev, generator=g)).to(torch.bfloat16)
    dpre = torch.empty(rows, H, dtype=torch.bfloat16, device=dev)

    def own():
        rc = L.rnnt_b200_dh_gemm(ds.data_ptr(), w2t.data_ptr(), h.data_ptr(), V, H, tiles, 0, tiles, 0, dpre.data_ptr(), st)
        assert rc == 0, rc

    def lib():
        return torch.mm(ds, w2[:Vds])

    res = {}
    for name, fn in (("own", own), ("lib", lib), ("own", own), ("lib", lib)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        res.setdefault(name, []).append(e0.elapsed_time(e1) / 10)
    flop = 2.0 * rows * Vds * H
    ref = (lib().float() * (1 - h.float() ** 2))
    err = float((dpre.float() - ref).norm() / ref.norm())
    print(f"tiles {tiles}: own {res['own']} ms = {flop / min(res['own']) / 1e9:.0f} TF/s | lib {res['lib']} ms = "
          f"{flop / min(res['lib']) / 1e9:.0f} TF/s | rel-L2 own vs lib(bf16 dh)*(1-h^2) {err:.2e}", flush=True)
