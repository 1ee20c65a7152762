"""hidden_kernel of two builds of the library on the same inputs: the h stashes must be bit-identical (used when the
kernel's arithmetic is re-expressed, e.g. on the packed f32x2 pipe). Usage: check_hidden_bits.py other_lib.so"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rnnt_transducer_b200 import _lib
A = _lib.lib()
Bl = ctypes.CDLL(sys.argv[1])
for name, (res, args) in _lib.SIGNATURES.items():
    if hasattr(Bl, name):
        getattr(Bl, name).restype, getattr(Bl, name).argtypes = res, args
dev = "cuda:0"
B, T, U1, H = 6, 83, 41, 512
g = torch.Generator().manual_seed(1)
pe = (torch.randn(B, T, H, generator=g) * 3).to(dev)
pd = (torch.randn(B, U1, H, generator=g) * 3).to(dev)
pe[0, 0, :8] = torch.tensor([50., -50., 10., 9.99, -10., 0., 1e-8, 20.])
tl = torch.tensor([83, 1, 40, 83, 7, 60], dtype=torch.int32, device=dev)
ul = torch.tensor([40, 0, 13, 40, 40, 5], dtype=torch.int32, device=dev)
out = []
for L in (A, Bl):
    cap = L.rnnt_b200_max_tiles(B, T, U1)
    tu = torch.empty(cap, dtype=torch.int32, device=dev); tc = torch.empty(cap, dtype=torch.int32, device=dev)
    nt = torch.empty(2, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    assert L.rnnt_b200_build_tiles(tl.data_ptr(), ul.data_ptr(), B, T, U1, cap, tu.data_ptr(), tc.data_ptr(), nt.data_ptr(), st) == 0
    h = torch.zeros(cap * 128, L.rnnt_b200_h_pitch(H), dtype=torch.bfloat16, device=dev)
    assert L.rnnt_b200_joint_hidden(pe.data_ptr(), pd.data_ptr(), tl.data_ptr(), ul.data_ptr(), tu.data_ptr(), tc.data_ptr(),
                                    nt.data_ptr(), B, T, U1, H, cap, h.data_ptr(), st) == 0
    torch.cuda.synchronize()
    out.append(h)
same = torch.equal(out[0].view(torch.int16), out[1].view(torch.int16))
print("hidden stash bit-identical:", same, "elements", out[0].numel(), "nonzero", int((out[0] != 0).sum()))
sys.exit(0 if same else 1)
