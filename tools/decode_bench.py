"""Greedy search: batched on-device `greedy_decode` against the per-utterance loop of
Transducer.recognize (oracle/greedy_oracle.py restates rnnt/model.py:74-111), AISHELL joint dims.
Usage: python tools/decode_bench.py [B T]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rnnt_transducer_b200 as R
from oracle.greedy_oracle import greedy_search
from tests.helpers import make_decoder

B, T = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (32, 500)
E = D = 320; H = 512; V = 4232
dev = "cuda:0"
torch.manual_seed(0)
joint = R.JointNet(E + D, H, V).to(dev)
dec = make_decoder(512, V, D, 1, 0).to(dev).eval()
enc = torch.randn(B, T, E, device=dev) * 3
lens = torch.full((B,), T)
with torch.no_grad():
    joint.project_layer.weight.mul_(12); joint.project_layer.bias.mul_(3)
    d0, _ = dec(torch.zeros(1, 1, dtype=torch.long, device=dev))
    lg = joint.project_layer(torch.tanh(joint.forward_layer(
        torch.cat((enc.reshape(-1, E), d0.view(1, -1).expand(B * T, -1)), dim=-1))))
    joint.project_layer.bias[0] += (lg[:, 1:].max(dim=1).values - lg[:, 0]).median()
for _ in range(2):
    tokens, counts = R.greedy_decode(enc, lens, dec, joint)
torch.cuda.synchronize()
t0 = time.perf_counter()
tokens, counts = R.greedy_decode(enc, lens, dec, joint)
torch.cuda.synchronize()
t_ours = time.perf_counter() - t0
gd = R.GreedyDecoder(dec, joint)
for _ in range(2):
    tg, cg = gd(enc, lens)
torch.cuda.synchronize()
t0 = time.perf_counter()
tg, cg = gd(enc, lens)
torch.cuda.synchronize()
t_graph = time.perf_counter() - t0
print(f"CUDA-graph replay {t_graph*1e3:.1f} ms ({B*T/t_graph:.0f} frames/s); same result: "
      f"{bool((tg == tokens).all()) and bool((cg == counts).all())}")
nref = min(B, 4)
t0 = time.perf_counter()
want = greedy_search(enc[:nref], lens[:nref], dec, joint)
torch.cuda.synchronize()
t_ref = (time.perf_counter() - t0) * B / nref
got = [tokens[i, :int(counts[i])].tolist() for i in range(nref)]
print(f"B={B} T={T}: batched {t_ours*1e3:.1f} ms ({B*T/t_ours:.0f} frames/s), per-utterance loop "
      f"{t_ref*1e3:.0f} ms extrapolated from {nref} utterances ({B*T/t_ref:.0f} frames/s), "
      f"x{t_ref/t_ours:.1f}; identical tokens: {got == want}; symbols {int(counts.sum())} of {B*T} frames")
