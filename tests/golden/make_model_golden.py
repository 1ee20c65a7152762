"""Generates tests/golden/ref_model_tiny.npz IN THE BUILD CONTAINER (needs /root/reference).

Run:  python tests/golden/make_model_golden.py
The reference's OWN ``Transducer`` (rnnt/model.py:40-111: BaseEncoder + BaseDecoder + JointNet),
imported unmodified with a stub for the absent ``warprnnt_pytorch`` (model.py:6; the stub loss is
constructed but never called), is built from a tiny config, and the fixture records
  * its complete state_dict (so that key names and shapes are pinned),
  * the encoder / decoder outputs on seeded inputs (rnnt/model.py:63-66),
  * the token lists its ``recognize`` (greedy search, model.py:74-111) returns.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
stub = types.ModuleType("warprnnt_pytorch")
stub.RNNTLoss = lambda *a, **k: None
sys.modules["warprnnt_pytorch"] = stub
sys.modules["editdistance"] = types.ModuleType("editdistance")  # imported by rnnt/utils.py:3, unused here
sys.path.insert(0, "/root/reference")
from rnnt.model import Transducer  # noqa: E402
from rnnt.utils import AttrDict  # noqa: E402

CFG = dict(enc=dict(type="lstm", hidden_size=16, output_size=12, n_layers=2, bidirectional=True),
           dec=dict(type="lstm", hidden_size=16, output_size=12, n_layers=1),
           joint=dict(input_size=24, inner_size=64), vocab_size=30, share_weight=False,
           share_embedding=False, feature_dim=20, dropout=0.0)


def main():
    torch.manual_seed(2019)
    model = Transducer(AttrDict(CFG)).eval()
    with torch.no_grad():  # spread the logits so that greedy search emits blanks and symbols
        model.joint.project_layer.weight.mul_(20.0)
        model.joint.project_layer.bias.mul_(4.0)
        model.joint.forward_layer.weight.mul_(4.0)
        model.encoder.output_proj.weight.mul_(4.0)
        model.decoder.output_proj.weight.mul_(4.0)
    B, T, U = 3, 14, 5
    inputs = torch.randn(B, T, CFG["feature_dim"]) * 2
    inputs_length = torch.tensor([14, 9, 11])
    targets = torch.randint(1, CFG["vocab_size"], (B, U))
    targets_length = torch.tensor([5, 3, 4])
    for b in range(B):
        targets[b, int(targets_length[b]):] = 0
    with torch.no_grad():  # blank bias calibrated so that about half of the frames emit nothing
        e, _ = model.encoder(inputs, inputs_length)
        d0, _ = model.decoder(torch.zeros(1, 1, dtype=torch.long))
        lg = model.joint.project_layer(torch.tanh(model.joint.forward_layer(
            torch.cat((e.reshape(-1, e.shape[-1]), d0.view(1, -1).expand(B * T, -1)), dim=-1))))
        model.joint.project_layer.bias[0] += (lg[:, 1:].max(dim=1).values - lg[:, 0]).median()
    with torch.no_grad():
        enc_state, _ = model.encoder(inputs, inputs_length)
        concat = torch.nn.functional.pad(targets, pad=(1, 0, 0, 0), value=0)
        dec_state, _ = model.decoder(concat, targets_length.add(1))
        logits = model.joint(enc_state, dec_state)
        tokens = model.recognize(inputs, inputs_length)
    n_sym = sum(len(t) for t in tokens)
    assert 0 < n_sym < int(inputs_length.sum()), (n_sym, tokens)
    out = {"sd__" + k: v.detach().numpy() for k, v in model.state_dict().items()}
    out.update(inputs=inputs.numpy(), inputs_length=inputs_length.numpy(), targets=targets.numpy(),
               targets_length=targets_length.numpy(), enc_state=enc_state.numpy(), dec_state=dec_state.numpy(),
               logits=logits.numpy(),
               tokens=np.array([t + [0] * (T - len(t)) for t in tokens], dtype=np.int64),
               token_counts=np.array([len(t) for t in tokens], dtype=np.int64))
    np.savez_compressed(os.path.join(HERE, "ref_model_tiny.npz"), **out)
    print("wrote ref_model_tiny.npz:", {k: v.shape for k, v in out.items() if k.startswith("sd__")}, tokens)


if __name__ == "__main__":
    main()
