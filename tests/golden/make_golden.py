"""Generates tests/golden/*.npz IN THE BUILD CONTAINER (needs /root/reference and torchaudio).

Run:  python tests/golden/make_golden.py
The fixtures are committed; the GPU box (no /root/reference) only reads them.

Sources of truth recorded per fixture:
  * joint logits  -- the reference's OWN ``JointNet`` class (rnnt/model.py:9-37), imported
    unmodified from /root/reference with a stub module standing in for the absent
    ``warprnnt_pytorch`` import at model.py:6 (the stub is never called).
  * loss / grads  -- ``torchaudio.functional.rnnt_loss`` (independent compiled CPU code; NOT
    the reference's dependency, which cannot be installed offline) in float32, and the
    float64 oracle (oracle/rnnt_oracle.py) for tight comparisons.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

stub = types.ModuleType("warprnnt_pytorch")
stub.RNNTLoss = lambda *a, **k: None
sys.modules["warprnnt_pytorch"] = stub
sys.path.insert(0, "/root/reference")
from rnnt.model import JointNet as RefJointNet  # noqa: E402

import torchaudio.functional as AF  # noqa: E402
from oracle import rnnt_oracle as O  # noqa: E402

CASES = {
    # name: (B, T, U, E, D, H, V, ragged)
    "tiny": (2, 5, 3, 6, 4, 8, 7, True),
    "timit_like": (3, 9, 5, 16, 16, 64, 49, True),
    "aishell_like": (2, 6, 4, 24, 24, 64, 300, True),
    "edge_u0_t1": (3, 4, 3, 8, 8, 64, 11, "edge"),
}


def main():
    for name, (B, T, U, E, D, H, V, ragged) in CASES.items():
        torch.manual_seed(2019)  # the reference's seed (config/aishell.yaml:46)
        joint = RefJointNet(E + D, H, V)
        enc = torch.randn(B, T, E)
        dec = torch.randn(B, U + 1, D)
        labels = torch.randint(1, V, (B, U), dtype=torch.int32)
        if ragged == "edge":
            act_lens = torch.tensor([T, 1, T - 1], dtype=torch.int32)[:B]
            label_lens = torch.tensor([U, 0, 0], dtype=torch.int32)[:B]
        else:
            act_lens = torch.randint((T + 1) // 2, T + 1, (B,), dtype=torch.int32)
            label_lens = torch.randint((U + 1) // 2, U + 1, (B,), dtype=torch.int32)
            act_lens[0], label_lens[0] = T, U
        for b in range(B):
            labels[b, int(label_lens[b]):] = 0  # zero padded like dataset.py:44
        enc.requires_grad_(True)
        dec.requires_grad_(True)
        logits = joint(enc, dec)  # reference code path (repeat + cat + 2 Linear)
        logits.retain_grad()
        loss_ta = AF.rnnt_loss(logits, labels, act_lens, label_lens, blank=0, reduction="none")
        loss_ta.sum().backward()
        sd = {k: v.detach().numpy() for k, v in joint.state_dict().items()}
        # float64 oracle on the reference's logits
        xd = logits.detach().double().requires_grad_(True)
        l64 = O.rnnt_loss(xd, labels, act_lens, label_lens)
        l64.sum().backward()
        np.savez_compressed(
            os.path.join(HERE, f"{name}.npz"),
            enc=enc.detach().numpy(), dec=dec.detach().numpy(), labels=labels.numpy(),
            act_lens=act_lens.numpy(), label_lens=label_lens.numpy(),
            w1=sd["forward_layer.weight"], b1=sd["forward_layer.bias"],
            w2=sd["project_layer.weight"], b2=sd["project_layer.bias"],
            ref_joint_logits=logits.detach().numpy(),
            torchaudio_loss=loss_ta.detach().numpy(), torchaudio_dlogits=logits.grad.numpy(),
            torchaudio_denc=enc.grad.numpy(), torchaudio_ddec=dec.grad.numpy(),
            oracle64_loss=l64.detach().numpy(), oracle64_dlogits=xd.grad.numpy(),
        )
        print(name, "loss(torchaudio)", loss_ta.detach().numpy(), "loss(oracle64)", l64.detach().numpy())


if __name__ == "__main__":
    main()
