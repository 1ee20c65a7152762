"""Shared test helpers: golden fixtures, synthetic inputs, error metrics."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    return {k: z[k] for k in z.files}


def synth_lengths(B, T, U, mode, gen):
    """mode: 'full' | 'ragged' (T_b ~ U{ceil(T/2)..T}, U_b ~ U{ceil(U/2)..U}, utterance 0 full)
    | 'edge' (adds U_b=0 and T_b=1 utterances)."""
    if mode == "full":
        return torch.full((B,), T, dtype=torch.int32), torch.full((B,), U, dtype=torch.int32)
    tl = torch.randint((T + 1) // 2, T + 1, (B,), generator=gen, dtype=torch.int32)
    ul = torch.randint((U + 1) // 2, U + 1, (B,), generator=gen, dtype=torch.int32)
    tl[0], ul[0] = T, U
    if mode == "edge" and B >= 3:
        ul[1] = 0
        tl[2] = 1
    return tl, ul


def synth_labels(B, U, V, label_lens, gen):
    y = torch.randint(1, V, (B, max(U, 0)), generator=gen, dtype=torch.int32)
    for b in range(B):
        y[b, int(label_lens[b]):] = 0
    return y


def rel_l2(a, b):
    a, b = torch.as_tensor(a).detach().double().cpu(), torch.as_tensor(b).detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


def max_abs(a, b):
    a, b = torch.as_tensor(a).detach().double().cpu(), torch.as_tensor(b).detach().double().cpu()
    return float((a - b).abs().max())


def make_decoder(hidden_size, vocab_size, output_size, n_layers=1, seed=0):
    """A prediction network with the interface of the reference's BaseDecoder
    (rnnt/decoder.py:5-44): embedding -> LSTM -> output_proj; forward(inputs, length=None,
    hidden=None) -> (outputs, hidden). Only the length=None path (decoding) is needed here."""
    import torch
    import torch.nn as nn

    class Decoder(nn.Module):
        def __init__(self):
            super().__init__()
            self.embedding = nn.Embedding(vocab_size, hidden_size, padding_idx=0)
            self.lstm = nn.LSTM(input_size=hidden_size, hidden_size=hidden_size, num_layers=n_layers,
                                batch_first=True)
            self.output_proj = nn.Linear(hidden_size, output_size)

        def forward(self, inputs, length=None, hidden=None):
            assert length is None
            outputs, hidden = self.lstm(self.embedding(inputs), hidden)
            return self.output_proj(outputs), hidden

    torch.manual_seed(seed)
    return Decoder()


REF_MODEL_TINY_CFG = dict(enc=dict(type="lstm", hidden_size=16, output_size=12, n_layers=2, bidirectional=True),
                          dec=dict(type="lstm", hidden_size=16, output_size=12, n_layers=1),
                          joint=dict(input_size=24, inner_size=64), vocab_size=30, share_weight=False,
                          share_embedding=False, feature_dim=20, dropout=0.0)  # = tests/golden/make_model_golden.py


def load_ref_model_tiny():
    """(Transducer with the reference's weights, fixture dict) from tests/golden/ref_model_tiny.npz."""
    import torch
    import rnnt_transducer_b200 as R
    g = load_golden("ref_model_tiny")
    model = R.Transducer(REF_MODEL_TINY_CFG).eval()
    sd = {k[4:]: torch.from_numpy(g[k]) for k in g if k.startswith("sd__")}
    model.load_state_dict(sd, strict=True)   # same key names and shapes as the reference's Transducer
    return model, g
