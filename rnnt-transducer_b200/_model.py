"""The callers either side of the hot path: a ``Transducer`` with the
reference's module tree -- ``encoder.{lstm,output_proj}``, ``decoder.{embedding,lstm,output_proj}``,
``joint.{forward_layer,project_layer}`` (rnnt/encoder.py:5-20, rnnt/decoder.py:5-22,
rnnt/model.py:9-16,40-59) -- so that reference checkpoints (rnnt/utils.py:74-85) load with
``strict=True``, whose training ``forward`` runs the fused loss on the LSTM outputs (output
projections folded into the first joint layer) and whose ``recognize`` is the batched on-device
greedy search. The LSTMs are stock ``nn.LSTM`` (north_star: they stay PyTorch modules).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence

from ._api import JointNet, JointRNNTLoss, ProjectedJointRNNTLoss, greedy_decode

__all__ = ["Encoder", "Decoder", "Transducer"]


def _run_lstm(lstm: nn.LSTM, x, lengths, hidden=None):
    """LSTM over a zero-padded batch; with ``lengths`` the padding never enters the recurrence
    (what the reference's sort -> pack -> LSTM -> unpack -> unsort does, rnnt/encoder.py:25-36,
    rnnt/decoder.py:28-40) and the output is trimmed to the longest sequence of the batch."""
    if lengths is None:
        return lstm(x, hidden)
    packed = pack_padded_sequence(x, lengths.detach().cpu().long(), batch_first=True, enforce_sorted=False)
    out, hidden = lstm(packed, hidden)
    out, _ = pad_packed_sequence(out, batch_first=True)
    return out, hidden


class Encoder(nn.Module):
    """Same parameters and ``forward`` contract as ``BaseEncoder`` (rnnt/encoder.py:5-40).
    ``rnn_out`` stops before ``output_proj`` (input of ``ProjectedJointRNNTLoss``). The returned
    hidden state is in batch order (the reference returns it in length-sorted order and never
    uses it)."""

    def __init__(self, input_size, hidden_size, output_size, n_layers, dropout=0.2, bidirectional=True):
        super().__init__()
        self.lstm = nn.LSTM(input_size=input_size, hidden_size=hidden_size, num_layers=n_layers, batch_first=True,
                            dropout=dropout, bidirectional=bidirectional)
        self.output_proj = nn.Linear(2 * hidden_size if bidirectional else hidden_size, output_size, bias=True)

    def rnn_out(self, inputs, input_lengths=None):
        if inputs.dim() != 3:
            raise ValueError("encoder inputs must be (B,T,F)")
        return _run_lstm(self.lstm, inputs, input_lengths)

    def forward(self, inputs, input_lengths=None):
        out, hidden = self.rnn_out(inputs, input_lengths)
        return self.output_proj(out), hidden


class Decoder(nn.Module):
    """Same parameters and ``forward`` contract as ``BaseDecoder`` (rnnt/decoder.py:5-44):
    embedding (padding_idx 0) -> LSTM -> output_proj, optional weight tying."""

    def __init__(self, hidden_size, vocab_size, output_size, n_layers, dropout=0.2, share_weight=False):
        super().__init__()
        self.embedding = nn.Embedding(vocab_size, hidden_size, padding_idx=0)
        self.lstm = nn.LSTM(input_size=hidden_size, hidden_size=hidden_size, num_layers=n_layers, batch_first=True,
                            dropout=dropout if n_layers > 1 else 0)
        self.output_proj = nn.Linear(hidden_size, output_size)
        if share_weight:
            self.embedding.weight = self.output_proj.weight

    def rnn_out(self, inputs, length=None, hidden=None):
        return _run_lstm(self.lstm, self.embedding(inputs), length, hidden)

    def forward(self, inputs, length=None, hidden=None):
        out, hidden = self.rnn_out(inputs, length, hidden)
        return self.output_proj(out), hidden


class Transducer(nn.Module):
    """``Transducer(cfg)(inputs, inputs_length, targets, targets_length) -> loss`` and
    ``.recognize(inputs, inputs_length) -> list of token lists`` as rnnt/model.py:40-111, with
    ``cfg`` the ``model`` section of the reference's YAML as a plain dict
    (config/aishell.yaml:21-39). ``fold_projections=False`` keeps ``output_proj`` as separate
    GEMMs ahead of ``JointRNNTLoss``."""

    def __init__(self, cfg: dict, fold_projections: bool = True):
        super().__init__()
        enc, dec, joint = cfg["enc"], cfg["dec"], cfg["joint"]
        drop = cfg.get("dropout", 0.0) or 0.0
        self.encoder = Encoder(cfg["feature_dim"], enc["hidden_size"], enc["output_size"], enc["n_layers"], drop,
                               enc.get("bidirectional", True))
        self.decoder = Decoder(dec["hidden_size"], cfg["vocab_size"], dec["output_size"], dec["n_layers"], drop,
                               bool(cfg.get("share_weight", False)))
        self.joint = JointNet(joint["input_size"], joint["inner_size"], cfg["vocab_size"])
        if cfg.get("share_embedding", False):  # rnnt/model.py:55-57
            if self.decoder.embedding.weight.shape != self.joint.project_layer.weight.shape:
                raise ValueError("share_embedding needs dec.hidden_size == joint.inner_size")
            self.joint.project_layer.weight = self.decoder.embedding.weight
        self.fold_projections = fold_projections

    @property
    def crit(self):
        """The loss (rnnt/model.py:59). Built on demand so that it does not register the modules it
        reads a second time: the state_dict keys stay exactly the reference's."""
        if self.fold_projections:
            return ProjectedJointRNNTLoss(self.joint, self.encoder.output_proj, self.decoder.output_proj)
        return JointRNNTLoss(self.joint)

    def forward(self, inputs, inputs_length, targets, targets_length):
        # rnnt/model.py:61-72: blank-prefixed targets feed the prediction network, lengths + 1
        concat_targets = F.pad(targets, pad=(1, 0, 0, 0), value=0)
        if self.fold_projections:
            enc, _ = self.encoder.rnn_out(inputs, inputs_length)
            dec, _ = self.decoder.rnn_out(concat_targets, targets_length.add(1))
        else:
            enc, _ = self.encoder(inputs, inputs_length)
            dec, _ = self.decoder(concat_targets, targets_length.add(1))
        labels = targets[:, :dec.shape[1] - 1].int().contiguous()  # outputs are trimmed to the longest sequence
        # Lengths that arrive on the host (as a DataLoader yields them, train.py:24-31) size the
        # tile loops exactly and cost no device read-back anywhere in the step (the reference
        # syncs twice per step on .max().item(), train.py:30-33).
        host_lens = (inputs_length, targets_length) if not (inputs_length.is_cuda or targets_length.is_cuda) else None
        return self.crit(enc.contiguous(), dec.contiguous(), labels, inputs_length.int().to(enc.device),
                         targets_length.int().to(enc.device), host_lens=host_lens)

    @torch.no_grad()
    def recognize(self, inputs, inputs_length):
        enc_states, _ = self.encoder(inputs, inputs_length)
        tokens, counts = greedy_decode(enc_states.contiguous(), inputs_length, self.decoder, self.joint)
        tokens, counts = tokens.cpu(), counts.cpu()
        return [tokens[i, :int(counts[i])].tolist() for i in range(tokens.shape[0])]
