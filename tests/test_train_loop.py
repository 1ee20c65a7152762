"""The loop around the hot path: host logic on CPU (optimizer
mirror, trimming, edit distance, checkpoint layout, the loop's bookkeeping on a stand-in model);
on the GPU a few epochs of the assembled model over the Kaldi fixture directory."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

import rnnt_transducer_b200 as R
from tests.golden.make_features_golden_cfg import VARIANTS, data_config
from tests.helpers import REF_MODEL_TINY_CFG

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIX = os.path.join(ROOT, "tests", "golden", "kaldi_tiny")


def test_edit_distance_known_answers():
    assert R.edit_distance([], []) == 0 and R.edit_distance([1, 2, 3], []) == 3
    assert R.edit_distance([1, 2, 3], [1, 3]) == 1 and R.edit_distance([1, 2, 3, 4], [2, 3, 4, 5]) == 2
    assert R.edit_distance(list("kitten"), list("sitting")) == 3
    rng = np.random.default_rng(0)
    for _ in range(50):  # symmetry and the triangle inequality on random strings
        a, b, c = (rng.integers(0, 4, rng.integers(0, 9)).tolist() for _ in range(3))
        assert R.edit_distance(a, b) == R.edit_distance(b, a)
        assert R.edit_distance(a, c) <= R.edit_distance(a, b) + R.edit_distance(b, c)
    assert R.compute_cer([[1, 2], [3]], [[1, 2, 3], [4]]) == (2, 4)


def test_optimizer_mirror_counts_and_decay():
    w = nn.Parameter(torch.ones(3))
    opt = R.Optimizer([w], dict(type="sgd", lr=0.5, momentum=0.0, decay_ratio=0.5, weight_decay=0))
    assert opt.global_step == 1 and opt.current_epoch == 0  # rnnt/optim.py:8-9
    w.grad = torch.ones(3)
    opt.step()
    assert opt.global_step == 2 and torch.allclose(w.detach(), torch.full((3,), 0.5))
    opt.decay_lr()
    assert opt.lr == 0.25 and opt.optimizer.param_groups[0]["lr"] == 0.25
    opt.zero_grad()
    assert w.grad is None
    for kind in ("adam", "adadelta"):
        R.Optimizer([w], dict(type=kind, lr=1e-3))
    with pytest.raises(NotImplementedError):
        R.Optimizer([w], dict(type="lamb", lr=1e-3))


def test_trim_batch_uses_host_lengths():
    x, y = torch.zeros(2, 9, 4), torch.zeros(2, 6, dtype=torch.long)
    xt, yt = R.trim_batch(x, torch.tensor([5, 7]), y, torch.tensor([2, 3]))
    assert xt.shape == (2, 7, 4) and yt.shape == (2, 3)


class _Quadratic(nn.Module):
    """Stand-in with the Transducer call signature: loss = mean((w - mean(inputs))^2)."""

    def __init__(self):
        super().__init__()
        self.w = nn.Parameter(torch.zeros(()))

    def forward(self, inputs, inputs_length, targets, targets_length):
        return (self.w - inputs.mean()) ** 2


def test_train_epoch_bookkeeping_on_cpu():
    model = _Quadratic()
    opt = R.Optimizer(model.parameters(), dict(type="sgd", lr=0.25, momentum=0.0))
    batches = [(torch.full((2, 6, 3), 2.0), torch.tensor([4, 6]), torch.zeros(2, 5, dtype=torch.long), torch.tensor([1, 2]))] * 4
    logs = []
    avg = R.train_epoch(model, batches, opt, max_grad_norm=100.0, show_interval=2, log=logs.append)
    # w: 0 -> 1 -> 1.5 -> 1.75 -> 1.875; losses 4, 1, .25, .0625
    assert abs(avg - (4 + 1 + 0.25 + 0.0625) / 4) < 1e-6 and abs(float(model.w) - 1.875) < 1e-6
    assert opt.global_step == 5 and opt.current_epoch == 1
    assert [l["global_step"] for l in logs] == [2, 4] and abs(logs[1]["loss"] - 0.25) < 1e-6
    assert abs(logs[0]["grad_norm"] - 4.0) < 1e-6 and abs(logs[1]["avg_loss"] - 1.75) < 1e-6


def test_checkpoint_layout_round_trip(tmp_path):
    torch.manual_seed(1)
    a, b = R.Transducer(REF_MODEL_TINY_CFG), R.Transducer(REF_MODEL_TINY_CFG)
    opt = R.Optimizer(a.parameters(), dict(type="sgd", lr=0.1, momentum=0.9))
    opt.epoch()
    path = str(tmp_path / "m.chkpt")
    R.save_checkpoint(a, opt, path)
    ckpt = torch.load(path)
    assert set(ckpt) == {"encoder", "decoder", "joint", "optimizer", "epoch", "step"}  # rnnt/utils.py:76-83
    assert set(ckpt["joint"]) == {"forward_layer.weight", "forward_layer.bias", "project_layer.weight", "project_layer.bias"}
    assert R.load_checkpoint(b, path, parts=("encoder",)) == 1
    assert torch.equal(a.encoder.lstm.weight_ih_l0, b.encoder.lstm.weight_ih_l0)
    assert not torch.equal(a.joint.project_layer.weight, b.joint.project_layer.weight)
    R.load_checkpoint(b, path)
    for (k, p), (_, q) in zip(a.state_dict().items(), b.state_dict().items()):
        assert torch.equal(p, q), k


@pytest.mark.gpu
def test_training_and_validation_over_kaldi_directory():
    torch.manual_seed(2019)
    name = "fm"
    cfg = data_config(name, VARIANTS[name][1], VARIANTS[name][2])
    ds = R.KaldiFeatureSet(cfg, "train", root=FIX)
    mcfg = dict(enc=dict(type="lstm", hidden_size=32, output_size=24, n_layers=1, bidirectional=True),
                dec=dict(type="lstm", hidden_size=32, output_size=24, n_layers=1),
                joint=dict(input_size=48, inner_size=64), vocab_size=7, feature_dim=32, dropout=0.0)
    model = R.Transducer(mcfg).cuda()
    opt = R.Optimizer(model.parameters(), dict(type="adam", lr=1e-2))
    logs, avgs = [], []
    for epoch in range(30):
        avgs.append(R.train_epoch(model, ds.loader(batch_size=3), opt, max_grad_norm=200.0, show_interval=1,
                                  log=logs.append))
    assert all(np.isfinite(avgs)) and avgs[-1] < 0.5 * avgs[0], avgs
    assert opt.global_step == 1 + 30 * 2 and len(logs) == 60 and logs[-1]["grad_norm"] > 0
    cer = R.eval_epoch(model, ds.loader(batch_size=3))
    assert 0.0 <= cer <= 300.0
    # the loop's loss is the model's own: one more forward on the first batch agrees with the log
    model.eval()
    x, xl, y, yl = next(iter(ds.loader(batch_size=3)))
    assert torch.isfinite(model(x, xl, y, yl)).all()


def test_checkpoint_optimizer_entry_has_no_implementation_only_keys():
    # 'fused' / 'foreach' in the saved param groups would switch the reference's plain optimiser
    # (rnnt/optim.py) onto another code path when it loads the checkpoint (train.py:190, -mode continue)
    import torch
    import rnnt_transducer_b200 as R
    lin = torch.nn.Linear(3, 2)
    opt = R.Optimizer(lin.parameters(), dict(type="sgd", lr=0.5, momentum=0.9))
    lin(torch.ones(1, 3)).sum().backward()
    opt.step()
    opt.decay_lr()
    sd = opt.state_dict()
    for g in sd["param_groups"]:
        assert not {"fused", "foreach", "capturable", "differentiable"} & set(g)
    plain = torch.optim.SGD(lin.parameters(), lr=0.1, momentum=0.9)     # what the reference constructs
    plain.load_state_dict(sd)
    assert plain.param_groups[0]["lr"] == opt.lr
    other = R.Optimizer(lin.parameters(), dict(type="sgd", lr=0.5, momentum=0.9))
    other.load_state_dict(plain.state_dict())                           # and back, from a reference checkpoint
    assert other.lr == opt.lr and other.optimizer.param_groups[0]["lr"] == opt.lr
