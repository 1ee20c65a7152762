"""Pins the oracle (oracle/) BEFORE it is trusted as the parity checker: golden vectors made
by the reference's own JointNet and by torchaudio's independent loss, closed forms, and
agreement between its three restatements (torch autograd, explicit Python, C)."""
import math

import numpy as np
import pytest
import torch

from oracle import c_oracle as C
from oracle import rnnt_oracle as O
from tests.helpers import load_golden, max_abs, rel_l2

CASES = ["tiny", "timit_like", "aishell_like", "edge_u0_t1"]


@pytest.mark.parametrize("name", CASES)
def test_joint_restatement_matches_reference_jointnet(name):
    g = load_golden(name)
    t = lambda k: torch.from_numpy(g[k])
    logits = O.joint_logits(t("enc"), t("dec"), t("w1"), t("b1"), t("w2"), t("b2"))
    assert max_abs(logits, g["ref_joint_logits"]) < 5e-6


@pytest.mark.parametrize("name", CASES)
def test_loss_oracle_matches_torchaudio_golden(name):
    g = load_golden(name)
    x = torch.from_numpy(g["ref_joint_logits"]).double().requires_grad_(True)
    y, tl, ul = (torch.from_numpy(g[k]) for k in ("labels", "act_lens", "label_lens"))
    loss = O.rnnt_loss(x, y, tl, ul)
    loss.sum().backward()
    assert np.allclose(loss.detach().numpy(), g["torchaudio_loss"], rtol=1e-5)
    assert np.allclose(loss.detach().numpy(), g["oracle64_loss"], rtol=1e-12)
    assert max_abs(x.grad, g["torchaudio_dlogits"]) < 1e-5
    assert max_abs(x.grad, g["oracle64_dlogits"]) < 1e-12


@pytest.mark.parametrize("name", CASES)
def test_c_oracle_matches_torch_oracle(name):
    g = load_golden(name)
    loss, grad = C.rnnt_loss(g["ref_joint_logits"], g["labels"], g["act_lens"], g["label_lens"])
    assert np.allclose(loss, g["oracle64_loss"], rtol=1e-12)
    assert max_abs(grad, g["oracle64_dlogits"]) < 1e-7  # grad is stored as float32


def test_explicit_coefficients_match_autograd():
    torch.manual_seed(1)
    B, T, U, V = 3, 6, 4, 9
    x = torch.randn(B, T, U + 1, V)
    y = torch.randint(1, V, (B, U), dtype=torch.int32)
    tl = torch.tensor([6, 4, 1], dtype=torch.int32)
    ul = torch.tensor([4, 0, 2], dtype=torch.int32)
    xd = x.double().requires_grad_(True)
    O.rnnt_loss(xd, y, tl, ul).sum().backward()
    loss_e, grad_e, cb, cl = O.rnnt_loss_and_grad_explicit(x, y, tl, ul)
    assert max_abs(grad_e, xd.grad) < 1e-12
    _, _, cbc, clc = C.rnnt_loss(x.numpy(), y.numpy(), tl.numpy(), ul.numpy(), want_coeffs=True)
    assert max_abs(cb, cbc) < 1e-12 and max_abs(cl, clc) < 1e-12
    # outside the valid window the gradient is exactly zero
    assert float(xd.grad[1, 4:].abs().max()) == 0.0 and float(xd.grad[1, :, 1:].abs().max()) == 0.0


def test_known_answers():
    # T=1, U=0: loss = -log_softmax(x)[blank]
    x = torch.randn(1, 1, 1, 5)
    l = O.rnnt_loss(x, torch.zeros(1, 0, dtype=torch.int32), torch.tensor([1]), torch.tensor([0]))
    assert abs(float(l) + float(torch.log_softmax(x.double(), -1)[0, 0, 0, 0])) < 1e-12
    # uniform logits: (T+U) log V - log C(T+U-1, U)
    for (T, U, V) in [(5, 3, 9), (2, 2, 5), (7, 0, 4), (1, 4, 6)]:
        x = torch.zeros(1, T, U + 1, V)
        y = torch.randint(1, V, (1, U), dtype=torch.int32)
        l = O.rnnt_loss(x, y, torch.tensor([T]), torch.tensor([U]))
        assert abs(float(l) - O.uniform_closed_form(T, U, V)) < 1e-10
        lc, _ = C.rnnt_loss(x.numpy(), y.numpy(), np.array([T]), np.array([U]))
        assert abs(lc[0] - O.uniform_closed_form(T, U, V)) < 1e-10


def test_gradient_rows_sum_to_zero():
    torch.manual_seed(3)
    x = torch.randn(2, 5, 4, 8)
    y = torch.randint(1, 8, (2, 3), dtype=torch.int32)
    _, grad = C.rnnt_loss(x.numpy(), y.numpy(), np.array([5, 3]), np.array([3, 2]))
    assert np.abs(grad.astype(np.float64).sum(-1)).max() < 1e-6


def test_reductions_and_bf16_quantised_variant():
    g = load_golden("timit_like")
    t = lambda k: torch.from_numpy(g[k])
    args = (t("enc"), t("dec"), t("w1"), t("b1"), t("w2"), t("b2"), t("labels"), t("act_lens"), t("label_lens"))
    none = O.joint_rnnt_loss(*args, reduction="none")
    assert np.allclose(none.numpy(), g["oracle64_loss"], rtol=1e-6)
    assert abs(float(O.joint_rnnt_loss(*args, reduction="mean")) - float(none.sum() / 3)) < 1e-12
    assert abs(float(O.joint_rnnt_loss(*args, reduction="sum")) - float(none.sum())) < 1e-12
    q = O.joint_rnnt_loss(*args, reduction="none", quantize_bf16=True)
    assert rel_l2(q, none) < 2e-2 and rel_l2(q, none) > 0


def test_c_oracle_rejects_bad_arguments():
    x = np.zeros((1, 2, 2, 4), np.float32)
    with pytest.raises(ValueError):
        C.rnnt_loss(x, np.array([[9]]), np.array([2]), np.array([1]))   # label >= V
    with pytest.raises(ValueError):
        C.rnnt_loss(x, np.array([[1]]), np.array([3]), np.array([1]))   # act_len > T
    with pytest.raises(ValueError):
        C.rnnt_loss(x, np.array([[1]]), np.array([2]), np.array([2]))   # label_len > U


def test_hypothesis_ragged_shapes_c_oracle_vs_torch_oracle():
    # randomly drawn ragged shapes; the two independent restatements
    # (C, OpenMP float64 / torch autograd float64) must agree and keep the structural properties
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=25, deadline=None, derandomize=True)
    @given(st.integers(1, 3), st.integers(1, 7), st.integers(0, 5), st.integers(2, 11), st.integers(0, 2 ** 16))
    def check(B, T, U, V, seed):
        g = torch.Generator().manual_seed(seed)
        x = torch.randn(B, T, U + 1, V, generator=g)
        y = torch.randint(1, V, (B, max(U, 1)), generator=g, dtype=torch.int32)[:, :U]
        tl = torch.randint(1, T + 1, (B,), generator=g, dtype=torch.int32)
        ul = torch.randint(0, U + 1, (B,), generator=g, dtype=torch.int32)
        xd = x.double().requires_grad_(True)
        lo = O.rnnt_loss(xd, y, tl, ul)
        lo.sum().backward()
        lc, gc = C.rnnt_loss(x.numpy(), y.numpy().reshape(B, U), tl.numpy(), ul.numpy())
        assert np.allclose(lc, lo.detach().numpy(), rtol=1e-10)
        assert max_abs(gc, xd.grad) < 1e-6
        for b in range(B):
            tb, ub = int(tl[b]), int(ul[b])
            assert float(np.abs(gc[b, tb:]).max(initial=0.0)) == 0.0
            assert float(np.abs(gc[b, :, ub + 1:]).max(initial=0.0)) == 0.0
            assert float(np.abs(gc[b, :tb, :ub + 1].sum(-1)).max()) < 1e-5   # softmax-Jacobian rows sum to 0
    check()


@pytest.mark.parametrize("dims,mode", [((3, 13, 6, 16, 12, 64, 50), "edge"), ((2, 9, 4, 8, 8, 32, 20), "ragged"),
                                       ((1, 1, 0, 8, 8, 32, 5), "full")])
@pytest.mark.parametrize("quantize", [False, True])
def test_large_shape_oracle_agrees_with_the_materialising_oracle(dims, mode, quantize):
    # oracle/large_oracle.py (per utterance, t-chunked, bounded memory) is the checker at the BASELINE shapes;
    # here it is pinned on the oracle that materialises the logits, at shapes both can run
    from oracle import large_oracle as LO
    from tests.gpu_checks import _joint_inputs
    enc, dec, w1, b1, w2, b2, y, tl, ul = _joint_inputs(*dims, mode, 3)
    ol = [t.double().requires_grad_(True) for t in (enc, dec, w1, b1, w2, b2)]
    lo = O.joint_rnnt_loss(*ol, y, tl, ul, reduction="none", quantize_bf16=quantize)
    lo.sum().backward()
    l2, g = LO.joint_rnnt_loss_and_grads(enc, dec, w1, b1, w2, b2, y, tl, ul, quantize_bf16=quantize,
                                         cells_per_chunk=20)
    assert float((l2 - lo.detach()).abs().max()) < 1e-11
    # (quantised mode: the fp32 addition of the two projections rounds the gradient that passes through it)
    for n, t in zip(["enc", "dec", "w1", "b1", "w2", "b2"], ol):
        assert max_abs(g[n], t.grad) < (1e-6 if quantize else 1e-11), n
    l3, none = LO.joint_rnnt_loss_and_grads(enc, dec, w1, b1, w2, b2, y, tl, ul, quantize_bf16=quantize, grads=False)
    assert none is None and float((l3 - l2).abs().max()) < 1e-11
