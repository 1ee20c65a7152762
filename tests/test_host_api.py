"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol the
header declares, pure shape helpers agree with their definition, and the Python mirror of
the reference interface validates arguments and refuses to run without CUDA (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

import rnnt_transducer_b200 as R
from rnnt_transducer_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "rnnt_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rnnt_b200_\w+)\s*\(", src)))


def test_library_exports_every_header_symbol():
    assert os.path.exists(_lib.SO_PATH), "run __graft_entry__.build() first"
    handle = ctypes.CDLL(_lib.SO_PATH)
    syms = _header_symbols()
    assert len(syms) >= 16
    for s in syms:
        assert hasattr(handle, s), f"{s} declared in include/rnnt_b200.h but not exported"
    assert set(syms) == set(_lib.SIGNATURES), "ctypes signature table out of sync with the header"


def test_shape_helpers_without_gpu():
    L = _lib.lib()
    assert L.rnnt_b200_version() == 100
    assert L.rnnt_b200_plane_elems(2, 5, 4) == 2 * (5 + 4 - 1) * 4
    assert L.rnnt_b200_max_tiles(32, 500, 101) == 32 * ((500 * 101 + 127) // 128)
    assert L.rnnt_b200_vocab_tile(4232) == 256 and L.rnnt_b200_vocab_pad(4232) == 4352
    assert L.rnnt_b200_vocab_tile(49) == 64 and L.rnnt_b200_vocab_pad(49) == 64
    assert L.rnnt_b200_h_pitch(512) == 512 and L.rnnt_b200_h_pitch(320) == 352
    # dS pitch: V rounded up to the UMMA N granularity (16), never above the W2 pack
    assert L.rnnt_b200_ds_pitch(4232) == 4288 and L.rnnt_b200_ds_pitch(49) == 64
    assert L.rnnt_b200_ds_pitch(256) == 256 and L.rnnt_b200_ds_pitch(270) == 320
    assert L.rnnt_b200_status_string(-3) == b"blank index not in [0,V)"


def test_c_abi_rejects_bad_arguments_without_launching():
    L = _lib.lib()
    assert L.rnnt_b200_logits_to_planes(None, None, None, None, 1, 1, 1, 4, 0, None, None, None, None) == -1
    one = ctypes.c_void_p(16)
    assert L.rnnt_b200_logits_to_planes(one, one, one, one, 1, 1, 1, 4, 7, one, one, one, None) == -3
    assert L.rnnt_b200_lattice(one, one, one, one, 1, 4, 2000, one, one, one, one, one, None, None) == -2
    assert L.rnnt_b200_joint_fwd(one, one, one, one, one, one, one, one, one,
                                 1, 4, 2, 10, 100, 0, 1, one, one, one, None) == -2   # H % 64
    assert L.rnnt_b200_joint_hidden(one, one, one, one, one, one, one, 1, 4, 2, 100, 1, one, None) == -2
    al = ctypes.c_void_p(256)
    assert L.rnnt_b200_joint_hidden(al, al, one, one, one, one, one, 1, 4, 2, 64, 99, al, None) == -2  # tile_cap
    assert L.rnnt_b200_joint_hidden(al, al, one, one, one, one, one, 1, 4, 2, 64, 1, one, None) == -4  # h_stash alignment


def test_rnntloss_is_dropin_constructible_and_validates():
    crit = R.RNNTLoss()  # no-arg, like rnnt/model.py:59
    assert crit.blank == 0 and crit.reduction == "mean"
    acts = torch.zeros(1, 2, 2, 4)
    y = torch.zeros(1, 1, dtype=torch.int32)
    tl = torch.tensor([2], dtype=torch.int32)
    ul = torch.tensor([1], dtype=torch.int32)
    with pytest.raises(RuntimeError, match="no CPU path"):
        crit(acts, y, tl, ul)
    with pytest.raises(ValueError, match="int32"):
        crit(acts, y.long(), tl, ul)
    with pytest.raises(ValueError, match="float32"):
        crit(acts.double(), y, tl, ul)
    with pytest.raises(ValueError):
        R.RNNTLoss(reduction="bogus")


def test_warprnnt_pytorch_alias_resolves_to_this_package():
    import warprnnt_pytorch
    assert warprnnt_pytorch.RNNTLoss is R.RNNTLoss


def test_jointnet_matches_reference_state_dict_layout_and_math():
    from tests.helpers import load_golden, max_abs
    g = load_golden("timit_like")
    E, D = g["enc"].shape[-1], g["dec"].shape[-1]
    H, V = g["w1"].shape[0], g["w2"].shape[0]
    j = R.JointNet(E + D, H, V)
    assert sorted(j.state_dict()) == ["forward_layer.bias", "forward_layer.weight",
                                      "project_layer.bias", "project_layer.weight"]
    j.load_state_dict({"forward_layer.weight": torch.from_numpy(g["w1"]), "forward_layer.bias": torch.from_numpy(g["b1"]),
                       "project_layer.weight": torch.from_numpy(g["w2"]), "project_layer.bias": torch.from_numpy(g["b2"])})
    out = j(torch.from_numpy(g["enc"]), torch.from_numpy(g["dec"]))
    assert max_abs(out, g["ref_joint_logits"]) < 5e-6
    # equal-rank branch used by greedy decoding (rnnt/model.py:28-29,90)
    e1, d1 = torch.from_numpy(g["enc"])[0, 0], torch.from_numpy(g["dec"])[0, 0]
    assert max_abs(j(e1, d1), g["ref_joint_logits"][0, 0, 0]) < 5e-6


def test_fused_entry_validates_shapes_on_cpu():
    j = R.JointNet(16, 100, 9)   # H = 100 is not a multiple of 64
    f = R.JointRNNTLoss(j)
    y = torch.zeros(1, 1, dtype=torch.int32)
    tl = torch.tensor([2], dtype=torch.int32)
    ul = torch.tensor([1], dtype=torch.int32)
    with pytest.raises(RuntimeError, match="no CPU path"):
        f(torch.zeros(1, 2, 8), torch.zeros(1, 2, 8), y, tl, ul)


def test_fold_output_projections_algebra():
    # SURVEY 8(f)-1, float64 on CPU: forward_layer([enc_proj(x); dec_proj(y)]) == folded form
    import torch
    from rnnt_transducer_b200 import fold_output_projections
    g = torch.Generator().manual_seed(3)
    He, Hd, E, D, H = 12, 10, 6, 5, 7
    r = lambda *s: torch.randn(*s, generator=g, dtype=torch.float64)
    Weo, beo, Wdo, bdo, W1, b1 = r(E, He), r(E), r(D, Hd), r(D), r(H, E + D), r(H)
    x, y = r(4, He), r(4, Hd)
    ref = torch.cat((x @ Weo.t() + beo, y @ Wdo.t() + bdo), dim=1) @ W1.t() + b1
    we, wd, b = fold_output_projections(Weo, beo, Wdo, bdo, W1, b1)
    assert torch.allclose(x @ we.t() + y @ wd.t() + b, ref, atol=1e-12)
    import pytest
    with pytest.raises(ValueError):
        fold_output_projections(Weo, beo, Wdo, bdo, W1[:, :-1], b1)


def test_dw2_gemm_rejects_bad_arguments():
    L = _lib.lib()
    a = 256  # aligned fake pointer: argument checks run before anything is dereferenced
    assert L.rnnt_b200_dw2_gemm(None, a, 128, 100, 512, 0, a, a, None) == -1
    assert L.rnnt_b200_dw2_gemm(a, a, 100, 100, 512, 0, a, a, None) == -2      # rows % 64
    assert L.rnnt_b200_dw2_gemm(a, a, 128, 100, 192, 0, a, a, None) == -2      # H not in {128, 256, 512}
    assert L.rnnt_b200_dw2_gemm(a + 2, a, 128, 100, 512, 0, a, a, None) == -4  # alignment
    assert L.rnnt_b200_dw2_gemm(a, a, 0, 100, 512, 0, a, a, None) == 0         # nothing to do
    assert L.rnnt_b200_dw2_splits(100, 100, 512, 0) == 0 and L.rnnt_b200_dw2_splits(128, 100, 192, 0) == 0
    assert L.rnnt_b200_dw2_splits(64, 100, 512, 0) == 1 and L.rnnt_b200_dw2_splits(64000, 4232, 512, 37) == 2                       # one K block: one split


def test_dh_gemm_and_reduce_reject_bad_arguments():
    L = _lib.lib()
    a = 256
    assert L.rnnt_b200_dh_gemm_supported(512) == 1 and L.rnnt_b200_dh_gemm_supported(192) == 0
    assert L.rnnt_b200_dh_gemm(None, a, a, 100, 512, 4, 0, 4, 0, a, None) == -1
    assert L.rnnt_b200_dh_gemm(a, a, a, 100, 192, 4, 0, 4, 0, a, None) == -2      # H not in {128, 256, 512}
    assert L.rnnt_b200_dh_gemm(a, a, a, 100, 512, 4, 2, 5, 0, a, None) == -2      # tile_hi > tile_cap
    assert L.rnnt_b200_dh_gemm(a + 16, a, a, 100, 512, 4, 0, 4, 0, a, None) == -4  # alignment
    assert L.rnnt_b200_dh_gemm(a, a, a, 100, 512, 4, 2, 2, 0, a, None) == 0       # nothing to do
    assert L.rnnt_b200_dpre_reduce_scratch_elems(32, 500, 101, 512) == 32 * 63 * 101 * 512
    assert L.rnnt_b200_dpre_reduce(a, a, a, 1, 4, 2, 100, a, a, a, None) == -2   # H % 64
    assert L.rnnt_b200_dpre_reduce(a, None, a, 1, 4, 2, 128, a, a, a, None) == -1
    assert L.rnnt_b200_sum_partials(a, 0, 10, a, None) == -2
    assert L.rnnt_b200_build_tiles(a, a, 1, 4, 2, 99, a, a, a, None) == -2       # tile_cap above the padded bound


def test_stored_exponentials_entry_points_reject_bad_arguments():
    # the backward without the second vocabulary GEMM (include/rnnt_b200.h): argument checks come before any launch
    L = _lib.lib()
    a = 256
    fwd = lambda **kw: L.rnnt_b200_joint_fwd_store(
        a, a, a, a, a, a, a, a, a, kw.get("B", 2), 8, 5, kw.get("V", 100), kw.get("H", 512), kw.get("blank", 0),
        kw.get("cap", 2), a, a, a, kw.get("kref", a), kw.get("e", a), None, None)
    assert fwd(kref=None) == -1 and fwd(e=None) == -1
    assert fwd(H=100) == -2 and fwd(cap=999) == -2 and fwd(blank=100) == -3
    assert fwd(e=a + 2) == -4
    ps = lambda **kw: L.rnnt_b200_patch_scale(a, a, a, a, a, a, a, a, a, a, a, a, a, a, 2, 8, 5, 100, kw.get("H", 512), 0,
                                              kw.get("cap", 2), kw.get("lo", 0), kw.get("hi", 2), 0,
                                              kw.get("e", a), kw.get("rs", a), a, None)
    assert ps(e=None) == -1 and ps(rs=None) == -1
    assert ps(H=192) == -2 and ps(H=64) == -2 and ps(cap=0) == -2       # H with tail columns in the stash: unsupported
    assert ps(lo=1, hi=3) == -2 and ps(lo=2, hi=1) == -2 and ps(lo=1, hi=1) == 0   # tile range; empty: nothing to do
    assert L.rnnt_b200_dh_gemm_scaled(a, a, a, None, 100, 512, 4, 0, 4, 0, a, None) == -1
    assert L.rnnt_b200_dh_gemm_scaled(a, a, a, a, 100, 512, 4, 2, 5, 0, a, None) == -2
    assert L.rnnt_b200_dh_gemm_scaled(a, a, a, a, 100, 512, 4, 2, 2, 0, a, None) == 0    # nothing to do
    assert L.rnnt_b200_dw2_gemm_weighted(a, a, None, 128, 100, 512, 0, a, a, None) == -1
    assert L.rnnt_b200_dw2_gemm_weighted(a, a, a, 128, 100, 256, 0, a, a, None) == -2   # H = 512 only
    assert L.rnnt_b200_dw2_gemm_weighted(a, a, a, 100, 100, 512, 0, a, a, None) == -2   # rows % 64


def test_backward_mode_selection(monkeypatch):
    from rnnt_transducer_b200 import _api
    monkeypatch.delenv("RNNT_B200_BWD", raising=False)
    assert _api._store_exponentials(512, 12625, 4288)             # AISHELL step: 13.9 GB of exponentials
    assert not _api._store_exponentials(256, 100, 320)            # kernels of the stored path cover H = 512
    assert not _api._store_exponentials(512, 60000, 4288)         # 65.9 GB: over the default budget -> recompute
    monkeypatch.setenv("RNNT_B200_BWD", "recompute")
    assert not _api._store_exponentials(512, 100, 4288)
    monkeypatch.setenv("RNNT_B200_BWD", "stored")
    assert _api._store_exponentials(512, 60000, 4288) and not _api._store_exponentials(128, 10, 64)


def test_host_tile_count_applies_the_device_clamps():
    from rnnt_transducer_b200._api import _tile_count
    # t = 0 counts as one frame, lengths above the padded dims are cut: same clamps as build_tiles_kernel
    assert _tile_count(([0, 700, 5], [3, 50, 900]), 500, 101) == 1 + (500 * 51 + 127) // 128 + (5 * 101 + 127) // 128
    assert _tile_count(None, 500, 101) is None


def test_transducer_loads_reference_checkpoint_and_matches_its_encoder_decoder_joint():
    # the module tree mirrors rnnt/model.py:40-59, so the reference's own state_dict
    # (fixture made by tests/golden/make_model_golden.py from the unmodified reference) loads strictly
    # and the stock-PyTorch parts reproduce the reference's outputs on CPU
    import torch
    from tests.helpers import load_ref_model_tiny, max_abs
    model, g = load_ref_model_tiny()
    t = lambda k: torch.from_numpy(g[k])
    with torch.no_grad():
        enc, _ = model.encoder(t("inputs"), t("inputs_length"))
        concat = torch.nn.functional.pad(t("targets"), pad=(1, 0, 0, 0), value=0)
        dec, _ = model.decoder(concat, t("targets_length").add(1))
        logits = model.joint(enc, dec)
    assert max_abs(enc, g["enc_state"]) < 1e-5 and max_abs(dec, g["dec_state"]) < 1e-5
    assert max_abs(logits, g["logits"]) < 1e-3 * float(abs(g["logits"]).max())
