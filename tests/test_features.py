"""Data path ahead of the hot path. CPU part: the numpy oracle
reproduces, bit for bit, what the reference's own ``kaldi_io.read_mat`` and ``AudioDataset``
returned for the fixture directories (tests/golden/kaldi_tiny*, generated from /root/reference by
tests/golden/make_features_golden.py), and the host logic (scp/ark parsing, filtering, ordering,
targets, splice table) agrees with them. GPU part: ``rnnt_b200_features_batch`` through
``KaldiFeatureSet.batch`` against the same golden arrays and, at AISHELL sizes, the oracle."""
import os
import sys

import numpy as np
import pytest
import torch

import rnnt_transducer_b200 as R
from rnnt_transducer_b200 import _data

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import features_oracle as FO  # noqa: E402  (checker only)
from tests.golden.make_features_golden_cfg import VARIANTS, data_config  # noqa: E402

FIX = os.path.join(ROOT, "tests", "golden", "kaldi_tiny")
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "kaldi_tiny.npz"))


def _same_bits(a, b):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and a.dtype == b.dtype and a.tobytes() == b.tobytes()


def _scp(name):
    with open(os.path.join(FIX, name, "train", "feats.scp")) as f:
        return dict(line.split() for line in f if line.strip())


@pytest.mark.parametrize("name", list(VARIANTS))
def test_oracle_matrix_decode_matches_reference_reader(name):
    for key, rx in _scp(name).items():
        got = FO.read_matrix(os.path.join(FIX, rx))
        assert _same_bits(got, GOLD[f"{name}/mat/{key}"]), (name, key)


@pytest.mark.parametrize("name", list(VARIANTS))
def test_oracle_items_match_reference_dataset(name):
    fmt, stats_dtype, over = VARIANTS[name]
    cfg = data_config(name, stats_dtype, over)
    scp = _scp(name)
    stats = {}
    if stats_dtype is not None:
        with open(os.path.join(FIX, name, "train", "cmvn.scp")) as f:
            stats = {k: FO.read_matrix(os.path.join(FIX, rx)) for k, rx in (l.split() for l in f if l.strip())}
    for i, utt in enumerate(GOLD[f"{name}/order"]):
        st = stats[utt.split("-")[0]] if stats else None
        feats, n = FO.utterance_features(os.path.join(FIX, scp[utt]), st, cfg["left_context_width"],
                                         cfg["right_context_width"], cfg["frame_rate"], cfg["max_input_length"])
        assert n == int(GOLD[f"{name}/{i}/inputs_length"])
        assert _same_bits(feats, GOLD[f"{name}/{i}/inputs"]), (name, i, utt)


@pytest.mark.parametrize("name", list(VARIANTS))
def test_ark_table_and_host_decode(name):
    table = _data.ArkTable(os.path.join(FIX, name, "train", "feats.scp"), root=FIX)
    assert len(table) == 7
    for key in table.keys:
        m = table[key]
        ref = GOLD[f"{name}/mat/{key}"]
        assert (m.rows, m.cols) == ref.shape
        assert _same_bits(np.array(m.to_numpy()), ref), (name, key)


@pytest.mark.parametrize("name", list(VARIANTS))
def test_feature_set_filtering_order_and_targets(name):
    fmt, stats_dtype, over = VARIANTS[name]
    ds = R.KaldiFeatureSet(data_config(name, stats_dtype, over), "train", root=FIX)
    assert len(ds) == int(GOLD[f"{name}/len"]) == 6  # the 13-label utterance is dropped (dataset.py:184)
    assert ds.utt_ids[:len(ds)] == list(GOLD[f"{name}/order"])
    for i in range(len(ds)):
        t = ds.targets_dict[ds.utt_ids[i]]
        assert len(t) == int(GOLD[f"{name}/{i}/targets_length"])
        assert t == GOLD[f"{name}/{i}/targets"][:len(t)].tolist()  # '<unk>' for the unknown unit included
    if stats_dtype is not None:
        assert ds.cmvn_mean.dtype == (np.float32 if stats_dtype == "f32" else np.float64)


def test_splice_table_follows_concat_frame():
    assert _data.splice_table(3, 0) == ([-3, -2, -1, 0], [_data.NO_SHIFT] * 4)
    prim, over = _data.splice_table(2, 2)
    assert prim == [-2, -1, 0, _data.NO_SHIFT, _data.NO_SHIFT] and over[3:] == [1, 2]
    prim, over = _data.splice_table(2, 1)  # the right block lands on the centre block (dataset.py:92-96)
    assert prim == [-2, -1, 0, _data.NO_SHIFT] and over == [_data.NO_SHIFT, _data.NO_SHIFT, 1, _data.NO_SHIFT]
    assert int(GOLD["right_gt_left_raises"]) == 1
    with pytest.raises(ValueError):
        _data.splice_table(0, 1)


def test_cm_column_params_and_desc_layout():
    hdr = np.array([[0, 100, 40000, 65535], [5, 6, 7, 8]], np.uint16)
    cp = _data.cm_column_params(hdr, -3.5, 11.25)
    assert cp.shape == (2, 6) and cp.dtype == np.float32
    p = hdr.astype(np.float32) * np.float32(11.25) * np.float32(1.52590218966964e-05) + np.float32(-3.5)
    assert _same_bits(cp[:, 0], p[:, 0]) and _same_bits(cp[:, 5], (p[:, 3] - p[:, 2]) / np.float32(63))
    assert _data.UTT_DESC_DTYPE.itemsize == 32 and _data.UTT_DESC_DTYPE.fields["rows"][1] == 16


def test_features_batch_refuses_without_cuda():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    ds = R.KaldiFeatureSet(data_config("fm", "f64", VARIANTS["fm"][2]), "train", root=FIX)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ds.batch([0, 1])


# ----------------------------------------------------------------------------- GPU parity
@pytest.mark.gpu
@pytest.mark.parametrize("name", list(VARIANTS))
def test_gpu_batches_match_reference_dataset(name):
    fmt, stats_dtype, over = VARIANTS[name]
    cfg = data_config(name, stats_dtype, over)
    ds = R.KaldiFeatureSet(cfg, "train", root=FIX)
    # the reference's padded layout, whole set in one batch
    x, xl, y, yl = ds.batch(range(len(ds)), pad_to_max=True)
    assert x.is_cuda and y.is_cuda and not xl.is_cuda and xl.dtype == torch.int64 and y.dtype == torch.int64
    xc, yc = x.cpu().numpy(), y.cpu().numpy()
    for i in range(len(ds)):
        assert int(xl[i]) == int(GOLD[f"{name}/{i}/inputs_length"]) and int(yl[i]) == int(GOLD[f"{name}/{i}/targets_length"])
        assert _same_bits(xc[i], GOLD[f"{name}/{i}/inputs"]), (name, i)
        assert np.array_equal(yc[i], GOLD[f"{name}/{i}/targets"])
    # loader order and batch-trimmed layout = what train.py:30-33 leaves
    seen = 0
    for x, xl, y, yl in ds.loader(batch_size=cfg["batch_size"]):
        assert x.shape[1] == int(xl.max()) and y.shape[1] == max(1, int(yl.max()))
        for b in range(x.shape[0]):
            g = GOLD[f"{name}/{seen}/inputs"]
            assert _same_bits(x[b].cpu().numpy(), g[:x.shape[1]]) and not g[x.shape[1]:].any()
            seen += 1
    assert seen == len(ds)


def _write_ark(path, mats, fmt):
    from tests.golden.make_features_golden_cfg import compress
    scp = []
    with open(path, "wb") as fd:
        for k, m in mats.items():
            fd.write((k + " ").encode())
            off = fd.tell()
            if fmt == "cm":
                fd.write(compress(m))
            else:
                dt = "<f4" if fmt == "f32" else "<f8"
                fd.write(b"\0B" + (b"FM " if fmt == "f32" else b"DM ") + b"\x04" + np.int32(m.shape[0]).tobytes()
                         + b"\x04" + np.int32(m.shape[1]).tobytes() + m.astype(dt).tobytes())
            scp.append(f"{k} {path}:{off}")
    with open(path[:-4] + ".scp", "w") as f:
        f.write("\n".join(scp) + "\n")
    return path[:-4] + ".scp"


@pytest.mark.gpu
@pytest.mark.parametrize("fmt,F,left,right,rate", [("f32", 40, 3, 0, 30), ("cm", 40, 3, 0, 30), ("f64", 13, 2, 2, 20),
                                                  ("f32", 7, 0, 0, 10)])
def test_gpu_features_match_oracle_at_aishell_sizes(tmp_path, fmt, F, left, right, rate):
    rng = np.random.default_rng(7)
    rows = [1500, 1, 2, 3, 700, 1499, 64, 65] + [int(r) for r in rng.integers(4, 1500, 24)]
    mats = {f"u{i:02d}": (rng.standard_normal((r, F)) * 4 + rng.standard_normal(F)).astype(np.float32)
            for i, r in enumerate(rows)}
    scp = _write_ark(str(tmp_path / "feats.ark"), mats, fmt)
    table = _data.ArkTable(scp)
    stats = []
    for s in range(3):
        st = np.zeros((2, F + 1))
        x = rng.standard_normal((500, F)) * 4
        st[0, :-1], st[1, :-1], st[0, -1] = x.sum(0), (x * x).sum(0), 500
        stats.append(st)
    mean, sd = _data.cmvn_tables(stats)
    spk = [i % 3 for i in range(len(rows))]
    out, lens = _data.features_batch([table[k] for k in table.keys], spk, torch.from_numpy(mean).cuda(),
                                     torch.from_numpy(sd).cuda(), left, right, int(rate / 10))
    out = out.cpu().numpy()
    with open(scp) as f:
        rx = dict(l.split() for l in f)
    for b, k in enumerate(table.keys):
        want, n = FO.utterance_features(rx[k], stats[spk[b]], left, right, rate, out.shape[1])
        assert n == int(lens[b])
        assert _same_bits(out[b], want), (fmt, k, rows[b])


def test_loader_sharding_covers_every_utterance_once():
    """rank r of world_size N takes utterances r::N of every global batch (train.py:136 sizes the
    global batch as batch_size x num_gpu); checked on the index level, no device needed."""
    ds = R.KaldiFeatureSet(data_config("fm", "f64", VARIANTS["fm"][2]), "train", root=FIX)
    ds.batch = lambda idx, device=None, pad_to_max=False: list(idx)
    one = [b for b in ds.loader(batch_size=4)]
    assert one == [[0, 1, 2, 3], [4, 5]]
    parts = [[b for b in ds.loader(batch_size=2, rank=r, world_size=2)] for r in range(2)]
    assert parts[0] == [[0, 2], [4]] and parts[1] == [[1, 3], [5]]
    g = torch.Generator().manual_seed(3)
    shuffled = [i for b in ds.loader(batch_size=4, shuffle=True, generator=g) for i in b]
    assert sorted(shuffled) == list(range(6))
