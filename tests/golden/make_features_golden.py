"""Generates tests/golden/kaldi_tiny/ (Kaldi data directories) and tests/golden/kaldi_tiny.npz
IN THE BUILD CONTAINER (needs /root/reference).

Run:  python tests/golden/make_features_golden.py
The matrices are written with the reference's own ``kaldi_io.write_mat`` (rnnt/kaldi_io.py:557-601;
the compressed variant with a small compressor below, since the reference only READS that format)
and the fixture records what the UNMODIFIED reference returns for them:
  * ``kaldi_io.read_mat`` of every feature matrix (float32, float64, compressed),
  * ``AudioDataset(config, 'train')[i]`` (rnnt/dataset.py:129-158) for every item of every variant:
    CMVN + splice + subsample + pad, targets, lengths, and the item order.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "kaldi_tiny")
sys.modules["editdistance"] = types.ModuleType("editdistance")  # imported by rnnt/utils.py:3, unused here
sys.path.insert(0, "/root/reference")
sys.path.insert(0, HERE)
import rnnt.kaldi_io as kaldi_io  # noqa: E402
from rnnt.dataset import AudioDataset  # noqa: E402
from rnnt.utils import AttrDict  # noqa: E402

from make_features_golden_cfg import F, UTTS, TEXT, VOCAB, VARIANTS, compress, data_config  # noqa: E402


def write_dir(name, fmt, stats_dtype, rng):
    d = os.path.join(ROOT, name, "train")
    os.makedirs(d, exist_ok=True)
    ark, scp = os.path.join(d, "feats.ark"), []
    feats = {}
    with open(ark, "wb") as fd:
        for key, spk, rows in UTTS:
            m = (rng.standard_normal((rows, F)) * 3.0 + np.linspace(-2, 5, F)).astype(np.float32)
            feats[key] = m
            off = fd.tell() + len(key) + 1
            if fmt == "cm":
                fd.write((key + " ").encode() + compress(m))
            else:
                kaldi_io.write_mat(fd, m.astype(np.float64) if fmt == "f64" else m, key=key)
            scp.append("%s %s/train/feats.ark:%d" % (key, name, off))
    with open(os.path.join(d, "feats.scp"), "w") as f:
        f.write("\n".join(scp) + "\n")
    with open(os.path.join(d, "utt2spk"), "w") as f:
        f.write("".join("%s %s\n" % (k, s) for k, s, _ in UTTS))
    with open(os.path.join(d, "text"), "w") as f:
        f.write("".join("%s %s\n" % (k, TEXT[k]) for k, _, _ in UTTS))
    if stats_dtype is not None:
        cark, cscp = os.path.join(d, "cmvn.ark"), []
        with open(cark, "wb") as fd:
            for spk in ("spkA", "spkB"):
                rows = np.concatenate([feats[k] for k, s, _ in UTTS if s == spk]).astype(np.float64)
                st = np.zeros((2, F + 1))
                st[0, :-1], st[1, :-1], st[0, -1] = rows.sum(0), (rows * rows).sum(0), rows.shape[0]
                off = fd.tell() + len(spk) + 1
                kaldi_io.write_mat(fd, st.astype(np.float32 if stats_dtype == "f32" else np.float64), key=spk)
                cscp.append("%s %s/train/cmvn.ark:%d" % (spk, name, off))
        with open(os.path.join(d, "cmvn.scp"), "w") as f:
            f.write("\n".join(cscp) + "\n")


def main():
    os.makedirs(ROOT, exist_ok=True)
    with open(os.path.join(ROOT, "vocab"), "w") as f:
        f.write("\n".join(VOCAB) + "\n")
    rng = np.random.default_rng(2019)
    out = {}
    os.chdir(ROOT)  # the scp entries are relative to the fixture root
    for name, (fmt, stats_dtype, over) in VARIANTS.items():
        write_dir(name, fmt, stats_dtype, rng)
        cfg = data_config(name, stats_dtype, over)
        ds = AudioDataset(AttrDict(cfg), "train")
        out[name + "/len"] = np.int64(len(ds))
        order = []
        for i in range(len(ds)):
            feats, n, tgt, tn = ds[i]
            utt = ds.sorted_list[i][0] if ds.sorted_list is not None else ds.feats_list[i]
            order.append(utt)
            out["%s/%d/inputs" % (name, i)] = feats
            out["%s/%d/inputs_length" % (name, i)] = n
            out["%s/%d/targets" % (name, i)] = tgt
            out["%s/%d/targets_length" % (name, i)] = tn
        out[name + "/order"] = np.array(order)
        for key, path in ds.feats_dict.items():
            out["%s/mat/%s" % (name, key)] = np.array(kaldi_io.read_mat(path))
        assert feats.dtype == np.float32 and feats.shape[0] == cfg["max_input_length"]
    # right context wider than the left one: the reference's slice assignment fails
    try:
        ds.left_context_width, ds.right_context_width = 0, 1
        ds.concat_frame(np.zeros((4, F), np.float32))
        out["right_gt_left_raises"] = np.int64(0)
    except ValueError:
        out["right_gt_left_raises"] = np.int64(1)
    np.savez_compressed(os.path.join(HERE, "kaldi_tiny.npz"), **out)
    print("wrote", len(out), "arrays;", {k: int(v) for k, v in out.items() if k.endswith("/len")},
          "right>left raises:", int(out["right_gt_left_raises"]))


if __name__ == "__main__":
    main()
