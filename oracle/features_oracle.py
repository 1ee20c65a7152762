"""TEST INFRASTRUCTURE, not product code: CPU restatement (numpy) of the reference's feature path for
one utterance -- Kaldi matrix decode, CMVN, context splicing, subsampling, padding -- used by the
tests as the checker of ``rnnt_b200_features_batch``. Only tests/, smoke() and bench.py's CPU legs
may import it.

Pinned: tests/golden/kaldi_tiny.npz holds what the reference's own ``AudioDataset.__getitem__``
(rnnt/dataset.py:129-158) and ``kaldi_io.read_mat`` (rnnt/kaldi_io.py:448-554) returned for the
fixture directories under tests/golden/kaldi_tiny/ (generator: tests/golden/make_features_golden.py);
tests/test_features.py requires this restatement to reproduce them bit for bit.
"""
import numpy as np


def read_matrix(path_offset):
    """``path[:offset]`` -> ndarray, as rnnt/kaldi_io.py:448-492 for binary FM / DM / CM matrices."""
    path, _, off = path_offset.rpartition(":")
    if not path or not off.isdigit():
        path, off = path_offset, "0"
    with open(path, "rb") as fd:
        fd.seek(int(off))
        if fd.read(2) != b"\0B":
            raise ValueError("binary Kaldi matrices only")
        tag = fd.read(3)
        if tag in (b"FM ", b"DM "):  # kaldi_io.py:473-492
            dt = np.dtype("<f4") if tag == b"FM " else np.dtype("<f8")
            hdr = np.frombuffer(fd.read(10), dtype=np.dtype([("a", "i1"), ("r", "<i4"), ("b", "i1"), ("c", "<i4")]))[0]
            rows, cols = int(hdr["r"]), int(hdr["c"])
            return np.frombuffer(fd.read(rows * cols * dt.itemsize), dtype=dt).reshape(rows, cols)
        if tag != b"CM ":
            raise ValueError("unknown matrix header %r" % tag)
        return _read_compressed(fd)


def _read_compressed(fd):
    """Kaldi CompressedMatrix, one-byte format (rnnt/kaldi_io.py:512-554): per column four 16-bit
    percentile codes; a byte d maps linearly onto [p0,p25] for d<=64, [p25,p75] for 65..192,
    [p75,p100] above. All arithmetic in float32; the three masked segments are summed."""
    gmin, grange = np.frombuffer(fd.read(8), dtype="<f4")
    rows, cols = (int(v) for v in np.frombuffer(fd.read(8), dtype="<i4"))
    codes = np.frombuffer(fd.read(8 * cols), dtype="<u2").reshape(cols, 4)
    pct = (codes.astype(np.float32) * grange * np.float32(1.52590218966964e-05) + gmin).astype(np.float32)
    d = np.frombuffer(fd.read(cols * rows), dtype=np.uint8).reshape(cols, rows)  # column-major on disk
    p0, p25, p75, p100 = (pct[:, i:i + 1] for i in range(4))
    low, high = d <= 64, d > 192
    mid = ~(low | high)
    out = np.zeros((cols, rows), np.float32)
    out += (p0 + (p25 - p0) / np.float32(64) * d) * low.astype(np.float32)
    out += (p25 + (p75 - p25) / np.float32(128) * (d - np.uint8(64))) * mid.astype(np.float32)
    out += (p75 + (p100 - p75) / np.float32(63) * (d - np.uint8(192))) * high.astype(np.float32)
    return out.T


def cmvn(mat, stats):
    """rnnt/dataset.py:72-75."""
    count = stats[0, -1]
    mean = stats[0, :-1] / count
    var = stats[1, :-1] / count - mean * mean
    return (mat - mean) / np.sqrt(var)


def splice(feats, left, right):
    """rnnt/dataset.py:77-98: blocks [t-left .. t-1 | t | ...], zeros where the frame does not
    exist. The right-context blocks are addressed with ``right`` (not ``left``) as in the
    reference, and written last."""
    T, F = feats.shape
    out = np.zeros((T, F * (1 + left + right)), np.float32)
    out[:, left * F:(left + 1) * F] = feats
    for i in range(left):
        blk = left - i - 1
        out[i + 1:, blk * F:(blk + 1) * F] = feats[:T - i - 1]
    for i in range(right):
        blk = right + i + 1
        out[:T - i - 1, blk * F:(blk + 1) * F] = feats[i + 1:]
    return out


def subsample(feats, frame_rate):
    """rnnt/dataset.py:100-109: every ``int(frame_rate/10)``-th frame unless frame_rate == 10."""
    if frame_rate == 10:
        return feats
    return feats[::int(frame_rate / 10)]


def pad_rows(feats, length):
    """rnnt/dataset.py:38-56 (2-D case): zero rows up to ``length``; longer input is an error."""
    if feats.shape[0] > length:
        raise ValueError("longer than the padded length")
    out = np.zeros((length, feats.shape[1]), np.float32)
    out[:feats.shape[0]] = feats
    return out


def utterance_features(path_offset, stats, left, right, frame_rate, pad_to):
    """One item of rnnt/dataset.py:129-158 -> (features (pad_to, F') float32, length)."""
    x = read_matrix(path_offset)
    if stats is not None:
        x = cmvn(x, stats)
    x = subsample(splice(x, left, right), frame_rate)
    return pad_rows(x, pad_to), x.shape[0]
