"""Shared between tests/golden/make_features_golden.py (the generator, needs /root/reference) and
tests/test_features.py (which must not): the fixture's utterances, transcripts, variants, the data
section of the config each variant is read with, and a small Kaldi CompressedMatrix encoder."""
import numpy as np

F = 8
UTTS = [("spkA-u1", "spkA", 37), ("spkB-u1", "spkB", 5), ("spkA-u2", "spkA", 61), ("spkB-u2", "spkB", 1),
        ("spkA-u3", "spkA", 24), ("spkB-u3", "spkB", 48), ("spkA-u4", "spkA", 9)]
TEXT = {"spkA-u1": "a b c", "spkB-u1": "d", "spkA-u2": "a a b zz c d e", "spkB-u2": "b",
        "spkA-u3": "c d", "spkB-u3": "e d c b a", "spkA-u4": "a b c d e a b c d e a b c"}  # last: too long, dropped
VOCAB = ["<blk> 0", "<unk> 1", "a 2", "b 3", "c 4", "d 5", "e 6"]

# variant -> (feature dtype/format, stats dtype, data-section overrides)
VARIANTS = {
    "fm": ("f32", "f64", dict(left_context_width=3, right_context_width=0, frame_rate=30)),
    "dm": ("f64", "f64", dict(left_context_width=2, right_context_width=1, frame_rate=20)),
    "cm": ("cm", "f64", dict(left_context_width=3, right_context_width=0, frame_rate=30)),
    "fm_f32stats": ("f32", "f32", dict(left_context_width=2, right_context_width=2, frame_rate=10)),
    "fm_nocmvn": ("f32", None, dict(left_context_width=1, right_context_width=0, frame_rate=40, short_first=False)),
}


def compress(mat):
    """Kaldi one-byte CompressedMatrix bytes of a float matrix (enough of the encoder to give the
    reference's reader realistic input; the decoded values are whatever the READER says)."""
    gmin, gmax = np.float32(mat.min()), np.float32(mat.max())
    grange = np.float32(gmax - gmin)
    rows, cols = mat.shape
    out = b"\0B" + b"CM " + np.array([gmin, grange], "<f4").tobytes() + np.array([rows, cols], "<i4").tobytes()
    hdrs, data = [], []
    for c in range(cols):
        col = np.sort(mat[:, c])
        q = [col[0], col[rows // 4], col[(3 * rows) // 4], col[-1]]
        code = [int(round(float((v - gmin) / grange) * 65535.0)) for v in q]
        for k in range(1, 4):  # strictly increasing percentile codes
            code[k] = max(code[k], code[k - 1] + 1)
        code = [min(c_, 65535) for c_ in code]
        hdrs.append(np.array(code, "<u2").tobytes())
        p = [gmin + grange * np.float32(1.52590218966964e-05) * c_ for c_ in code]
        x = mat[:, c].astype(np.float64)
        b = np.where(x < p[1], (x - p[0]) / (p[1] - p[0]) * 64.0 + 0.5,
                     np.where(x < p[2], 64 + (x - p[1]) / (p[2] - p[1]) * 128.0 + 0.5,
                              192 + (x - p[2]) / (p[3] - p[2]) * 63.0 + 0.5))
        data.append(np.clip(np.floor(b), 0, 255).astype(np.uint8).tobytes())
    return out + b"".join(hdrs) + b"".join(data)


def data_config(name, stats_dtype, over):
    cfg = dict(name="kaldi_tiny", vocab="vocab", left_context_width=0, right_context_width=0, frame_rate=10,
               apply_cmvn=stats_dtype is not None, max_input_length=64, max_target_length=8, batch_size=3,
               encoding=True, short_first=True, shuffle=False, text_flag="text", train="%s/train" % name)
    cfg.update(over)
    return cfg
