"""Data path feeding the hot path: Kaldi ``feats.scp`` / ``cmvn.scp``
/ ``utt2spk`` / ``text`` directories as the reference's ``AudioDataset`` reads them
(rnnt/dataset.py:8-207), with the per-utterance numpy work moved onto the GPU:

* ark files are memory-mapped once; an utterance is a zero-copy view of its matrix payload
  (the reference opens, seeks and reads the file for every item, rnnt/kaldi_io.py:58-92,448-492);
* a batch is ONE pinned staging buffer (descriptors + raw payload bytes, float32 / float64 /
  Kaldi-compressed), ONE host-to-device copy and ONE launch of ``rnnt_b200_features_batch``
  (csrc/features.cu) that decompresses, applies CMVN, splices context frames, subsamples and
  zero-pads -- bit-identical to dataset.py:72-109 + :38-56;
* batches are padded to the longest utterance of the batch, which is what train.py:30-33 trims
  them to anyway (``pad_to_max=True`` gives the reference's ``max_input_length`` layout), and the
  lengths stay on the host, so the step needs no ``.max().item()`` read-back.

Host side is Python because the reference's is; there is no CPU fallback for the feature kernel.
"""
from __future__ import annotations

import ctypes
import mmap
import os
import re
from typing import Dict, Iterable, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib

__all__ = ["ArkTable", "ArkMatrix", "KaldiFeatureSet", "splice_table", "cmvn_tables", "cm_column_params",
           "features_batch", "UTT_DESC_DTYPE"]

FEAT_F32, FEAT_F64, FEAT_CM8 = 0, 1, 2
NO_SHIFT = -2 ** 31
# rnnt_b200_utt_desc of include/rnnt_b200.h
UTT_DESC_DTYPE = np.dtype([("payload_off", "<i8"), ("col_off", "<i8"), ("rows", "<i4"), ("format", "<i4"),
                           ("out_rows", "<i4"), ("spk", "<i4")])
assert UTT_DESC_DTYPE.itemsize == 32


class ArkMatrix:
    """One matrix inside a memory-mapped ark file: ``payload`` is a uint8 view of the matrix data
    (no copy); ``col_header`` (cols,4) uint16 and ``gmin``/``grange`` exist for compressed ones."""
    __slots__ = ("format", "rows", "cols", "payload", "col_header", "gmin", "grange")

    def __init__(self, fmt, rows, cols, payload, col_header=None, gmin=None, grange=None):
        self.format, self.rows, self.cols, self.payload = fmt, int(rows), int(cols), payload
        self.col_header, self.gmin, self.grange = col_header, gmin, grange

    def to_numpy(self) -> np.ndarray:
        """Decoded matrix on the host (statistics matrices, tests); features go to the GPU raw."""
        if self.format == FEAT_F32:
            return self.payload.view("<f4").reshape(self.rows, self.cols)
        if self.format == FEAT_F64:
            return self.payload.view("<f8").reshape(self.rows, self.cols)
        cp = cm_column_params(self.col_header, self.gmin, self.grange)
        d = self.payload.reshape(self.cols, self.rows)
        k = d.astype(np.float32)
        lo = cp[:, 0:1] + cp[:, 1:2] * k
        mid = cp[:, 2:3] + cp[:, 3:4] * (k - np.float32(64))
        hi = cp[:, 4:5] + cp[:, 5:6] * (k - np.float32(192))
        return (np.where(d <= 64, lo, np.where(d <= 192, mid, hi)) + np.float32(0)).T


def cm_column_params(col_header: np.ndarray, gmin, grange) -> np.ndarray:
    """(cols,6) float32 {p0, s0, p25, s1, p75, s2} of a Kaldi compressed matrix: percentile values
    ``u16 * range * 2^-16-ish + min`` and the slopes of the three segments, each step rounded to
    float32 as rnnt/kaldi_io.py:531-549 evaluates them."""
    gmin, grange = np.float32(gmin), np.float32(grange)
    p = col_header.astype(np.float32) * grange * np.float32(1.52590218966964e-05) + gmin
    out = np.empty((p.shape[0], 6), np.float32)
    out[:, 0], out[:, 2], out[:, 4] = p[:, 0], p[:, 1], p[:, 2]
    out[:, 1] = (p[:, 1] - p[:, 0]) / np.float32(64)
    out[:, 3] = (p[:, 2] - p[:, 1]) / np.float32(128)
    out[:, 5] = (p[:, 3] - p[:, 2]) / np.float32(63)
    return out


class ArkTable:
    """``key -> matrix`` over a Kaldi scp file (lines ``key path[:offset]``, optional ``ark:``
    prefix; rnnt/dataset.py:58-66, rnnt/kaldi_io.py:58-92). Binary ``FM``/``DM``/``CM`` matrices;
    text-mode matrices, gzip and pipes (kaldi_io.py:71-79,495-509) are refused loudly."""

    _RX_PREFIX = re.compile(r"^(ark|scp)(,scp|,b|,t|,n?f|,n?p|,b?o|,n?s|,n?cs)*:")

    def __init__(self, scp_path: str, root: Optional[str] = None):
        self.root = root if root is not None else os.path.dirname(os.path.abspath(scp_path))
        self.keys: List[str] = []
        self._where: Dict[str, Tuple[str, int]] = {}
        self._maps: Dict[str, np.ndarray] = {}
        with open(scp_path, "r") as fid:
            for line in fid:
                line = line.strip()
                if not line:
                    continue
                key, rx = line.split(" ", 1)
                rx = self._RX_PREFIX.sub("", rx.strip())
                if rx.endswith("|") or rx.startswith("|") or rx.endswith(".gz"):
                    raise NotImplementedError(f"{scp_path}: pipe / gzip rxfilename '{rx}' is not supported")
                m = re.search(r":([0-9]+)$", rx)
                off = int(m.group(1)) if m else 0
                path = rx[:m.start()] if m else rx
                self.keys.append(key)
                self._where[key] = (path, off)

    def __len__(self):
        return len(self.keys)

    def __contains__(self, key):
        return key in self._where

    def _bytes(self, path: str) -> np.ndarray:
        buf = self._maps.get(path)
        if buf is None:
            full = path if os.path.isabs(path) or os.path.exists(path) else os.path.join(self.root, path)
            with open(full, "rb") as f:
                mm = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ)
            buf = np.frombuffer(mm, dtype=np.uint8)
            self._maps[path] = buf
        return buf

    def __getitem__(self, key: str) -> ArkMatrix:
        path, off = self._where[key]
        buf = self._bytes(path)
        if bytes(buf[off:off + 2]) != b"\0B":
            raise NotImplementedError(f"{path}:{off}: not a binary Kaldi matrix (text mode is not supported)")
        tag = bytes(buf[off + 2:off + 5])
        if tag in (b"FM ", b"DM "):
            if buf[off + 5] != 4 or buf[off + 10] != 4:
                raise ValueError(f"{path}:{off}: bad dimension markers")
            rows, cols = (int(buf[off + 6:off + 10].view("<i4")[0]), int(buf[off + 11:off + 15].view("<i4")[0]))
            size = 4 if tag == b"FM " else 8
            start = off + 15
            return ArkMatrix(FEAT_F32 if size == 4 else FEAT_F64, rows, cols, buf[start:start + rows * cols * size])
        if tag == b"CM ":
            gmin, grange = buf[off + 5:off + 13].view("<f4")
            rows, cols = (int(v) for v in buf[off + 13:off + 21].view("<i4"))
            hdr = buf[off + 21:off + 21 + 8 * cols].view("<u2").reshape(cols, 4)
            start = off + 21 + 8 * cols
            return ArkMatrix(FEAT_CM8, rows, cols, buf[start:start + rows * cols], hdr, gmin, grange)
        raise ValueError(f"{path}:{off}: unknown matrix header {tag!r}")  # kaldi_io.py:477-478 (CM2/CM3 too)


def cmvn_tables(stats: Sequence[np.ndarray]) -> Tuple[np.ndarray, np.ndarray]:
    """Per-speaker mean and sqrt(variance) rows from Kaldi CMVN accumulators (2, F+1): the first
    two lines of rnnt/dataset.py:72-75, evaluated in the statistics' own dtype."""
    mean = [s[0, :-1] / s[0, -1] for s in stats]
    sd = [np.sqrt(s[1, :-1] / s[0, -1] - np.square(m)) for s, m in zip(stats, mean)]
    return np.stack(mean), np.stack(sd)


def splice_table(left: int, right: int) -> Tuple[List[int], List[int]]:
    """Frame shift written into each of the ``1+left+right`` output blocks by ``concat_frame``
    (rnnt/dataset.py:77-98): centre block ``left``; left-context ``i`` (shift ``-(i+1)``) in block
    ``left-i-1``; right-context ``i`` (shift ``+(i+1)``) in block ``right+i+1`` -- the reference
    indexes the right blocks by ``right_context_width``, written last, so for ``right < left`` they
    land on (and override, where the source frame exists) earlier blocks. ``right > left`` makes
    the reference's slice assignment fail; it is refused here too."""
    if left < 0 or right < 0:
        raise ValueError("context widths must be >= 0")
    if right > left:
        raise ValueError("right_context_width > left_context_width: the reference's concat_frame "
                         "(rnnt/dataset.py:92-96) cannot place those blocks")
    n = 1 + left + right
    prim, over = [NO_SHIFT] * n, [NO_SHIFT] * n
    prim[left] = 0
    for i in range(left):
        prim[left - i - 1] = -(i + 1)
    for i in range(right):
        over[right + i + 1] = i + 1
    return prim, over


class _Staging:
    """Two pinned host buffers used in turn, each guarded by the event of the copy that last read
    it, so filling batch k+1 overlaps the transfer and the kernel of batch k."""

    def __init__(self):
        self.bufs: List[Optional[torch.Tensor]] = [None, None]
        self.events: List[Optional[torch.cuda.Event]] = [None, None]
        self.turn = 0

    def take(self, nbytes: int) -> Tuple[torch.Tensor, int]:
        i = self.turn
        self.turn ^= 1
        if self.events[i] is not None:
            self.events[i].synchronize()
        if self.bufs[i] is None or self.bufs[i].numel() < nbytes:
            self.bufs[i] = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, pin_memory=True)
        return self.bufs[i], i

    def release(self, i: int):
        ev = torch.cuda.Event()
        ev.record()
        self.events[i] = ev


def _align(n: int, a: int = 16) -> int:
    return (n + a - 1) // a * a


def features_batch(mats: Sequence[ArkMatrix], spk: Sequence[int], cmvn_mean: Optional[torch.Tensor],
                   cmvn_sd: Optional[torch.Tensor], left: int, right: int, interval: int,
                   pad_to: Optional[int] = None, device=None, staging: Optional[_Staging] = None):
    """Raw ark matrices of one batch -> ``(inputs (B, t_pad, (1+left+right)*F) float32 on the
    device, lengths (B) int64 on the host)``. ``spk[b]`` is the row of the CMVN tables (device
    tensors, float64 or float32) or -1. One H2D copy, one kernel launch, nothing read back."""
    if not torch.cuda.is_available():
        raise RuntimeError("rnnt_b200: features_batch needs a CUDA device (there is no CPU fallback)")
    B = len(mats)
    if B == 0:
        raise ValueError("empty batch")
    F_ = mats[0].cols
    if any(m.cols != F_ for m in mats):
        raise ValueError("all utterances of a batch must have the same feature dimension")
    prim, over = splice_table(left, right)
    nblk = len(prim)
    if nblk > 32:
        raise ValueError("at most 32 spliced blocks")
    if interval < 1:
        raise ValueError("subsampling interval must be >= 1")
    use_cmvn = cmvn_mean is not None and any(s >= 0 for s in spk)
    cmvn_f64 = bool(use_cmvn and cmvn_mean.dtype == torch.float64)
    if use_cmvn and not cmvn_f64 and any(m.format == FEAT_F64 for m in mats):
        raise ValueError("float64 features with float32 CMVN statistics: convert the tables to float64 "
                         "(numpy promotes that product to float64)")
    out_rows = [-(-m.rows // interval) for m in mats]
    t_pad = max(out_rows) if pad_to is None else int(pad_to)
    if max(out_rows) > t_pad:
        raise ValueError(f"utterance of {max(out_rows)} frames exceeds max_input_length {t_pad}")  # dataset.py:50
    # staging layout: descriptors | CM column parameters | payloads (each 16-byte aligned)
    desc = np.zeros(B, UTT_DESC_DTYPE)
    col_base = _align(B * UTT_DESC_DTYPE.itemsize)
    n_col = sum(6 * m.cols for m in mats if m.format == FEAT_CM8)
    cursor = _align(col_base + 4 * n_col)
    col_cursor = 0
    for b, m in enumerate(mats):
        desc[b] = (cursor, col_cursor, m.rows, m.format, out_rows[b], spk[b] if use_cmvn else -1)
        cursor = _align(cursor + m.payload.size)
        if m.format == FEAT_CM8:
            col_cursor += 6 * m.cols
    total = cursor
    staging = staging if staging is not None else _DEFAULT_STAGING
    host, slot = staging.take(total)
    hv = host.numpy()
    hv[:B * UTT_DESC_DTYPE.itemsize] = desc.view(np.uint8)
    for b, m in enumerate(mats):
        o = int(desc["payload_off"][b])
        hv[o:o + m.payload.size] = m.payload
        if m.format == FEAT_CM8:
            c = col_base + 4 * int(desc["col_off"][b])
            hv[c:c + 24 * m.cols] = cm_column_params(m.col_header, m.gmin, m.grange).reshape(-1).view(np.uint8)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    with torch.cuda.device(dev):
        blob = torch.empty(total, dtype=torch.uint8, device=dev)
        blob.copy_(host[:total], non_blocking=True)
        staging.release(slot)
        out = torch.empty((B, t_pad, nblk * F_), dtype=torch.float32, device=dev)
        arr = (ctypes.c_int * nblk)
        base = blob.data_ptr()
        _lib.check(_lib.lib().rnnt_b200_features_batch(
            base, base, base + col_base, cmvn_mean.data_ptr() if use_cmvn else None,
            cmvn_sd.data_ptr() if use_cmvn else None, int(cmvn_f64), B, F_, nblk, arr(*prim), arr(*over),
            interval, t_pad, out.data_ptr(), torch.cuda.current_stream().cuda_stream), "rnnt_b200_features_batch")
    return out, torch.tensor(out_rows, dtype=torch.int64)


_DEFAULT_STAGING = _Staging()


def _cfg(config, name, default=None):
    if isinstance(config, dict):
        return config.get(name, default)
    return getattr(config, name, default)


class KaldiFeatureSet:
    """The reference's ``AudioDataset(config.data, type)`` (rnnt/dataset.py:112-207) as a batch
    source: same directory layout (``<dir>/feats.scp``, ``utt2spk``, ``cmvn.scp``, ``<text_flag>``,
    vocabulary file), same filtering (utterances without a transcript or with more than
    ``max_target_length`` labels are dropped, :177-203) and ordering (``short_first`` sorts the
    training set by label count, :124-127). ``config`` is the ``data`` section of the YAML as a
    dict or an attribute object. ``batch(indices)`` returns what the DataLoader yields after
    train.py:26-33 -- ``inputs`` on the device and trimmed to the batch, ``targets`` on the device,
    both length vectors int64 on the host -- with the feature work done by one kernel launch."""

    def __init__(self, config, type: str, root: Optional[str] = None):
        self.type = type
        self.left = int(_cfg(config, "left_context_width", 0))
        self.right = int(_cfg(config, "right_context_width", 0))
        frame_rate = _cfg(config, "frame_rate", 10)
        self.interval = int(frame_rate / 10) if frame_rate != 10 else 1  # dataset.py:101-102
        self.apply_cmvn = bool(_cfg(config, "apply_cmvn", False))
        self.max_input_length = int(_cfg(config, "max_input_length"))
        self.max_target_length = int(_cfg(config, "max_target_length"))
        self.dir = _cfg(config, type)
        if root is not None and not os.path.isabs(self.dir):
            self.dir = os.path.join(root, self.dir)
        self.feats = ArkTable(os.path.join(self.dir, "feats.scp"), root=root)
        self._cmvn_dev: Dict[torch.device, Tuple[torch.Tensor, torch.Tensor]] = {}
        self.spk_index: Dict[str, int] = {}
        self.utt2spk: Dict[str, str] = {}
        if self.apply_cmvn:
            with open(os.path.join(self.dir, "utt2spk"), "r") as fid:
                for line in fid:
                    parts = line.strip().split()
                    if parts:
                        self.utt2spk[parts[0]] = parts[1]
            table = ArkTable(os.path.join(self.dir, "cmvn.scp"), root=root)
            stats = []
            for k in table.keys:
                self.spk_index[k] = len(stats)
                stats.append(table[k].to_numpy())
            if len({s.dtype for s in stats}) > 1:
                stats = [s.astype(np.float64) for s in stats]
            self.cmvn_mean, self.cmvn_sd = cmvn_tables(stats)
        vocab = _cfg(config, "vocab")
        if root is not None and vocab and not os.path.isabs(vocab):
            vocab = os.path.join(root, vocab)
        self.unit2idx = self._vocab(vocab) if _cfg(config, "encoding", False) else None
        self.targets_dict = self._targets(os.path.join(self.dir, _cfg(config, "text_flag", "text")))
        ids = [k for k in self.feats.keys if k in self.targets_dict]  # dataset.py:198-203
        if _cfg(config, "short_first", False) and type == "train":
            # dataset.py:124-125 sorts the transcript table itself (stable, by label count)
            order = sorted(self.targets_dict.items(), key=lambda kv: len(kv[1]))
            self.utt_ids = [k for k, _ in order]
            self._n = len(ids)
        else:
            self.utt_ids = ids
            self._n = len(ids)
        self._staging = _Staging()

    @staticmethod
    def _vocab(path):
        table = {}
        with open(path, "r", encoding="utf-8") as fid:
            for line in fid:
                parts = line.strip().split()
                if parts:
                    table[parts[0]] = int(parts[1])
        return table

    def _targets(self, path):
        out = {}
        with open(path, "r", encoding="utf-8") as fid:
            for line in fid:
                parts = line.strip().split(" ")
                units = parts[1:]
                if len(units) > self.max_target_length:
                    continue
                if self.unit2idx is not None:
                    unk = self.unit2idx.get("<unk>")
                    out[parts[0]] = [self.unit2idx.get(u, unk) for u in units]
                else:
                    out[parts[0]] = [int(u) for u in units]
        return out

    def __len__(self):
        return self._n  # dataset.py:128,160: the number of feature entries that have a transcript

    def _cmvn_on(self, dev: torch.device):
        if not self.apply_cmvn:
            return None, None
        if dev not in self._cmvn_dev:
            self._cmvn_dev[dev] = (torch.from_numpy(np.ascontiguousarray(self.cmvn_mean)).to(dev),
                                   torch.from_numpy(np.ascontiguousarray(self.cmvn_sd)).to(dev))
        return self._cmvn_dev[dev]

    def batch(self, indices: Iterable[int], device=None, pad_to_max: bool = False):
        """-> ``inputs (B,T',F') float32 cuda, inputs_length (B) int64 cpu, targets (B,U') int64 cuda,
        targets_length (B) int64 cpu``; ``pad_to_max`` pads to ``max_input_length`` /
        ``max_target_length`` as dataset.py:152-156 does instead of to the batch maxima."""
        if not torch.cuda.is_available():
            raise RuntimeError("rnnt_b200: KaldiFeatureSet.batch needs a CUDA device (there is no CPU fallback)")
        ids = [self.utt_ids[i] for i in indices]
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        mats = [self.feats[k] for k in ids]
        f64_feats = any(m.format == FEAT_F64 for m in mats)
        mean, sd = self._cmvn_on(dev)
        if mean is not None and f64_feats and mean.dtype != torch.float64:
            mean, sd = mean.double(), sd.double()
        spk = [self.spk_index[self.utt2spk[k]] for k in ids] if self.apply_cmvn else [-1] * len(ids)
        inputs, in_len = features_batch(mats, spk, mean, sd, self.left, self.right, self.interval,
                                        pad_to=self.max_input_length if pad_to_max else None, device=dev,
                                        staging=self._staging)
        if not pad_to_max and int(in_len.max()) > self.max_input_length:
            raise ValueError(f"utterance of {int(in_len.max())} frames exceeds max_input_length")
        labels = [self.targets_dict[k] for k in ids]
        t_len = torch.tensor([len(l) for l in labels], dtype=torch.int64)
        width = self.max_target_length if pad_to_max else max(1, int(t_len.max()))
        tgt = torch.zeros((len(ids), width), dtype=torch.int64).pin_memory()
        for b, l in enumerate(labels):
            tgt[b, :len(l)] = torch.tensor(l, dtype=torch.int64)
        return inputs, in_len, tgt.to(dev, non_blocking=True), t_len

    def loader(self, batch_size: int, shuffle: bool = False, device=None, pad_to_max: bool = False,
               generator: Optional[torch.Generator] = None, rank: int = 0, world_size: int = 1
               ) -> Iterator[Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]]:
        """Batches in the order ``DataLoader(dataset, batch_size, shuffle)`` would draw them
        (train.py:135-137; the last short batch is kept). With ``world_size > 1`` every rank takes
        utterances ``rank::world_size`` of each global batch of ``batch_size * world_size`` (the
        reference sizes the global batch the same way, train.py:136)."""
        n = len(self)
        order = torch.randperm(n, generator=generator).tolist() if shuffle else list(range(n))
        step = batch_size * world_size
        for lo in range(0, n, step):
            mine = order[lo:lo + step][rank::world_size]
            if mine:
                yield self.batch(mine, device=device, pad_to_max=pad_to_max)
