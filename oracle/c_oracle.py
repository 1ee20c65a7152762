"""ctypes binding of oracle/rnnt_oracle.c (TEST INFRASTRUCTURE ONLY).

Used by tests/ (parity checker), __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs. Never imported by the product package.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "librnnt_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "rnnt_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "librnnt_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.rnnt_oracle_loss.restype = ctypes.c_int
        _lib.rnnt_oracle_loss.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int] * 5 + [ctypes.c_void_p] * 4
        _lib.rnnt_oracle_num_threads.restype = ctypes.c_int
    return _lib


def num_threads() -> int:
    return int(lib().rnnt_oracle_num_threads())


def rnnt_loss(logits, labels, act_lens, label_lens, blank=0, want_grad=True, want_coeffs=False):
    """numpy in / numpy out. logits (B,T,U1,V) float32. Returns (loss f64 (B,), grad f32 or None
    [, c_blank, c_label f64])."""
    x = np.ascontiguousarray(logits, dtype=np.float32)
    B, T, U1, V = x.shape
    y = np.ascontiguousarray(labels, dtype=np.int32).reshape(B, max(U1 - 1, 0))
    tl = np.ascontiguousarray(act_lens, dtype=np.int32)
    ul = np.ascontiguousarray(label_lens, dtype=np.int32)
    loss = np.zeros(B, dtype=np.float64)
    grad = np.empty_like(x) if want_grad else None
    cb = np.zeros((B, T, U1), dtype=np.float64) if want_coeffs else None
    cl = np.zeros((B, T, U1), dtype=np.float64) if want_coeffs else None
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p) if a is not None else None
    if y.size == 0:
        y = np.zeros(1, dtype=np.int32)
    rc = lib().rnnt_oracle_loss(p(x), p(y), p(tl), p(ul), B, T, U1, V, blank,
                                p(loss), p(grad), p(cb), p(cl))
    if rc != 0:
        raise ValueError(f"rnnt_oracle_loss: invalid argument (code {rc})")
    if want_coeffs:
        return loss, grad, cb, cl
    return loss, grad
