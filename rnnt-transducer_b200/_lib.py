"""ctypes loader of the C-ABI library (include/rnnt_b200.h). No CPU fallback: every
compute entry point raises if the library or a CUDA device is missing."""
from __future__ import annotations

import ctypes
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# RNNT_B200_LIB: another build of the same library (A/B timing of compile-time variants in one call)
SO_PATH = os.environ.get("RNNT_B200_LIB") or os.path.join(_HERE, "librnnt_b200.so")
CSRC = os.path.join(_HERE, "csrc")
_lock = threading.Lock()
_lib = None

_i = ctypes.c_int
_p = ctypes.c_void_p

# name -> (restype, argtypes); mirrors include/rnnt_b200.h one to one
SIGNATURES = {
    "rnnt_b200_version": (_i, []),
    "rnnt_b200_status_string": (ctypes.c_char_p, [_i]),
    "rnnt_b200_plane_elems": (ctypes.c_size_t, [_i, _i, _i]),
    "rnnt_b200_logits_to_planes": (_i, [_p, _p, _p, _p, _i, _i, _i, _i, _i, _p, _p, _p, _p]),
    "rnnt_b200_lattice": (_i, [_p, _p, _p, _p, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p]),
    "rnnt_b200_lattice_static": (_i, [_p, _p, _p, _p, _i, _i, _i, _p, _p, _p, _p, _p]),
    "rnnt_b200_planes_to_dlogits": (_i, [_p, _p, _p, _p, _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p]),
    "rnnt_b200_max_tiles": (_i, [_i, _i, _i]),
    "rnnt_b200_build_tiles": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _p, _p]),
    "rnnt_b200_vocab_pad": (_i, [_i]),
    "rnnt_b200_vocab_tile": (_i, [_i]),
    "rnnt_b200_pack_vocab": (_i, [_p, _p, _i, _i, _p, _p, _p, _p]),
    "rnnt_b200_h_pitch": (_i, [_i]),
    "rnnt_b200_ds_pitch": (_i, [_i]),
    "rnnt_b200_joint_hidden": (_i, [_p] * 7 + [_i] * 5 + [_p, _p]),
    "rnnt_b200_joint_fwd": (_i, [_p] * 9 + [_i] * 7 + [_p] * 4),
    "rnnt_b200_joint_bwd_ds": (_i, [_p] * 15 + [_i] * 9 + [_p, _p]),
    "rnnt_b200_joint_step_argmax": (_i, [_p, ctypes.c_longlong, _p, _p, _p, _i, _i, _i, _p, _p, _p]),
    "rnnt_b200_dw2_splits": (_i, [ctypes.c_longlong, _i, _i, _i]),
    "rnnt_b200_dw2_gemm": (_i, [_p, _p, ctypes.c_longlong, _i, _i, _i, _p, _p, _p]),
    "rnnt_b200_sum_partials": (_i, [_p, _i, ctypes.c_longlong, _p, _p]),
    "rnnt_b200_dh_gemm_supported": (_i, [_i]),
    "rnnt_b200_dh_gemm": (_i, [_p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _p]),
    "rnnt_b200_joint_fwd_store": (_i, [_p] * 9 + [_i] * 7 + [_p] * 7),
    "rnnt_b200_patch_scale": (_i, [_p] * 14 + [_i] * 10 + [_p] * 4),
    "rnnt_b200_dh_gemm_scaled": (_i, [_p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _p]),
    "rnnt_b200_dw2_gemm_weighted": (_i, [_p, _p, _p, ctypes.c_longlong, _i, _i, _i, _p, _p, _p]),
    "rnnt_b200_dpre_reduce_scratch_elems": (ctypes.c_size_t, [_i, _i, _i, _i]),
    "rnnt_b200_dpre_reduce": (_i, [_p, _p, _p, _i, _i, _i, _i, _p, _p, _p, _p]),
    "rnnt_b200_features_batch": (_i, [_p] * 5 + [_i] * 4 + [_p, _p, _i, _i, _p, _p]),
    "rnnt_b200_selftest_gemm": (_i, [_p, _p, _p, _i, _i, _i, _p]),
    "rnnt_b200_selftest_gemm_ts": (_i, [_p, _p, _p, _i, _i, _i, _i, _p]),
}


def build(verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a into the in-tree shared library."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", CSRC, "-j8"], stdout=out)
    return SO_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(SO_PATH):
                raise RuntimeError(
                    f"rnnt_b200: CUDA extension missing ({SO_PATH}); run `python -c 'import "
                    "__graft_entry__ as g; g.build()'` or `make -C rnnt-transducer_b200/csrc`. "
                    "There is no CPU fallback.")
            handle = ctypes.CDLL(SO_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(handle, name)  # AttributeError = header/library mismatch: be loud
                fn.restype = res
                fn.argtypes = args
            _lib = handle
    return _lib


class RnntB200Error(RuntimeError):
    pass


def check(status: int, what: str) -> None:
    if status != 0:
        msg = lib().rnnt_b200_status_string(status).decode()
        raise RnntB200Error(f"{what} failed: status {status} ({msg})")
