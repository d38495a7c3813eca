"""ctypes binding of the C-ABI library (include/nonmyopia_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` /
``python -m nonmyopia_b200.build`` into ``nonmyopia_b200/lib/``.  There is no CPU
fallback: if the library is missing or no CUDA device is present every op raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import torch

LIB_PATH = os.environ.get("NMB_LIB_PATH") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib",
                                                          "libnonmyopia_b200.so")   # NMB_LIB_PATH: A/B builds (tools/)

NMB_F64, NMB_F32 = 0, 1
KERNEL_IDS = {"rbf": 0, "matern12": 1, "matern32": 2, "matern52": 3}

_lib = None
_lock = threading.Lock()


class NmbError(RuntimeError):
    pass


class PathStruct(C.Structure):
    """``nmb_path_t``"""
    _fields_ = [
        ("d", C.c_int32), ("n_freq", C.c_int32), ("n_train", C.c_int32), ("n_paths", C.c_int32),
        ("kernel_id", C.c_int32), ("_pad", C.c_int32),
        ("feat_scale", C.c_double), ("amplitude", C.c_double), ("mean_const", C.c_double),
        ("w_phase", C.c_void_p), ("w_feat", C.c_void_p), ("inv_ls", C.c_void_p),
        ("train_x", C.c_void_p), ("v_update", C.c_void_p),
    ]


def lib() -> C.CDLL:
    """Load the shared library (once)."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.isfile(LIB_PATH):
                    raise NmbError(
                        f"{LIB_PATH} not built - run `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(no CPU fallback exists)")
                l = C.CDLL(LIB_PATH)
                l.nmb_last_error.restype = C.c_char_p
                _lib = l
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().nmb_last_error().decode(errors="replace")
        raise NmbError(f"{what}: rc={rc}: {msg}")


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float64:
        return NMB_F64
    if dt == torch.float32:
        return NMB_F32
    raise TypeError(f"nonmyopia_b200 kernels take float64 or float32 tensors, got {dt}")


def ptr(t):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return C.c_void_p(0)
    if not t.is_cuda:
        raise NmbError("nonmyopia_b200 ops need CUDA tensors (no CPU fallback)")
    if not t.is_contiguous():
        raise NmbError("tensor must be contiguous")
    return C.c_void_p(t.data_ptr())


def stream_ptr() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
