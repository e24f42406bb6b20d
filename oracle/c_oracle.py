"""ctypes front-end of ``oracle/c/pva_oracle.c``.  TEST INFRASTRUCTURE (see ``oracle/__init__.py``)."""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpva_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "c", "pva_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "libpva_oracle.so"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.pva_oracle_clip.restype = ctypes.c_int
    return _lib


def build_lut(mean: Sequence[float], std: Sequence[float]) -> np.ndarray:
    c = len(mean)
    m = np.asarray(mean, dtype=np.float32)
    s = np.asarray(std, dtype=np.float32)
    out = np.empty((c, 256), dtype=np.float32)
    lib().pva_oracle_build_lut(m.ctypes.data_as(ctypes.c_void_p), s.ctypes.data_as(ctypes.c_void_p),
                               ctypes.c_int(c), out.ctypes.data_as(ctypes.c_void_p))
    return out


def clip(src: np.ndarray, idx1: np.ndarray, idx2: Optional[np.ndarray], oh: int, ow: int, top: int, left: int,
         flip: bool, crop: int, lut: np.ndarray) -> Tuple[np.ndarray, Optional[np.ndarray]]:
    """Run one uint8 ``C x T x H x W`` clip through the C restatement; returns fp32 ``(fast, slow)`` in NCDHW."""
    src = np.ascontiguousarray(src, dtype=np.uint8)
    c, t, h, w = src.shape
    idx1 = np.ascontiguousarray(idx1, dtype=np.int32)
    n_fast = idx1.shape[0]
    fast = np.empty((c, n_fast, crop, crop), dtype=np.float32)
    if idx2 is not None:
        idx2 = np.ascontiguousarray(idx2, dtype=np.int32)
        n_slow = idx2.shape[0]
        slow = np.empty((c, n_slow, crop, crop), dtype=np.float32)
        slow_p, idx2_p = slow.ctypes.data_as(ctypes.c_void_p), idx2.ctypes.data_as(ctypes.c_void_p)
    else:
        n_slow, slow, slow_p, idx2_p = 0, None, None, None
    lut = np.ascontiguousarray(lut, dtype=np.float32)
    rc = lib().pva_oracle_clip(
        src.ctypes.data_as(ctypes.c_void_p), c, t, h, w,
        idx1.ctypes.data_as(ctypes.c_void_p), n_fast, idx2_p, n_slow,
        int(oh), int(ow), int(top), int(left), int(bool(flip)), int(crop),
        lut.ctypes.data_as(ctypes.c_void_p), fast.ctypes.data_as(ctypes.c_void_p), slow_p)
    if rc != 0:
        raise ValueError("pva_oracle_clip: window or index outside the frame")
    return fast, slow


def footprint(h: int, w: int, oh: int, ow: int, top: int, left: int, crop: int) -> Tuple[int, int]:
    fh, fw = ctypes.c_int(0), ctypes.c_int(0)
    lib().pva_oracle_footprint(h, w, oh, ow, top, left, crop, ctypes.byref(fh), ctypes.byref(fw))
    return fh.value, fw.value
