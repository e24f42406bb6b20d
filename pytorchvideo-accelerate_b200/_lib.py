"""ctypes binding of ``libpva_preproc.so`` (C ABI: ``include/pva_preproc.h``).

There is no CPU fallback and no alternative backend: if the shared library is missing
or does not export the ABI, importing callers get a ``PvaError`` telling them to build it
(``python -c "import __graft_entry__ as g; g.build()"``).
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

PVA_MAX_CHANNELS = 4

# enum values of include/pva_preproc.h
PVA_OK = 0
PVA_ERR_INVALID_ARGUMENT = -1
PVA_ERR_CROP_TOO_LARGE = -2
PVA_ERR_UNSUPPORTED = -3
PVA_ERR_CUDA = -4

PVA_F32, PVA_F16, PVA_BF16 = 0, 1, 2
PVA_LAYOUT_CTHW, PVA_LAYOUT_THWC = 0, 1
PVA_NORM_EXACT, PVA_NORM_AFFINE = 0, 1
PVA_KERNEL_AUTO, PVA_KERNEL_GENERIC, PVA_KERNEL_TILED = 0, 1, 2
PVA_SRC_U8, PVA_SRC_F32 = 0, 1
PVA_FLAG_CLIPS_READY = 1
PVA_FLAG_OUTPUTS_DISJOINT = 2

EXPORTED_SYMBOLS = (
    "pva_preprocess",
    "pva_build_norm_lut",
    "pva_source_footprint",
    "pva_upload_selected_frames",
    "pva_last_launch_info",
    "pva_version",
    "pva_last_error_string",
)


class PvaError(RuntimeError):
    """Raised for library-level failures (missing .so, CUDA errors)."""


class PvaClipDesc(ctypes.Structure):
    _fields_ = [
        ("src", ctypes.c_void_p),
        ("frame_idx", ctypes.POINTER(ctypes.c_int32)),
        ("t_in", ctypes.c_int32),
        ("h", ctypes.c_int32),
        ("w", ctypes.c_int32),
        ("oh", ctypes.c_int32),
        ("ow", ctypes.c_int32),
        ("top", ctypes.c_int32),
        ("left", ctypes.c_int32),
        ("flip", ctypes.c_int32),
    ]


class PvaPlan(ctypes.Structure):
    _fields_ = [
        ("channels", ctypes.c_int32),
        ("n_fast", ctypes.c_int32),
        ("n_slow", ctypes.c_int32),
        ("crop", ctypes.c_int32),
        ("slow_idx", ctypes.POINTER(ctypes.c_int32)),
        ("mean", ctypes.c_float * PVA_MAX_CHANNELS),
        ("std", ctypes.c_float * PVA_MAX_CHANNELS),
        ("out_dtype", ctypes.c_int32),
        ("out_layout", ctypes.c_int32),
        ("norm_mode", ctypes.c_int32),
        ("kernel", ctypes.c_int32),
        ("src_type", ctypes.c_int32),
        ("flags", ctypes.c_int32),
        ("max_ctas", ctypes.c_int32),
        ("reserved", ctypes.c_int32),
    ]


class PvaLaunchInfo(ctypes.Structure):
    _fields_ = [
        ("kernel", ctypes.c_int32),
        ("launches", ctypes.c_int32),
        ("grid", ctypes.c_int32),
        ("block", ctypes.c_int32),
        ("smem_bytes", ctypes.c_int32),
        ("rows_per_tile", ctypes.c_int32),
        ("tiles", ctypes.c_int32),
        ("reserved", ctypes.c_int32),
    ]


_LIB: Optional[ctypes.CDLL] = None


def library_path() -> str:
    """The in-tree library; PVA_LIBRARY points at another build of the same ABI (kernel A/B runs)."""
    return os.environ.get("PVA_LIBRARY") or os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                                         "libpva_preproc.so")


def load_library() -> ctypes.CDLL:
    """Load (once) and type the C ABI.  Loud failure when the library is absent."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if not os.path.exists(path):
        raise PvaError(
            f"{path} not found: the CUDA library has not been built. "
            "Run `python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc). "
            "There is no CPU fallback for this path."
        )
    lib = ctypes.CDLL(path)
    for name in EXPORTED_SYMBOLS:
        if not hasattr(lib, name):
            raise PvaError(f"{path} does not export {name}; rebuild it")
    lib.pva_preprocess.restype = ctypes.c_int
    lib.pva_preprocess.argtypes = [
        ctypes.POINTER(PvaClipDesc), ctypes.c_int32, ctypes.POINTER(PvaPlan),
        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
    ]
    lib.pva_build_norm_lut.restype = ctypes.c_int
    lib.pva_build_norm_lut.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p]
    lib.pva_source_footprint.restype = ctypes.c_int
    lib.pva_source_footprint.argtypes = [ctypes.c_int32] * 7 + [ctypes.POINTER(ctypes.c_int32)] * 2
    lib.pva_upload_selected_frames.restype = ctypes.c_int
    lib.pva_upload_selected_frames.argtypes = [ctypes.c_void_p, ctypes.c_void_p] + [ctypes.c_int32] * 4 + [
        ctypes.POINTER(ctypes.c_int32), ctypes.c_int32, ctypes.c_void_p]
    lib.pva_last_launch_info.restype = ctypes.c_int
    lib.pva_last_launch_info.argtypes = [ctypes.POINTER(PvaLaunchInfo)]
    lib.pva_version.restype = ctypes.c_int32
    lib.pva_version.argtypes = []
    lib.pva_last_error_string.restype = ctypes.c_char_p
    lib.pva_last_error_string.argtypes = []
    _LIB = lib
    return lib


def last_error() -> str:
    return load_library().pva_last_error_string().decode("utf-8", "replace")


def raise_for_status(rc: int) -> None:
    """Map a PvaStatus to the exception type the reference ops would raise."""
    if rc == PVA_OK:
        return
    msg = last_error()
    if rc in (PVA_ERR_CROP_TOO_LARGE, PVA_ERR_INVALID_ARGUMENT):
        raise ValueError(msg)
    if rc == PVA_ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise PvaError(msg)


def source_footprint(h: int, w: int, oh: int, ow: int, top: int, left: int, crop: int):
    fh, fw = ctypes.c_int32(0), ctypes.c_int32(0)
    raise_for_status(load_library().pva_source_footprint(h, w, oh, ow, top, left, crop, ctypes.byref(fh), ctypes.byref(fw)))
    return fh.value, fw.value


def last_launch_info() -> dict:
    info = PvaLaunchInfo()
    raise_for_status(load_library().pva_last_launch_info(ctypes.byref(info)))
    return {name: getattr(info, name) for name, _ in PvaLaunchInfo._fields_ if name != "reserved"}
