"""Importable name of the package that lives in ``pytorchvideo-accelerate_b200/``.

The directory name required for this repo contains a hyphen, which Python cannot
import; this shim points the package search path at it, so
``import pva_b200.transforms`` loads ``pytorchvideo-accelerate_b200/transforms.py``.
"""
import os as _os

_impl = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "pytorchvideo-accelerate_b200")
if not _os.path.isdir(_impl):
    raise ImportError(f"package directory missing: {_impl}")
__path__.insert(0, _impl)

from .geometry import (  # noqa: E402
    ClipParams,
    center_crop_offsets,
    pack_indices,
    resized_hw,
    sample_train_params,
    temporal_indices,
    val_params,
)
from .transforms import (  # noqa: E402
    GpuClipTransform,
    GpuPreprocess,
    PackPathway,
    SampleParams,
    collate_raw_clips,
    make_transform,
)
from .staging import StagedPreprocess  # noqa: E402
from .sharding import batches_for_rank, rank_and_world, rank_seed  # noqa: E402
from ._lib import PvaError, library_path, load_library  # noqa: E402

__all__ = [
    "ClipParams", "center_crop_offsets", "pack_indices", "resized_hw", "sample_train_params",
    "temporal_indices", "val_params", "GpuClipTransform", "GpuPreprocess", "PackPathway", "SampleParams",
    "collate_raw_clips", "make_transform", "StagedPreprocess", "batches_for_rank", "rank_and_world", "rank_seed", "PvaError", "library_path", "load_library",
]
