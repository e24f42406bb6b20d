"""Restated ``pytorchvideo.transforms`` wrappers used at ``run.py:10-18``.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  pytorchvideo is an unpinned
third-party dependency of the reference (``requirements.txt:5``) whose source is
absent from ``/root/reference`` and from this image.  These seven classes restate
its published behaviour over the installed torch / torchvision; each docstring
names the reference call site it serves.  The restatement of
``UniformTemporalSubsample`` is corroborated by torchvision's own port, which
cites the pytorchvideo line it copies
(``torchvision/transforms/v2/functional/_temporal.py``).
"""
from __future__ import annotations

import math
from typing import Callable, Dict

import torch
import torchvision


class ApplyTransformToKey:
    """``run.py:102`` -- apply ``transform`` to ``x[key]`` and hand back the same dict."""

    def __init__(self, key: str, transform: Callable):
        self._key = key
        self._transform = transform

    def __call__(self, x: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        x[self._key] = self._transform(x[self._key])
        return x


class RemoveKey(torch.nn.Module):
    """``run.py:102`` -- drop ``key`` from the sample dict when present."""

    def __init__(self, key: str):
        super().__init__()
        self._key = key

    def __call__(self, x: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        if self._key in x:
            del x[self._key]
        return x


def uniform_temporal_subsample(x: torch.Tensor, num_samples: int, temporal_dim: int = -3) -> torch.Tensor:
    """Frame gather used by ``UniformTemporalSubsample`` (``run.py:82``).

    fp32 ``linspace`` over ``[0, T-1]``, clamped, truncated to int64, then
    ``index_select`` along the temporal axis.
    """
    t = x.shape[temporal_dim]
    assert num_samples > 0 and t > 0
    indices = torch.linspace(0, t - 1, num_samples)
    indices = torch.clamp(indices, 0, t - 1).long()
    return torch.index_select(x, temporal_dim, indices)


class UniformTemporalSubsample(torch.nn.Module):
    """``run.py:82``."""

    def __init__(self, num_samples: int, temporal_dim: int = -3):
        super().__init__()
        self._num_samples = num_samples
        self._temporal_dim = temporal_dim

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return uniform_temporal_subsample(x, self._num_samples, self._temporal_dim)


class Div255(torch.nn.Module):
    """``run.py:83`` -- true division by 255.0; uint8 promotes to fp32."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return x / 255.0


class Normalize(torchvision.transforms.Normalize):
    """``run.py:84`` -- torchvision ``Normalize`` applied with (C,T) swapped to (T,C) and back."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        vid = x.permute(1, 0, 2, 3)
        vid = super().forward(vid)
        vid = vid.permute(1, 0, 2, 3)
        return vid


def short_side_scale(x: torch.Tensor, size: int, interpolation: str = "bilinear") -> torch.Tensor:
    """Spatial resize used by ``ShortSideScale`` / ``RandomShortSideScale`` (``run.py:89,95``).

    The short side becomes ``size``; the long side is ``floor((long/short)*size)``
    evaluated in Python doubles.  Interpolation is
    ``F.interpolate(..., align_corners=False)`` without antialiasing, frames acting
    as the channel dimension of a 4-D ``C x T x H x W`` batch.
    """
    assert len(x.shape) == 4
    assert x.dtype == torch.float32
    _c, _t, h, w = x.shape
    if w < h:
        new_h = int(math.floor((float(h) / w) * size))
        new_w = size
    else:
        new_h = size
        new_w = int(math.floor((float(w) / h) * size))
    return torch.nn.functional.interpolate(x, size=(new_h, new_w), mode=interpolation, align_corners=False)


class ShortSideScale(torch.nn.Module):
    """``run.py:95``."""

    def __init__(self, size: int, interpolation: str = "bilinear"):
        super().__init__()
        self._size = size
        self._interpolation = interpolation

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return short_side_scale(x, self._size, self._interpolation)


class RandomShortSideScale(torch.nn.Module):
    """``run.py:89`` -- one ``torch.randint(min, max + 1)`` draw (max inclusive), then ``short_side_scale``."""

    def __init__(self, min_size: int, max_size: int, interpolation: str = "bilinear"):
        super().__init__()
        self._min_size = min_size
        self._max_size = max_size
        self._interpolation = interpolation

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        size = torch.randint(self._min_size, self._max_size + 1, (1,)).item()
        return short_side_scale(x, size, self._interpolation)
