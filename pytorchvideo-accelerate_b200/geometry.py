"""Host-side geometry and random draws of the transform chain -- integer/bit-exact work.

Everything here is what the reference decides on the CPU *before* any pixel is touched:
which frames, what resized size, which crop window, flip or not.  The CUDA kernels take
these as explicit per-clip arguments, so a seeded run reproduces the reference's choices
exactly.  torch is used on purpose for the index tables and
the RNG calls: ``torch.linspace(...).long()`` is not closed-form in integers, and the draws
must consume the same global generator the reference's transforms consume.
"""
from __future__ import annotations

import functools
import math
from dataclasses import dataclass
from typing import Tuple

import torch


@dataclass(frozen=True)
class ClipParams:
    """Per-clip draws + derived geometry handed to the kernel.

    ``size`` is the short-side target (draw #1 of ``RandomShortSideScale``, ``run.py:89``),
    ``oh, ow`` the resized frame, ``top, left`` the crop origin inside it (``run.py:90,96``),
    ``flip`` the outcome of ``RandomHorizontalFlip`` (``run.py:91``).
    """

    size: int
    oh: int
    ow: int
    top: int
    left: int
    flip: bool

    def as_row(self):
        return [self.size, self.oh, self.ow, self.top, self.left, int(self.flip)]

    @staticmethod
    def from_row(row) -> "ClipParams":
        s, oh, ow, top, left, flip = (int(v) for v in row)
        return ClipParams(s, oh, ow, top, left, bool(flip))


@functools.lru_cache(maxsize=256)
def temporal_indices(t_in: int, num_samples: int) -> Tuple[int, ...]:
    """Frame indices of ``UniformTemporalSubsample(num_samples)`` on a clip of ``t_in`` frames
    (``run.py:82``): fp32 ``linspace(0, t-1, n)``, clamped, truncated."""
    if num_samples <= 0 or t_in <= 0:
        raise ValueError(f"num_samples={num_samples} and t_in={t_in} must be positive")
    idx = torch.clamp(torch.linspace(0, t_in - 1, num_samples), 0, t_in - 1).long()
    return tuple(int(v) for v in idx.tolist())


@functools.lru_cache(maxsize=256)
def pack_indices(n_fast: int, alpha: int) -> Tuple[int, ...]:
    """Slow-pathway frame indices of ``PackPathway(alpha)`` (``run.py:59-63``):
    ``linspace(0, T-1, T // alpha).long()`` over the already-subsampled frames."""
    n_slow = n_fast // alpha
    idx = torch.linspace(0, n_fast - 1, n_slow).long()
    return tuple(int(v) for v in idx.tolist())


def resized_hw(h: int, w: int, size: int) -> Tuple[int, int]:
    """Frame size after ``short_side_scale`` -- evaluated in Python doubles exactly like the
    reference op does; ``w * size // h`` would differ for some triples (App. A.4)."""
    if w < h:
        return int(math.floor((float(h) / w) * size)), size
    return size, int(math.floor((float(w) / h) * size))


def center_crop_offsets(h: int, w: int, crop: int) -> Tuple[int, int]:
    """``CenterCrop`` origin (``run.py:96``): Python ``round`` = banker's rounding, as in
    ``torchvision.transforms.functional.center_crop``."""
    return int(round((h - crop) / 2.0)), int(round((w - crop) / 2.0))


def _check_window(oh: int, ow: int, crop: int) -> None:
    if oh < crop or ow < crop:
        # same message as torchvision.transforms.RandomCrop.get_params
        raise ValueError(f"Required crop size {(crop, crop)} is larger than input image size {(oh, ow)}")


def val_params(h: int, w: int, size: int, crop: int) -> ClipParams:
    """Validation geometry (``run.py:93-97``): fixed short side, centre crop, no flip, no draws.

    torchvision's ``center_crop`` zero-pads frames smaller than the crop; with the
    reference's settings that branch is unreachable and it is rejected here."""
    oh, ow = resized_hw(h, w, size)
    _check_window(oh, ow, crop)
    top, left = center_crop_offsets(oh, ow, crop)
    return ClipParams(size, oh, ow, top, left, False)


def sample_train_params(h: int, w: int, min_size: int, max_size: int, crop: int, flip_p: float) -> ClipParams:
    """Training draws (``run.py:87-92``) from the *global* torch generator, with the same
    calls in the same order as the reference ops make them (App. A.3):

    1. ``torch.randint(min, max + 1, (1,))``                       RandomShortSideScale
    2. ``torch.randint(0, oh - crop + 1, size=(1,))``              RandomCrop.get_params (top)
    3. ``torch.randint(0, ow - crop + 1, size=(1,))``              RandomCrop.get_params (left)
       -- 2 and 3 are skipped when the resized frame already equals the crop
    4. ``torch.rand(1) < p``                                       RandomHorizontalFlip
    """
    size = int(torch.randint(min_size, max_size + 1, (1,)).item())
    oh, ow = resized_hw(h, w, size)
    _check_window(oh, ow, crop)
    if (oh, ow) == (crop, crop):
        top, left = 0, 0
    else:
        top = int(torch.randint(0, oh - crop + 1, size=(1,)).item())
        left = int(torch.randint(0, ow - crop + 1, size=(1,)).item())
    flip = bool(torch.rand(1) < flip_p)
    return ClipParams(size, oh, ow, top, left, flip)
