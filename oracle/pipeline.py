"""Portable CPU restatement of the reference transform pipeline.  TEST INFRASTRUCTURE.

This is what runs as the checker on the GPU box (where ``/root/reference`` does
not exist) and as the timed CPU baseline in ``bench.py``.  It restates
``make_transform`` (``run.py:68-102``) and ``PackPathway`` (``run.py:38-65``) over
the restated pytorchvideo wrappers (``oracle.pv_transforms``) and the *real*
torchvision ``RandomCrop`` / ``CenterCrop`` / ``RandomHorizontalFlip`` /
``Compose``.  ``tests/test_oracle.py`` checks it against the golden vectors that
were produced by the reference's own ``run.py`` (``tests/golden/``), and, in the
build container, against ``run.make_transform`` directly.

Two further forms serve the parity tests:

* ``apply_explicit`` -- the same op chain with the four random draws passed in
  as arguments (``size, top, left, flip``) instead of drawn from the global RNG.
* ``replay_draws`` -- draws ``(size, top, left, flip)`` from the global torch
  RNG with the same calls in the same order as the reference's train transform
  (SURVEY App. A.3), so a seeded run reproduces the reference's draws.
"""
from __future__ import annotations

import math
from typing import List, Sequence, Tuple, Union

import torch
from torchvision.transforms import CenterCrop, Compose, RandomCrop, RandomHorizontalFlip

from .pv_transforms import (
    ApplyTransformToKey,
    Div255,
    Normalize,
    RandomShortSideScale,
    RemoveKey,
    ShortSideScale,
    UniformTemporalSubsample,
    short_side_scale,
    uniform_temporal_subsample,
)

DEFAULT_MEAN = (0.45, 0.45, 0.45)
DEFAULT_STD = (0.225, 0.225, 0.225)


class PackPathway(torch.nn.Module):
    """Slow/fast split of ``run.py:38-65``: ``[frames[:, idx], frames]`` with
    ``idx = linspace(0, T-1, T // alpha).long()`` (slow pathway first)."""

    def __init__(self, alpha: int):
        super().__init__()
        self.alpha = alpha

    def forward(self, frames: torch.Tensor) -> List[torch.Tensor]:
        t = frames.shape[1]
        idx = torch.linspace(0, t - 1, t // self.alpha).long()
        return [torch.index_select(frames, 1, idx), frames]


def make_transform(
    min_short_side_scale: int = 256,
    max_short_side_scale: int = 320,
    mean: Sequence[float] = DEFAULT_MEAN,
    std: Sequence[float] = DEFAULT_STD,
    crop_size: int = 256,
    num_frames: int = 32,
    slowfast_alpha: int = 4,
    is_slowfast: bool = True,
    horizontal_flip_p: float = 0.5,
    training: bool = False,
):
    """Same op list, same order, same defaults as ``run.py:68-102``."""
    ops = [UniformTemporalSubsample(num_frames), Div255(), Normalize(list(mean), list(std))]
    if training:
        ops.append(RandomShortSideScale(min_short_side_scale, max_short_side_scale))
        ops.append(RandomCrop(crop_size))
        ops.append(RandomHorizontalFlip(p=horizontal_flip_p))
    else:
        ops.append(ShortSideScale(min_short_side_scale))
        ops.append(CenterCrop(crop_size))
    if is_slowfast:
        ops.append(PackPathway(slowfast_alpha))
    return Compose([ApplyTransformToKey(key="video", transform=Compose(ops)), RemoveKey("audio")])


def resized_hw(h: int, w: int, size: int) -> Tuple[int, int]:
    """Output (height, width) of ``short_side_scale`` -- Python-double arithmetic (App. A.4)."""
    if w < h:
        return int(math.floor((float(h) / w) * size)), size
    return size, int(math.floor((float(w) / h) * size))


def center_crop_offsets(h: int, w: int, crop: int) -> Tuple[int, int]:
    """``torchvision.transforms.functional.center_crop`` offsets (banker's rounding)."""
    return int(round((h - crop) / 2.0)), int(round((w - crop) / 2.0))


def replay_draws(h: int, w: int, min_size: int = 256, max_size: int = 320, crop: int = 256,
                 flip_p: float = 0.5) -> Tuple[int, int, int, bool]:
    """Draw ``(size, top, left, flip)`` from the global RNG in the reference's order.

    1. ``RandomShortSideScale``: ``torch.randint(min, max + 1, (1,))``
    2./3. ``RandomCrop.get_params``: ``randint(0, oh - crop + 1)`` then
       ``randint(0, ow - crop + 1)`` -- both skipped when the resized frame already
       equals the crop (torchvision ``transforms.py:650-651``)
    4. ``RandomHorizontalFlip``: ``torch.rand(1) < p``.
    """
    size = int(torch.randint(min_size, max_size + 1, (1,)).item())
    oh, ow = resized_hw(h, w, size)
    if oh < crop or ow < crop:
        raise ValueError(f"Required crop size {(crop, crop)} is larger than input image size {(oh, ow)}")
    if (oh, ow) == (crop, crop):
        top, left = 0, 0
    else:
        top = int(torch.randint(0, oh - crop + 1, size=(1,)).item())
        left = int(torch.randint(0, ow - crop + 1, size=(1,)).item())
    flip = bool(torch.rand(1) < flip_p)
    return size, top, left, flip


def apply_explicit(
    video: torch.Tensor,
    size: int,
    top: int,
    left: int,
    flip: bool,
    *,
    mean: Sequence[float] = DEFAULT_MEAN,
    std: Sequence[float] = DEFAULT_STD,
    crop_size: int = 256,
    num_frames: int = 32,
    slowfast_alpha: int = 4,
    is_slowfast: bool = True,
) -> Union[torch.Tensor, List[torch.Tensor]]:
    """The reference op chain (``run.py:81-100``) with the random draws given explicitly.

    Every arithmetic step is the same torch call the Compose would make:
    ``index_select`` -> ``/ 255.0`` -> ``Normalize`` -> ``F.interpolate`` ->
    window slice -> ``flip(-1)`` -> ``PackPathway``.
    """
    x = uniform_temporal_subsample(video, num_frames)
    x = Div255()(x)
    x = Normalize(list(mean), list(std))(x)
    x = short_side_scale(x, size)
    x = x[..., top:top + crop_size, left:left + crop_size]
    if flip:
        x = x.flip(-1)
    if is_slowfast:
        return PackPathway(slowfast_alpha)(x)
    return x


def val_draws(h: int, w: int, size: int = 256, crop: int = 256) -> Tuple[int, int, int, bool]:
    """The deterministic "draws" of the val transform: fixed size, centre crop, no flip."""
    oh, ow = resized_hw(h, w, size)
    top, left = center_crop_offsets(oh, ow, crop)
    return size, top, left, False
