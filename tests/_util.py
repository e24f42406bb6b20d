"""Helpers shared by the test modules (golden loading, oracle calls)."""
from __future__ import annotations

import json
import os

import numpy as np
import torch

from oracle import c_oracle, pipeline

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

REF_DEFAULTS = dict(min_short_side_scale=256, max_short_side_scale=320, mean=[0.45, 0.45, 0.45],
                    std=[0.225, 0.225, 0.225], crop_size=256, num_frames=32, slowfast_alpha=4, is_slowfast=True,
                    horizontal_flip_p=0.5, training=False)


def full_kwargs(kwargs):
    kw = dict(REF_DEFAULTS)
    kw.update(kwargs)
    return kw


def load_small():
    meta = json.load(open(os.path.join(GOLDEN, "golden_small.json")))
    data = np.load(os.path.join(GOLDEN, "golden_small.npz"))
    return meta, data


def load_large():
    return json.load(open(os.path.join(GOLDEN, "golden_large.json")))


def load_tables():
    return json.load(open(os.path.join(GOLDEN, "golden_tables.json")))


def synth_clip(shape, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, tuple(shape), dtype=torch.uint8, generator=g)


def draws_for_case(kw, h, w, seed):
    """Replay the reference's draws for a golden case (seed the global RNG like make_golden did)."""
    if kw["training"]:
        torch.manual_seed(seed)
        return pipeline.replay_draws(h, w, kw["min_short_side_scale"], kw["max_short_side_scale"], kw["crop_size"],
                                     kw["horizontal_flip_p"])
    return pipeline.val_draws(h, w, kw["min_short_side_scale"], kw["crop_size"])


def oracle_explicit(clip, draws, kw):
    """Portable torch oracle with explicit draws -> list [slow, fast] or [single]."""
    out = pipeline.apply_explicit(clip, *draws, mean=kw["mean"], std=kw["std"], crop_size=kw["crop_size"],
                                  num_frames=kw["num_frames"], slowfast_alpha=kw["slowfast_alpha"],
                                  is_slowfast=kw["is_slowfast"])
    return out if isinstance(out, list) else [out]


def c_oracle_explicit(clip, draws, kw):
    """C restatement with explicit draws -> list like oracle_explicit (numpy fp32, CTHW)."""
    size, top, left, flip = draws
    _c, t, h, w = clip.shape
    oh, ow = pipeline.resized_hw(h, w, size)
    idx1 = torch.clamp(torch.linspace(0, t - 1, kw["num_frames"]), 0, t - 1).long().numpy()
    idx2 = None
    if kw["is_slowfast"]:
        idx2 = torch.linspace(0, kw["num_frames"] - 1, kw["num_frames"] // kw["slowfast_alpha"]).long().numpy()
    lut = c_oracle.build_lut(kw["mean"], kw["std"])
    fast, slow = c_oracle.clip(clip.numpy(), idx1, idx2, oh, ow, top, left, flip, kw["crop_size"], lut)
    return [slow, fast] if kw["is_slowfast"] else [fast]
