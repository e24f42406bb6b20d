"""Seeded random shapes: the tiled kernel against the generic kernel (EXACT arithmetic is the same
in both, so the fp32 outputs must be identical bit for bit), and the generic kernel against the C
restatement of the reference on a sample of the cases.  Needs a B200: -m gpu."""
import numpy as np
import pytest
import torch

import pva_b200
from pva_b200 import _lib
from pva_b200.geometry import ClipParams, resized_hw

from _util import c_oracle_explicit, full_kwargs, synth_clip

pytestmark = pytest.mark.gpu


def random_case(rng):
    h = int(rng.integers(40, 300))
    w = int(rng.integers(40, 400))
    if rng.random() < 0.6:
        w -= w % 4  # the fast x-path needs rows that are multiples of 4 bytes; keep both kinds in the mix
    t_in = int(rng.integers(2, 20))
    n_fast = int(rng.choice([2, 4, 8]))
    alpha = int(rng.choice([1, 2, 4]))
    alpha = min(alpha, n_fast)
    crop = int(rng.integers(8, 160))
    short = min(h, w)
    size = int(rng.integers(max(crop, short // 3), max(crop + 1, int(short * 1.6))))
    oh, ow = resized_hw(h, w, size)
    if oh < crop or ow < crop:
        size = crop + int(rng.integers(0, 40))
        oh, ow = resized_hw(h, w, size)
    top = int(rng.integers(0, oh - crop + 1))
    left = int(rng.integers(0, ow - crop + 1))
    flip = bool(rng.random() < 0.5)
    return (3, t_in, h, w), n_fast, alpha, crop, ClipParams(size, oh, ow, top, left, flip)


@pytest.mark.parametrize("seed", range(6))
def test_random_shapes_tiled_equals_generic(seed, cuda_device):
    rng = np.random.default_rng(1000 + seed)
    for case in range(8):
        shape, n_fast, alpha, crop, p = random_case(rng)
        kw = full_kwargs(dict(num_frames=n_fast, slowfast_alpha=alpha, crop_size=crop,
                              mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225]))
        clip = synth_clip(shape, 50_000 + 100 * seed + case).cuda()
        outs = {}
        for kernel in ("generic", "tiled"):
            t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel=kernel)
            try:
                outs[kernel] = t.explicit(clip, p)
            except NotImplementedError:
                outs[kernel] = None  # a shape the tiled kernel declines is served by the generic one
            if kernel == "tiled" and outs[kernel] is not None:
                assert _lib.last_launch_info()["kernel"] == 2
        if outs["tiled"] is None:
            continue
        for a, b in zip(outs["generic"], outs["tiled"]):
            assert torch.equal(a, b), (shape, n_fast, alpha, crop, p)
        # the reference's C x T x S x S memory order from the tiled kernel: same values, contiguous
        tp = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel="tiled",
                                     channels_last=False)
        for a, b in zip(outs["generic"], tp.explicit(clip, p)):
            assert b.is_contiguous() and torch.equal(a, b), ("planar", shape, n_fast, alpha, crop, p)
        assert _lib.last_launch_info()["kernel"] == 2
        # production mode on the same case stays inside the fp16 budget of the exact result
        t16 = pva_b200.make_transform(**kw, kernel="tiled")
        for a, b in zip(outs["generic"], t16.explicit(clip, p)):
            assert (a - b.float()).abs().max().item() <= 1e-3
        if case < 2:  # and the exact result is the reference's arithmetic
            ref = c_oracle_explicit(clip.cpu(), (p.size, p.top, p.left, p.flip), kw)
            for a, r in zip(outs["generic"], ref):
                assert np.array_equal(a.cpu().numpy(), r)


def test_random_ragged_batches(cuda_device):
    rng = np.random.default_rng(7)
    for _ in range(4):
        n_fast, crop = 4, int(rng.choice([32, 48, 96]))
        kw = full_kwargs(dict(num_frames=n_fast, slowfast_alpha=2, crop_size=crop))
        clips, params = [], []
        for i in range(int(rng.integers(2, 7))):
            shape, _, _, _, _ = random_case(rng)
            h, w = max(shape[2], crop), max(shape[3], crop)
            size = max(crop, min(h, w))
            oh, ow = resized_hw(h, w, size)
            clips.append(synth_clip((3, shape[1], h, w), int(rng.integers(1 << 30))).cuda())
            params.append(ClipParams(size, oh, ow, int(rng.integers(0, oh - crop + 1)),
                                     int(rng.integers(0, ow - crop + 1)), bool(rng.random() < 0.5)))
        res = {}
        for kernel in ("generic", "tiled"):
            t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel=kernel)
            res[kernel] = t.run_batch(clips, params)
        for a, b in zip(res["generic"], res["tiled"]):
            assert torch.equal(a, b)
        tp = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel="tiled",
                                     channels_last=False)
        for a, b in zip(res["generic"], tp.run_batch(clips, params)):
            assert b.is_contiguous() and torch.equal(a, b)
