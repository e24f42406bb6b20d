"""Edge cases of the CUDA path: launch chunking, ragged index tables, odd crops, strong
down-scaling, channel counts, and tiled-vs-generic equality.  Needs a B200: -m gpu."""
import os

import numpy as np
import pytest
import torch

import pva_b200
from oracle import pipeline
from pva_b200 import _lib
from pva_b200.geometry import ClipParams

from _util import c_oracle_explicit, full_kwargs, synth_clip

pytestmark = pytest.mark.gpu


def oracle_clip(clip, p, kw):
    return c_oracle_explicit(clip, (p.size, p.top, p.left, p.flip), kw)


def test_more_clips_than_one_parameter_block_holds(cuda_device):
    """> 128 clips and > 8 distinct frame tables force several launches inside one ABI call."""
    kw = full_kwargs(dict(min_short_side_scale=32, max_short_side_scale=40, crop_size=32, num_frames=4,
                          slowfast_alpha=2, training=True))
    t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact")
    torch.manual_seed(5)
    clips, params = [], []
    for i in range(150):
        t_in = 4 + (i % 11)  # 11 distinct temporal tables
        c = synth_clip((3, t_in, 40, 52), 7000 + i)
        clips.append(c)
        params.append(t.sample_params(40, 52))
    slow, fast = t.run_batch([c.cuda() for c in clips], params)
    torch.cuda.synchronize()
    assert _lib.last_launch_info()["launches"] >= 2
    assert fast.shape == (150, 3, 4, 32, 32) and slow.shape == (150, 3, 2, 32, 32)
    for i in (0, 7, 63, 127, 128, 149):
        rs, rf = oracle_clip(clips[i], params[i], kw)
        assert np.array_equal(fast[i].cpu().numpy(), rf) and np.array_equal(slow[i].cpu().numpy(), rs)


@pytest.mark.parametrize("crop", [30, 33, 45])
def test_crops_not_divisible_by_four_take_the_plain_store_path(crop, cuda_device):
    kw = full_kwargs(dict(min_short_side_scale=48, max_short_side_scale=56, crop_size=crop, num_frames=4,
                          slowfast_alpha=2, training=True))
    clip = synth_clip((3, 8, 52, 68), crop)
    torch.manual_seed(crop)
    for dtype in (torch.float32, torch.float16):
        t = pva_b200.make_transform(**kw, out_dtype=dtype, norm_mode="exact", kernel="tiled")
        p = t.sample_params(52, 68)
        slow, fast = t.explicit(clip.cuda(), p)
        rs, rf = oracle_clip(clip, p, kw)
        want_f = torch.from_numpy(rf).to(dtype).float().numpy()
        want_s = torch.from_numpy(rs).to(dtype).float().numpy()
        assert np.array_equal(fast.float().cpu().numpy(), want_f)
        assert np.array_equal(slow.float().cpu().numpy(), want_s)


@pytest.mark.parametrize("kernel", ["generic", "tiled"])
def test_strong_downscale_of_a_large_frame(kernel, cuda_device):
    """720x1280 frames to a 224 crop: ~3x down-scaling, byte-tap x-pass, small row tiles."""
    kw = full_kwargs(dict(min_short_side_scale=224, max_short_side_scale=256, crop_size=224, num_frames=2,
                          is_slowfast=False, training=True))
    clip = synth_clip((3, 3, 720, 1280), 31)
    torch.manual_seed(31)
    t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel=kernel)
    p = t.sample_params(720, 1280)
    out = t.explicit(clip.cuda(), p)
    (rf,) = oracle_clip(clip, p, kw)
    assert np.array_equal(out.cpu().numpy(), rf)


@pytest.mark.parametrize("channels", [1, 2, 4])
def test_other_channel_counts_use_the_generic_kernel(channels, cuda_device):
    mean, std = [0.4, 0.5, 0.45, 0.3][:channels], [0.2, 0.25, 0.225, 0.3][:channels]
    kw = full_kwargs(dict(min_short_side_scale=40, crop_size=36, num_frames=4, slowfast_alpha=2, mean=mean, std=std))
    clip = synth_clip((channels, 6, 44, 60), channels)
    t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact")
    p = t.sample_params(44, 60)
    slow, fast = t.explicit(clip.cuda(), p)
    assert _lib.last_launch_info()["kernel"] == 1
    rs, rf = oracle_clip(clip, p, kw)
    assert np.array_equal(fast.cpu().numpy(), rf) and np.array_equal(slow.cpu().numpy(), rs)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("norm_mode", ["exact", "affine"])
def test_tiled_and_generic_agree(dtype, norm_mode, cuda_device):
    """EXACT: both kernels evaluate the same arithmetic -> identical bits.  AFFINE: the tiled kernel
    folds constants differently (documented) -> equal to a few fp32 ulps before the output rounding."""
    kw = full_kwargs(dict(num_frames=8, training=True, mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225]))
    clips = torch.stack([synth_clip((3, 16, 256, 340), 50 + i) for i in range(3)]).cuda()
    torch.manual_seed(8)
    outs = {}
    for kernel in ("generic", "tiled"):
        t = pva_b200.make_transform(**kw, out_dtype=dtype, norm_mode=norm_mode, kernel=kernel)
        if kernel == "generic":
            params = [t.sample_params(256, 340) for _ in range(3)]
        outs[kernel] = [o.float() for o in t.run_batch(clips, params)]
    for a, b in zip(outs["generic"], outs["tiled"]):
        if norm_mode == "exact":
            assert torch.equal(a, b)
        else:
            tol = {torch.float32: 4e-6, torch.float16: 2e-3, torch.bfloat16: 1.6e-2}[dtype]  # <= 1 output ulp
            assert (a - b).abs().max().item() <= tol
            assert (a != b).float().mean().item() < 0.01 or dtype == torch.float32


def test_extreme_windows_at_the_frame_corners(cuda_device):
    """Crop windows touching every edge (clamped taps on the last source row / column), both flips."""
    kw = full_kwargs(dict(num_frames=4, slowfast_alpha=2))
    clip = synth_clip((3, 4, 256, 340), 77)
    t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel="tiled")
    for size in (256, 320):
        oh, ow = pipeline.resized_hw(256, 340, size)
        for top, left in ((0, 0), (oh - 256, ow - 256), (0, ow - 256), (oh - 256, 0)):
            for flip in (False, True):
                p = ClipParams(size, oh, ow, top, left, flip)
                slow, fast = t.explicit(clip.cuda(), p)
                rs, rf = oracle_clip(clip, p, kw)
                assert np.array_equal(fast.cpu().numpy(), rf), (size, top, left, flip)
                assert np.array_equal(slow.cpu().numpy(), rs)


@pytest.mark.parametrize("kernel", ["generic", "tiled"])
def test_maximum_frame_counts(kernel, cuda_device):
    """n_fast = 128 (the ABI's limit) out of a 300-frame clip, alpha = 1 so that n_slow = 128 as well."""
    kw = full_kwargs(dict(min_short_side_scale=32, crop_size=32, num_frames=128, slowfast_alpha=1))
    clip = synth_clip((3, 300, 36, 48), 4242)
    t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel=kernel)
    p = t.sample_params(36, 48)
    slow, fast = t.explicit(clip.cuda(), p)
    assert fast.shape == (3, 128, 32, 32) and slow.shape == (3, 128, 32, 32)
    rs, rf = oracle_clip(clip, p, kw)
    assert np.array_equal(fast.cpu().numpy(), rf) and np.array_equal(slow.cpu().numpy(), rs)
    with pytest.raises(NotImplementedError, match="n_fast"):
        pva_b200.make_transform(**dict(kw, num_frames=129), kernel=kernel).explicit(clip.cuda(), p)


def test_environment_cannot_switch_off_stores_or_fences(cuda_device):
    """PVA_TILED_DEBUG / PVA_TILED_PDL used to be read on every launch of the shipped library (1 = drop
    every store, 16 = drop the proxy fence, PDL=1 = weaker ordering).  They are compiled out now: with
    them set, a fresh process gets bit-identical output."""
    import subprocess
    import sys

    code = (
        "import sys, hashlib, torch; sys.path.insert(0, %r); import pva_b200\n"
        "g = torch.Generator().manual_seed(3)\n"
        "clips = torch.randint(0, 256, (2, 3, 16, 128, 172), dtype=torch.uint8, generator=g).cuda()\n"
        "t = pva_b200.make_transform(min_short_side_scale=128, max_short_side_scale=160, crop_size=128, num_frames=8,"
        " training=True, kernel='tiled')\n"
        "torch.manual_seed(5); params = [t.sample_params(128, 172) for _ in range(2)]\n"
        "slow, fast = t.run_batch(clips, params); torch.cuda.synchronize()\n"
        "print(hashlib.sha256(fast.cpu().numpy().tobytes() + slow.cpu().numpy().tobytes()).hexdigest())\n"
    ) % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    digests = []
    for extra in ({}, {"PVA_TILED_DEBUG": "1"}, {"PVA_TILED_DEBUG": "16", "PVA_TILED_PDL": "1", "PVA_TILED_RPS": "2"}):
        env = {k: v for k, v in os.environ.items() if not k.startswith("PVA_")}
        env.update(extra)
        out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True).stdout
        digests.append(out.strip().splitlines()[-1])
    assert digests[0] == digests[1] == digests[2]


@pytest.mark.parametrize("kernel", ["generic", "tiled"])
def test_clip_at_the_edges_of_its_allocation(kernel, cuda_device):
    """The band copies of the tiled kernel start at the 16-byte granule of their first byte and end at
    the granule of their last one: only granules that hold at least one byte of the clip are read.
    Allocation (and mapping) boundaries are aligned far more coarsely, so a clip that starts at the
    first byte or ends at the last byte of a cudaMalloc'ed block is safe.  Exercise both, with corner
    crop windows whose taps touch the clip's first and last bytes."""
    kw = full_kwargs(dict(num_frames=4, slowfast_alpha=2, crop_size=128, min_short_side_scale=128))
    t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel=kernel)
    shape = (3, 6, 130, 171)
    clip = synth_clip(shape, 4242)
    n = clip.numel()
    block_bytes = 32 << 20  # a whole number of 2 MiB pages: the caching allocator hands out the cudaMalloc block itself
    oh, ow = pipeline.resized_hw(130, 171, 128)
    windows = [ClipParams(128, oh, ow, 0, 0, False), ClipParams(128, oh, ow, oh - 128, ow - 128, True)]
    for where in ("start", "end"):
        block = torch.empty(block_bytes, dtype=torch.uint8, device="cuda")
        off = 0 if where == "start" else block_bytes - n
        dev = block[off:off + n].view(shape)
        dev.copy_(clip)
        for p in windows:
            slow, fast = t.run_batch([dev], [p])
            torch.cuda.synchronize()
            rs, rf = oracle_clip(clip, p, kw)
            assert np.array_equal(fast[0].cpu().numpy(), rf) and np.array_equal(slow[0].cpu().numpy(), rs)
        del dev, block


@pytest.mark.parametrize("kernel", ["generic", "tiled"])
@pytest.mark.parametrize("channels_last", [True, False])
def test_slow_table_with_repeated_frames(kernel, channels_last, cuda_device):
    """PackPathway's table never repeats a frame, but the C ABI takes any slow_idx: a fast frame that
    feeds several slow slots must be written to each of them (the tiled kernel walks an inverted table)."""
    from pva_b200.transforms import _IdxTable

    kw = full_kwargs(dict(num_frames=4, slowfast_alpha=2, crop_size=128, min_short_side_scale=128))
    t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel=kernel,
                                channels_last=channels_last)
    t.n_slow, t._slow_tab = 5, _IdxTable([0, 0, 3, 1, 3])
    clip = synth_clip((3, 6, 130, 171), 77)
    p = t.sample_params(130, 171)
    slow, fast = t.run_batch([clip.cuda()], [p])
    torch.cuda.synchronize()
    assert slow.shape == (1, 3, 5, 128, 128)
    assert torch.equal(slow[0], fast[0][:, [0, 0, 3, 1, 3]])
    (rf,) = c_oracle_explicit(clip, (p.size, p.top, p.left, p.flip), dict(kw, is_slowfast=False))
    assert np.array_equal(fast[0].cpu().numpy(), rf)
