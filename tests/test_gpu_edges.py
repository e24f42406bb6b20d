"""Edge cases of the CUDA path: launch chunking, ragged index tables, odd crops, strong
down-scaling, channel counts, and tiled-vs-generic equality.  Needs a B200: -m gpu."""
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


def test_column_cropped_tensor_map_loads_give_the_same_bits(cuda_device, monkeypatch):
    """PVA_TILED_COLCROP=1 (experimental load path: four tensor maps per clip, only the window's
    columns are fetched) must not change a single bit: frame corners, both flips, ragged batch,
    clips at addresses that are only 4-byte aligned, H not a multiple of 4."""
    kw = full_kwargs(dict(num_frames=4, slowfast_alpha=2))
    t = pva_b200.make_transform(**kw, out_dtype=torch.float32, norm_mode="exact", kernel="tiled")
    clips, params = [], []
    for i, (h, size, off) in enumerate(((256, 256, 0), (256, 320, 4), (258, 300, 8), (256, 288, 12))):
        c = synth_clip((3, 4 + (i % 2) * 2, h, 340), 900 + i)
        oh, ow = pipeline.resized_hw(h, 340, size)
        for top, left, flip in ((0, 0, False), (oh - 256, ow - 256, True), ((oh - 256) // 2, (ow - 256) // 3, False)):
            n = c.numel()
            dev = torch.empty(n + 16, dtype=torch.uint8, device="cuda")[off:off + n].view(c.shape)  # src % 16 == off
            dev.copy_(c)
            clips.append(dev)
            params.append(ClipParams(size, oh, ow, top, left, flip))
    monkeypatch.setenv("PVA_TILED_COLCROP", "0")
    ref = [o.clone() for o in t.run_batch(clips, params)]
    monkeypatch.setenv("PVA_TILED_COLCROP", "1")
    got = t.run_batch(clips, params)
    torch.cuda.synchronize()
    for a, b in zip(ref, got):
        assert torch.equal(a, b)
    # and the production mode on a full-size same-shape batch (the benchmark's shape)
    t16 = pva_b200.make_transform(**full_kwargs(dict(num_frames=32, slowfast_alpha=4)), kernel="tiled")
    batch = torch.stack([synth_clip((3, 64, 256, 340), 40 + i) for i in range(2)]).cuda()
    oh, ow = pipeline.resized_hw(256, 340, 297)
    ps = [ClipParams(297, oh, ow, 11, 23, False), ClipParams(297, oh, ow, 0, ow - 256, True)]
    monkeypatch.setenv("PVA_TILED_COLCROP", "0")
    ref = [o.clone() for o in t16.run_batch(batch, ps)]
    smem_linear = _lib.last_launch_info()["smem_bytes"]
    monkeypatch.setenv("PVA_TILED_COLCROP", "1")
    got = t16.run_batch(batch, ps)
    torch.cuda.synchronize()
    assert _lib.last_launch_info()["smem_bytes"] < smem_linear  # the cropped stages are smaller: the path was taken
    for a, b in zip(ref, got):
        assert torch.equal(a, b)
