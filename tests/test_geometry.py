"""Host-side geometry and draws of the product against the reference-derived tables (CPU)."""
import pytest
import torch

import pva_b200
from oracle import pipeline
from pva_b200 import geometry

from _util import load_tables


def test_temporal_and_pack_tables():
    tabs = load_tables()
    for key, want in tabs["temporal"].items():
        t, n = (int(v) for v in key.split("->"))
        assert list(geometry.temporal_indices(t, n)) == want
    for key, want in tabs["pack"].items():
        t, alpha = (int(v) for v in key.split("/"))
        assert list(geometry.pack_indices(t, alpha)) == want
    for n in range(1, 129):  # SampleParams(subsample_on_worker=True) relies on this identity
        assert list(geometry.temporal_indices(n, n)) == list(range(n))


def test_resized_and_center_crop():
    tabs = load_tables()
    for h, w, s, oh, ow in tabs["resized"]:
        assert geometry.resized_hw(h, w, s) == (oh, ow)
    for w, left in tabs["center_crop_left"].items():
        assert geometry.center_crop_offsets(256, int(w), 256) == (0, left)
    p = geometry.val_params(256, 340, 256, 256)
    assert (p.size, p.oh, p.ow, p.top, p.left, p.flip) == (256, 256, 340, 0, 42, False)


@pytest.mark.parametrize("hw", [(256, 340), (340, 256), (240, 320), (360, 480), (256, 256), (300, 300)])
def test_train_draws_replay_the_reference_order(hw):
    h, w = hw
    for seed in range(20):
        torch.manual_seed(seed)
        want = pipeline.replay_draws(h, w, 256, 320, 256, 0.5)
        after_oracle = torch.rand(1)
        torch.manual_seed(seed)
        got = geometry.sample_train_params(h, w, 256, 320, 256, 0.5)
        after_product = torch.rand(1)
        assert (got.size, got.top, got.left, got.flip) == want
        assert (got.oh, got.ow) == pipeline.resized_hw(h, w, got.size)
        assert torch.equal(after_oracle, after_product)  # consumed exactly the same RNG stream


def test_square_equal_to_crop_makes_no_crop_draws():
    torch.manual_seed(0)
    a = geometry.sample_train_params(256, 256, 256, 256, 256, 0.5)
    assert (a.top, a.left) == (0, 0)
    torch.manual_seed(0)
    torch.randint(256, 257, (1,))
    assert a.flip == bool(torch.rand(1) < 0.5)


def test_crop_too_large_raises_value_error():
    with pytest.raises(ValueError, match="larger than input image size"):
        geometry.val_params(100, 200, 64, 128)
    with pytest.raises(ValueError, match="larger than input image size"):
        geometry.sample_train_params(100, 200, 64, 72, 128, 0.5)


def test_params_row_roundtrip():
    p = geometry.ClipParams(300, 300, 398, 7, 99, True)
    assert geometry.ClipParams.from_row(p.as_row()) == p


def test_sample_params_and_collate_cpu_side():
    """Worker-side hand-over needs no GPU: draws, raw uint8 clips, int32 table."""
    t = pva_b200.make_transform(training=True, num_frames=8, crop_size=64, min_short_side_scale=64,
                                max_short_side_scale=80)
    sp = pva_b200.SampleParams(t, subsample_on_worker=True)
    samples = []
    torch.manual_seed(3)
    for i in range(3):
        clip = torch.randint(0, 256, (3, 16, 64, 85), dtype=torch.uint8)
        samples.append(sp({"video": clip, "audio": 1, "label": i}))
    assert all("audio" not in s and s["video"].dtype == torch.uint8 and s["video"].shape[1] == 8 for s in samples)
    batch = pva_b200.collate_raw_clips(samples)
    assert batch["video"].shape == (3, 3, 8, 64, 85) and batch["video"].dtype == torch.uint8
    assert batch["pva_params"].shape == (3, 6) and batch["pva_params"].dtype == torch.int32
    assert batch["label"].tolist() == [0, 1, 2]
    ragged = [dict(s) for s in samples]
    ragged[1]["video"] = ragged[1]["video"][:, :, :50]
    assert isinstance(pva_b200.collate_raw_clips(ragged)["video"], list)


def test_sample_params_narrows_decoder_style_float_frames_only_when_lossless():
    """pytorchvideo's decoder hands over float32 frames holding whole byte values."""
    t = pva_b200.make_transform(training=False, num_frames=4, crop_size=32, min_short_side_scale=32)
    raw = torch.randint(0, 256, (3, 8, 40, 52), dtype=torch.uint8)
    out = pva_b200.SampleParams(t)({"video": raw.float()})["video"]
    assert out.dtype == torch.uint8 and torch.equal(out, raw)
    for bad in (raw.float() + 0.25, raw.float() - 1.0, raw.float() * 1.5):  # fractional / below 0 / above 255
        kept = pva_b200.SampleParams(t)({"video": bad.clone()})["video"]
        assert kept.dtype == torch.float32 and torch.equal(kept, bad)
    kept = pva_b200.SampleParams(t, narrow_float_frames=False)({"video": raw.float()})["video"]
    assert kept.dtype == torch.float32
    sub = pva_b200.SampleParams(t, subsample_on_worker=True)({"video": raw.float()})["video"]
    assert sub.dtype == torch.uint8 and sub.shape[1] == 4


def test_transform_validates_before_any_launch():
    t = pva_b200.make_transform()
    with pytest.raises(TypeError):
        t({"video": torch.zeros(3, 8, 32, 32, dtype=torch.float64)})  # only uint8 / float32 frames
    with pytest.raises(ValueError):
        t({"video": torch.zeros(8, 32, 32, dtype=torch.uint8)})  # not 4-D
    with pytest.raises(ValueError):
        t({"video": torch.zeros(1, 8, 32, 32, dtype=torch.uint8)})  # channel count != len(mean)
    with pytest.raises(ValueError, match="larger than input image size"):
        pva_b200.make_transform(min_short_side_scale=16, crop_size=32)({"video": torch.zeros(3, 8, 40, 60, dtype=torch.uint8)})
