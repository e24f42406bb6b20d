"""Generate the golden vectors under tests/golden/ from the reference's own code.

Run in the BUILD CONTAINER only (needs /root/reference):

    python tests/golden/make_golden.py

It imports ``/root/reference/run.py`` unmodified through
``oracle.reference_loader`` (stub modules for the absent third-party packages),
calls ``run.make_transform(...)`` -- the reference's public entry for this path,
``run.py:68-102`` -- on seeded synthetic uint8 clips, and freezes inputs, the
RNG seed, and outputs.  The reference has no tests or fixtures of its own for
this path, so these files are the known-answer vectors the
oracle and the CUDA path are held to.

Files written:
  golden_small.npz   full fp32 outputs of 10 small transform cases
  golden_large.json  sha256 + strided samples of 3 realistic-size cases
                     (inputs are regenerated from the seed)
  golden_tables.json temporal index tables, resized sizes, centre-crop offsets
"""
from __future__ import annotations

import hashlib
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import reference_loader  # noqa: E402

SMALL_CASES = [
    # name, (C,T,H,W), kwargs, seed
    ("val_landscape", (3, 20, 48, 64), dict(min_short_side_scale=32, crop_size=32, num_frames=8, training=False), 11),
    ("val_portrait", (3, 16, 64, 48), dict(min_short_side_scale=40, crop_size=36, num_frames=8, training=False), 12),
    ("val_odd_width", (3, 9, 50, 67), dict(min_short_side_scale=48, crop_size=40, num_frames=4, training=False), 13),
    ("val_single_pathway", (3, 16, 48, 64), dict(min_short_side_scale=40, crop_size=36, num_frames=8, is_slowfast=False, training=False), 14),
    ("train_landscape", (3, 20, 48, 64), dict(min_short_side_scale=32, max_short_side_scale=40, crop_size=32, num_frames=8, training=True), 15),
    ("train_portrait", (3, 12, 64, 48), dict(min_short_side_scale=36, max_short_side_scale=48, crop_size=32, num_frames=8, training=True), 16),
    ("train_flip_always", (3, 12, 48, 64), dict(min_short_side_scale=32, max_short_side_scale=40, crop_size=32, num_frames=8, horizontal_flip_p=1.0, training=True), 17),
    ("train_square_nodraw", (3, 8, 40, 40), dict(min_short_side_scale=32, max_short_side_scale=32, crop_size=32, num_frames=4, training=True), 18),
    ("train_downscale", (3, 8, 90, 120), dict(min_short_side_scale=40, max_short_side_scale=56, crop_size=36, num_frames=8, slowfast_alpha=2, training=True), 19),
    ("train_single_pathway_meanstd", (3, 10, 48, 64), dict(min_short_side_scale=40, max_short_side_scale=48, crop_size=36, num_frames=5, is_slowfast=False, mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225], training=True), 20),
]

LARGE_CASES = [
    ("val_256x340", (3, 64, 256, 340), dict(num_frames=8, slowfast_alpha=4, training=False), 101),
    ("train_256x340", (3, 64, 256, 340), dict(num_frames=8, slowfast_alpha=4, training=True), 102),
    ("train_single_224", (3, 32, 256, 340), dict(num_frames=4, crop_size=224, is_slowfast=False, training=True), 103),
]


def synth_clip(shape, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, shape, dtype=torch.uint8, generator=g)


def run_reference(run, shape, kwargs, seed):
    clip = synth_clip(shape, seed)
    transform = run.make_transform(**kwargs)
    torch.manual_seed(seed)
    sample = {"video": clip.clone(), "audio": torch.zeros(4), "label": 7, "video_name": "synthetic"}
    out = transform(sample)
    assert out is sample and "audio" not in out and out["label"] == 7
    video = out["video"]
    if isinstance(video, list):
        return clip, [v.contiguous() for v in video]
    return clip, [video.contiguous()]


def main():
    run = reference_loader.load_reference_run()
    torch.set_num_threads(1)

    arrays = {}
    meta = {}
    for name, shape, kwargs, seed in SMALL_CASES:
        clip, outs = run_reference(run, shape, kwargs, seed)
        arrays[f"{name}/in"] = clip.numpy()
        for i, o in enumerate(outs):
            arrays[f"{name}/out{i}"] = o.numpy()
        meta[name] = dict(shape=list(shape), kwargs=kwargs, seed=seed, n_out=len(outs))
    np.savez_compressed(os.path.join(HERE, "golden_small.npz"), **arrays)
    with open(os.path.join(HERE, "golden_small.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)

    large = {}
    for name, shape, kwargs, seed in LARGE_CASES:
        clip, outs = run_reference(run, shape, kwargs, seed)
        entry = dict(shape=list(shape), kwargs=kwargs, seed=seed, outs=[])
        entry["in_sha256"] = hashlib.sha256(clip.numpy().tobytes()).hexdigest()
        for o in outs:
            a = o.numpy()
            flat = a.reshape(-1)
            stride = max(1, flat.size // 2048)
            entry["outs"].append(dict(
                shape=list(a.shape),
                sha256=hashlib.sha256(a.tobytes()).hexdigest(),
                sample_stride=stride,
                samples=[float(x) for x in flat[::stride][:2048]],
            ))
        large[name] = entry
    with open(os.path.join(HERE, "golden_large.json"), "w") as f:
        json.dump(large, f, indent=1, sort_keys=True)

    # Tables: produced by the same torch / torchvision calls the reference makes.
    from torchvision.transforms import functional as TF
    tables = {"temporal": {}, "pack": {}, "resized": [], "center_crop_left": {}}
    for t, n in [(64, 32), (80, 16), (64, 16), (64, 8), (16, 4), (8, 2), (5, 1), (3, 15), (300, 128), (32, 32), (250, 32)]:
        probe = torch.arange(t).view(1, t, 1, 1).float()
        sub = run.UniformTemporalSubsample(n)(probe)
        tables["temporal"][f"{t}->{n}"] = [int(v) for v in sub.flatten().tolist()]
    for t, alpha in [(32, 4), (16, 4), (8, 4), (8, 2), (32, 8), (5, 2), (64, 4)]:
        probe = torch.arange(t).view(1, t, 1, 1).float().expand(1, t, 1, 1)
        slow, fast = run.PackPathway(alpha)(probe)
        assert fast is probe
        tables["pack"][f"{t}/{alpha}"] = [int(v) for v in slow.flatten().tolist()]
    for (h, w, s) in [(256, 340, 256), (256, 340, 263), (256, 340, 288), (256, 340, 320), (120, 146, 300),
                      (340, 256, 300), (360, 480, 282), (240, 320, 256), (1080, 1920, 320), (255, 255, 256), (97, 131, 77)]:
        x = torch.zeros(1, 1, h, w)
        y = run.ShortSideScale(s)(x)
        tables["resized"].append([h, w, s, int(y.shape[-2]), int(y.shape[-1])])
    for w in [340, 341, 342, 343, 345, 347, 425, 256, 257]:
        probe = torch.arange(w).view(1, 1, w).expand(1, 256, w).float()
        c = TF.center_crop(probe, [256, 256])
        tables["center_crop_left"][str(w)] = int(c[0, 0, 0].item())
    with open(os.path.join(HERE, "golden_tables.json"), "w") as f:
        json.dump(tables, f, indent=1, sort_keys=True)
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
