"""the local SlowFast-R50 (run.py:105-110) and the training micro-step
(run.py:243-261) that consume the kernel's output.  Context components -- stock PyTorch -- but the drop-in
claim is only worth something if they accept what the kernel writes.  Needs a B200: -m gpu."""
import os
import sys

import numpy as np
import pytest
import torch

import pva_b200
from oracle import pipeline

from _util import synth_clip

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "examples"))

pytestmark = pytest.mark.gpu


def test_slowfast_r50_shape_contract_and_size():
    """CPU-side facts (no kernel): head width 2304 = 2048 + 256 (run.py:109), parameter count of the
    published architecture (pytorchvideo's model zoo lists 34.57 M for slowfast_r50 at 400 classes)."""
    from slowfast_r50 import SlowFastR50

    m400, m700 = SlowFastR50(400), SlowFastR50(700)
    assert m700.proj.in_features == 2304 and m700.proj.out_features == 700
    assert sum(p.numel() for p in m400.parameters()) == 34_566_488
    assert sum(p.numel() for p in m700.parameters()) == 34_566_488 + 300 * 2305


def test_slowfast_r50_takes_the_kernel_output(cuda_device):
    """The model consumes exactly what the kernel writes: fp16, channels_last_3d, [slow B x 3 x 8 x S x S,
    fast B x 3 x 32 x S x S] -- no conversion, no copy in between."""
    from slowfast_r50 import SlowFastR50

    t = pva_b200.make_transform(min_short_side_scale=64, max_short_side_scale=80, crop_size=64, training=True)
    clips = torch.stack([synth_clip((3, 64, 72, 96), 70 + i) for i in range(2)]).cuda()
    torch.manual_seed(3)
    slow, fast = t.run_batch(clips, [t.sample_params(72, 96) for _ in range(2)])
    assert slow.shape == (2, 3, 8, 64, 64) and fast.shape == (2, 3, 32, 64, 64) and fast.dtype == torch.float16
    assert fast.is_contiguous(memory_format=torch.channels_last_3d)
    model = SlowFastR50(700).cuda().to(memory_format=torch.channels_last_3d).eval()
    seen = {}
    model.proj.register_forward_hook(lambda mod, inp, out: seen.update(width=inp[0].shape[-1]))
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
        logits = model([slow, fast])
    assert logits.shape == (2, 700) and torch.isfinite(logits).all()
    assert seen["width"] == 2304


def test_micro_step_gpu_feed_matches_cpu_reference_feed(cuda_device):
    """The one end-to-end parity check of the drop-in claim: one training micro-step (run.py:251-261) fed by
    SampleParams -> collate_raw_clips -> GpuPreprocess, and one fed by the CPU reference transform with
    default_collate, from the same seed (hence the same draws) and the same initial weights.  The two
    batches differ by fp16 rounding of the inputs only (<= 1e-3), so logits and loss must agree to fp16
    training precision and both losses must be finite."""
    from slowfast_r50 import SlowFastR50
    from torch.utils.data import default_collate
    from train_step_bench import micro_step

    kw = dict(min_short_side_scale=64, max_short_side_scale=80, crop_size=64, num_frames=32, slowfast_alpha=4,
              training=True)
    raw = [synth_clip((3, 64, 72, 96), 900 + i) for i in range(4)]
    labels = [3, 141, 592, 65]
    t = pva_b200.make_transform(**kw, device=cuda_device)

    torch.manual_seed(123)
    sp = pva_b200.SampleParams(t)
    gpu_batch = pva_b200.collate_raw_clips([sp({"video": c.clone(), "label": l}) for c, l in zip(raw, labels)])
    gpu_batch = pva_b200.GpuPreprocess(t, cuda_device)(gpu_batch)

    torch.manual_seed(123)
    cpu_t = pipeline.make_transform(**kw)
    cpu_batch = default_collate([cpu_t({"video": c.clone(), "label": l}) for c, l in zip(raw, labels)])

    for g, c in zip(gpu_batch["video"], cpu_batch["video"]):  # same draws: inputs within fp16 rounding
        assert g.shape == c.shape
        assert (g.float().cpu() - c).abs().max() <= 1e-3
    assert gpu_batch["label"].tolist() == cpu_batch["label"].tolist() == labels

    results = []
    for batch in (gpu_batch, cpu_batch):
        torch.manual_seed(7)
        model = SlowFastR50(700, dropout=0.0).to(cuda_device).to(memory_format=torch.channels_last_3d)
        opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
        scaler = torch.amp.GradScaler("cuda")
        video = [v.to(cuda_device) for v in batch["video"]]
        loss, logits = micro_step(model, opt, scaler, video, batch["label"].to(cuda_device), step=0, accum=4)
        grad_free = all(p.grad is None for p in model.parameters())  # step 0 steps and clears (run.py:258-261)
        results.append((float(loss), logits.float().cpu().numpy(), grad_free))
    (loss_g, logits_g, gf_g), (loss_c, logits_c, gf_c) = results
    assert np.isfinite(loss_g) and np.isfinite(loss_c) and gf_g and gf_c
    scale = np.abs(logits_c).max()
    assert np.abs(logits_g - logits_c).max() <= 0.02 * scale + 0.02, (np.abs(logits_g - logits_c).max(), scale)
    assert abs(loss_g - loss_c) <= 0.02 * abs(loss_c) + 1e-3
