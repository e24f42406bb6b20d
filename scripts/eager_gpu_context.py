#!/usr/bin/env python
"""CONTEXT number ("unfused GPU"): the reference's op chain written with stock
PyTorch ops on a cuda tensor, one clip at a time as the reference's transform sees them, timed beside
the fused kernel on the same clips and the same draws.

    python scripts/eager_gpu_context.py [--workload slowfast_train_b8] [--iters 30]

The chain is the one of run.py:81-100 (index_select -> /255 -> (x-mean)/std -> F.interpolate bilinear ->
crop window -> flip -> PackPathway index_select), fp32 throughout like the reference, outputs contiguous
C x T x S x S (the fused kernel writes fp16 channels-last, its production form; values are compared
through the logical C x T x S x S views).  It is stock-library code, not part of the product path,
and nothing here is a bench value: it says how much of the fused kernel's speed comes from fusion
rather than from "being on a GPU".
Also prints the largest difference between the two (fused output is fp16).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import pva_b200  # noqa: E402
from bench import WORKLOADS  # noqa: E402
from pva_b200 import geometry  # noqa: E402


def eager_clip(video, p, t, idx_fast, idx_slow, mean, std):
    """One clip through stock ops.  video: uint8 C x T x H x W on the GPU."""
    x = video.index_select(1, idx_fast)                       # UniformTemporalSubsample
    x = x / 255.0                                             # Div255 (uint8 promotes to fp32)
    x = (x - mean) / std                                      # Normalize
    x = F.interpolate(x, size=(p.oh, p.ow), mode="bilinear", align_corners=False)  # ShortSideScale
    x = x[..., p.top:p.top + t.crop_size, p.left:p.left + t.crop_size]            # crop (view)
    if p.flip:
        x = x.flip(-1)                                        # RandomHorizontalFlip
    x = x.contiguous()
    if idx_slow is None:
        return [x]
    return [x.index_select(1, idx_slow), x]                   # PackPathway: slow first


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="slowfast_train_b8", choices=sorted(WORKLOADS))
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    args = ap.parse_args()

    dev = torch.device("cuda", 0)
    wl = WORKLOADS[args.workload]
    B, shape, kw = wl["batch"], wl["shape"], wl["kwargs"]
    t = pva_b200.make_transform(**kw, device=dev, out_dtype=torch.float16, channels_last=True)  # production layout
    gen = torch.Generator().manual_seed(1234)
    torch.manual_seed(42)
    n_sets = 4
    sets = []
    for _ in range(n_sets):
        clips = torch.randint(0, 256, (B,) + tuple(shape), dtype=torch.uint8, generator=gen).to(dev)
        params = [t.sample_params(shape[2], shape[3]) for _ in range(B)]
        sets.append((clips, params, t.prepare_batch(clips, params, clips_ready=True)))

    idx_fast = torch.tensor(geometry.temporal_indices(shape[1], t.num_frames), device=dev)
    idx_slow = None
    if t.n_slow:
        idx_slow = torch.tensor(geometry.pack_indices(t.num_frames, t.slowfast_alpha), device=dev)
    mean = torch.tensor(t.mean, device=dev).view(-1, 1, 1, 1)
    std = torch.tensor(t.std, device=dev).view(-1, 1, 1, 1)

    raw_stream = torch.cuda.current_stream(dev).cuda_stream

    def eager_batch(i):
        clips, params, _ = sets[i % n_sets]
        outs = [eager_clip(clips[b], params[b], t, idx_fast, idx_slow, mean, std) for b in range(B)]
        return [torch.stack(o) for o in zip(*outs)]           # default_collate's stacking

    def fused_batch(i):
        sets[i % n_sets][2].launch_raw(raw_stream)            # what bench.py times: one C-ABI call
        return sets[i % n_sets][2].outputs

    def timed(fn):
        for i in range(args.warmup):
            fn(i)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / args.iters

    ms_eager = timed(eager_batch)
    ms_fused = timed(fused_batch)

    ref = eager_batch(0)
    got = fused_batch(0)
    torch.cuda.synchronize(dev)
    got = got if isinstance(got, (list, tuple)) else [got]
    diff = max(float((g.float() - r).abs().max()) for g, r in zip(got, ref))

    print(json.dumps({
        "workload": args.workload, "batch": B, "iters": args.iters,
        "eager_gpu_ms_per_batch": ms_eager, "eager_gpu_clips_per_s": 1000.0 * B / ms_eager,
        "fused_ms_per_batch": ms_fused, "fused_clips_per_s": 1000.0 * B / ms_fused,
        "fused_over_eager": ms_eager / ms_fused,
        "max_abs_diff_fused_fp16_vs_eager_fp32": diff,
        "note": "eager = stock PyTorch ops on cuda, fp32, per clip + stack; context only",
    }), flush=True)


if __name__ == "__main__":
    main()
