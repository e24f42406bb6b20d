#!/usr/bin/env python
"""End-to-end training step, for CONTEXT (BASELINE.json configs[4]).

Mirrors the reference's inner loop (``run.py:243-261``): forward on ``batch["video"]``, cross
entropy, ``loss / accum``, backward (DDP all-reduce every micro-step, as the reference never uses
``no_sync``), optimiser step every ``accum`` micro-steps, fp16 autocast + GradScaler (what
``--mixed_precision fp16`` makes accelerate do).  accelerate itself is absent here, so the launcher is
``torchrun`` and DDP is used directly.  Synthetic Kinetics-700-shaped data: uint8 64x256x340 clips,
700 classes, batch 8 per rank.

Two feeders are timed against the same model:

  --feed gpu   DataLoader workers only draw (SampleParams) and ship raw uint8 clips; the fused
               kernel preprocesses on the rank's GPU (GpuPreprocess)
  --feed cpu   the reference's way: workers run the whole CPU transform (oracle restatement) and
               ship fp32 [slow, fast]
  --feed none  model only, on a resident preprocessed batch (upper bound)

Prints one JSON line per run (rank 0).  Nothing here is part of the preprocessing path or of
bench.py's metric; the model and the loop are stock PyTorch.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from slowfast_r50 import SlowFastR50  # noqa: E402

CLIP_SHAPE = (3, 64, 256, 340)


class SyntheticClips(torch.utils.data.Dataset):
    """Map-style stand-in for LimitDataset(Kinetics(...)) (run.py:151-168): one synthetic decoded clip per
    worker, re-used for every index, so that only the transform is measured, not decoding."""

    def __init__(self, n, transform, num_classes=700):
        self.n, self.transform, self.num_classes, self._clip = n, transform, num_classes, None

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        if self._clip is None:
            g = torch.Generator().manual_seed(1234 + os.getpid())
            self._clip = torch.randint(0, 256, CLIP_SHAPE, dtype=torch.uint8, generator=g)
        return self.transform({"video": self._clip, "label": i % self.num_classes})


def micro_step(model, opt, scaler, video, label, step, accum):
    """One micro-step of the reference's inner loop (run.py:251-261): fp16-autocast forward on
    ``batch["video"]``, cross entropy, ``loss / accum``, scaled backward, optimiser step when
    ``step % accum == 0`` (which fires at step 0 too, as in the reference).  Returns (loss, logits)."""
    with torch.autocast("cuda", dtype=torch.float16):
        logits = model(video)
        loss = F.cross_entropy(logits, label) / accum
    scaler.scale(loss).backward()
    if step % accum == 0:  # run.py:258
        scaler.step(opt)
        scaler.update()
        opt.zero_grad(set_to_none=True)
    return loss.detach(), logits.detach()


def rank_batch_sampler(n_batches, batch, rank, world):
    """Index batches of this rank: accelerate's per-process batch sharding (run.py:196-198), stated
    by pva_b200.batches_for_rank -- rank r takes global batches r, r + world, ..."""
    import pva_b200

    return [[b * batch + j for j in range(batch)] for b in pva_b200.batches_for_rank(n_batches, rank, world)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--feed", choices=["gpu", "cpu", "none"], default="gpu")
    ap.add_argument("--steps", type=int, default=24, help="timed micro-steps")
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--accum", type=int, default=4)
    ap.add_argument("--workers", type=int, default=0, help="DataLoader workers per rank (0 = cores / ranks)")
    ap.add_argument("--read-in-place", dest="read_in_place", action="store_true",
                    help="--feed gpu: the kernel reads the pinned batch in place instead of a strided upload")
    args = ap.parse_args()

    import pva_b200

    rank, world, local = pva_b200.rank_and_world()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=dev)
    torch.manual_seed(pva_b200.rank_seed(42, rank))

    model = SlowFastR50(700).to(dev).to(memory_format=torch.channels_last_3d)
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)  # run.py:192
    scaler = torch.amp.GradScaler("cuda")

    cores = len(os.sched_getaffinity(0))
    workers = args.workers or max(1, cores // world)
    n_batches = (args.steps + args.warmup + 4) * world  # global batches; every rank takes its own share
    n_items = n_batches * args.batch
    batches = rank_batch_sampler(n_batches, args.batch, rank, world)

    kw = dict(num_frames=32, slowfast_alpha=4, training=True)  # run_slowfast_r50.sh
    t = pva_b200.make_transform(**kw, device=dev)
    to_device = None
    if args.feed == "gpu":
        ds = SyntheticClips(n_items, pva_b200.SampleParams(t))
        loader = torch.utils.data.DataLoader(ds, batch_sampler=batches, num_workers=workers, pin_memory=True,
                                             collate_fn=pva_b200.collate_raw_clips, persistent_workers=True,
                                             prefetch_factor=4)
        # --read-in-place: no H2D copy, the kernel reads the pinned batch over PCIe on 8 SMs
        to_device = pva_b200.GpuPreprocess(t, dev, upload_selected_frames=not args.read_in_place,
                                           read_in_place=args.read_in_place, max_ctas=8)
    elif args.feed == "cpu":
        from bench import reference_cpu_transform  # bench.py's CPU-baseline leg: the reference's transform

        ds = SyntheticClips(n_items, reference_cpu_transform(**kw))
        loader = torch.utils.data.DataLoader(ds, batch_sampler=batches, num_workers=workers, pin_memory=True,
                                             persistent_workers=True, prefetch_factor=2)
    else:
        g = torch.Generator().manual_seed(7)
        clips = torch.randint(0, 256, (args.batch,) + CLIP_SHAPE, dtype=torch.uint8, generator=g).to(dev)
        params = [t.sample_params(256, 340) for _ in range(args.batch)]
        resident = t.run_batch(clips, params)
        labels = torch.randint(0, 700, (args.batch,), device=dev)
        loader = [{"video": resident, "label": labels}] * (args.steps + args.warmup)

    def step_fn(step, batch):
        if to_device is not None:
            batch = to_device(batch)
        video = [v.to(dev, non_blocking=True) for v in batch["video"]]
        label = batch["label"].to(dev, non_blocking=True)
        micro_step(model, opt, scaler, video, label, step, args.accum)

    it = iter(loader)
    for s in range(args.warmup):
        step_fn(s, next(it))
    torch.cuda.synchronize(dev)
    if world > 1:
        torch.distributed.barrier()
    t0 = time.perf_counter()
    for s in range(args.steps):
        step_fn(args.warmup + s, next(it))
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    if world > 1:
        tt = torch.tensor([dt], device=dev, dtype=torch.float64)
        torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
        dt = float(tt.item())
    if rank == 0:
        print(json.dumps({"feed": args.feed, "n_gpus": world, "batch_per_gpu": args.batch, "accum": args.accum,
                          "micro_steps": args.steps, "ms_per_micro_step": 1000 * dt / args.steps,
                          "clips_per_s": world * args.batch * args.steps / dt, "workers_per_rank": workers,
                          "host_cores": cores, "read_in_place": bool(args.read_in_place and args.feed == "gpu"), "model": "SlowFast-R50 (local restatement, random init), fp16 autocast"}),
              flush=True)
    del it
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
