#!/usr/bin/env python
"""Benchmark of the fused preprocessing path (BASELINE.json metric: preprocessed clips/s and
% of HBM roofline, beside the CPU reference).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl b200|reference]

A *step* is one pass of the hot path over one batch of synthetic uint8 clips: one
``pva_preprocess`` call through the C ABI (one kernel launch per step).  Under torchrun every
rank runs the same per-rank batch on its own GPU (weak scaling, no data-path collective --
SURVEY.md section 8e); rank 0 prints ONE JSON line.

``value`` / ``ms_per_step`` / ``roofline``: K plain launches back to back on one stream between two
CUDA events (barrier + synchronize on both sides).  Consecutive launches are chained by programmatic
dependent launch, so a launch starts reading its clips while the previous one drains: elapsed / K is
the STEADY-STATE time per launch (a throughput figure), not the duration of one launch in isolation.
The isolated duration is reported next to it (``launch_us_isolated``: events around a single launch
that is queued behind a filler kernel, so that no host time sits between the events; median of 50),
and so is the host cost of one call (``host_us_per_call``).

Keys beyond the base contract:
  roofline     dominant kernel: algorithmic bytes per launch / steady-state time per launch against
               MEASURED_PEAKS.json hbm_gbs; ``traffic`` = measured steady-state DRAM bytes per launch
               (profiles/traffic.json: ncu over consecutive launches on rotating buffers)
  cpu_baseline the oracle's CPU pipeline (restated reference transform) on the host cores, bounded
               samples, rank 0 at N=1 only: one single-thread process per core (``value``), one
               process x one thread (``one_thread``), and the reference's own execution model, a
               ``DataLoader(num_workers=cores, batch_size=B)`` with default_collate (``dataloader``)
  e2e          same metric through the public API with HOST buffers: pinned uint8 clips ->
               fused kernel (reads them in place over PCIe) -> D2H of [slow, fast] into pinned
               memory, every step; ``ceiling`` = the same bytes moved with no preprocessing at all
  overlapped_launches
               informational: the same K steps issued round-robin on two streams
  rank_ms_per_step
               N > 1: min / median / max over ranks of the per-rank time per step

``--impl reference`` times the reference's CPU implementation of the path (oracle port: the
reference itself is pure Python over pytorchvideo, which is not installable here) on all host
cores and prints the same line with ``"impl": "reference"``.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "preprocessed_clips_per_sec"
UNIT = "clips/s"
L2_BYTES = 126 * 1024 * 1024
NOMINAL_HBM_GBS = 8000.0
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md

WORKLOADS = {
    # BASELINE.json configs[1]: the configuration the metric is quoted on
    "slowfast_train_b8": dict(batch=8, shape=(3, 64, 256, 340),
                              kwargs=dict(num_frames=32, slowfast_alpha=4, training=True)),
    # configs[0]
    "slowfast_val_b8": dict(batch=8, shape=(3, 64, 256, 340),
                            kwargs=dict(num_frames=32, slowfast_alpha=4, training=False)),
    # configs[3]
    "single_pathway_b64": dict(batch=64, shape=(3, 64, 256, 340),
                               kwargs=dict(num_frames=16, crop_size=224, is_slowfast=False, training=True)),
}


def workload_geometry(wl):
    """(crop, n_fast, n_slow) of a workload from the reference factory's defaults (run.py:68-79)."""
    kw = wl["kwargs"]
    crop = kw.get("crop_size", 256)
    n_fast = kw.get("num_frames", 32)
    n_slow = n_fast // kw.get("slowfast_alpha", 4) if kw.get("is_slowfast", True) else 0
    return crop, n_fast, n_slow


def rotation_sets(wl):
    """How many distinct input/output buffer sets the device-resident region rotates over, and the
    bytes of one set: the working set must exceed L2 (timing rule)."""
    B, shape = wl["batch"], wl["shape"]
    crop, n_fast, n_slow = workload_geometry(wl)
    in_bytes = B * shape[0] * shape[1] * shape[2] * shape[3]
    out_bytes = B * shape[0] * (n_fast + n_slow) * crop * crop * 2
    n_sets = max(4, -(-3 * L2_BYTES // (in_bytes + out_bytes)))
    return n_sets, in_bytes, out_bytes


def workload_config(name, wl, world):
    """The `config` object: what is measured.  Identical in the b200 and the reference arm."""
    n_sets, in_bytes, out_bytes = rotation_sets(wl)
    return {"workload": name, "batch_per_gpu": wl["batch"], "global_batch": wl["batch"] * world,
            "clip_shape": list(wl["shape"]), "transform": wl["kwargs"],
            "parallelism": f"dp{world} (clips sharded per rank, no collective)",
            "l2": f"GPU arm: inputs and outputs rotate over {n_sets} sets = {(in_bytes + out_bytes) * n_sets >> 20} MiB > L2"}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.lower().startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


# --------------------------------------------------------------------------------------
# CPU arm (oracle port of the reference transform), duration-bounded
# --------------------------------------------------------------------------------------
def reference_cpu_transform(**kwargs):
    """The reference's CPU transform (oracle restatement of run.py:68-102) for the CPU-baseline legs.
    bench.py's cpu_baseline / reference arm is the one place outside tests/ and smoke() that may
    execute oracle/; examples/train_step_bench.py's CPU feeder goes through here as well."""
    from oracle import pipeline

    return pipeline.make_transform(**kwargs)


def _cpu_worker(widx, shape, kwargs, counter, go, stop):
    import torch

    torch.set_num_threads(1)  # what a DataLoader worker gets (torch/utils/data/_utils/worker.py)
    torch.manual_seed(42 + widx)
    g = torch.Generator().manual_seed(1234 + widx)
    clip = torch.randint(0, 256, shape, dtype=torch.uint8, generator=g)
    transform = reference_cpu_transform(**kwargs)
    transform({"video": clip})  # warm the allocator
    go.wait()
    while not stop.is_set():
        out = transform({"video": clip, "label": 0})
        assert out["video"] is not None
        with counter.get_lock():
            counter.value += 1


def cpu_reference_throughput(workload, seconds, warm_seconds=2.0, max_procs=64, procs=None):
    """clips/s of the reference CPU pipeline using one single-threaded process per host core
    (``procs=1``: one process x one thread, i.e. one DataLoader worker)."""
    import multiprocessing as mp

    ctx = mp.get_context("spawn")
    cores = host_cores()
    procs_n = procs if procs else max(1, min(cores, max_procs))
    counter = ctx.Value("l", 0)
    go, stop = ctx.Event(), ctx.Event()
    wl = WORKLOADS[workload]
    ps = [ctx.Process(target=_cpu_worker, args=(i, wl["shape"], wl["kwargs"], counter, go, stop), daemon=True)
          for i in range(procs_n)]
    for p in ps:
        p.start()
    time.sleep(0.5)
    go.set()
    # wait until every worker is warm: the counter starts moving
    t_wait = time.time()
    while counter.value < procs_n and time.time() - t_wait < 120:
        time.sleep(0.05)
    time.sleep(warm_seconds)
    c0, t0 = counter.value, time.perf_counter()
    time.sleep(seconds)
    c1, t1 = counter.value, time.perf_counter()
    stop.set()
    for p in ps:
        p.join(timeout=30)
        if p.is_alive():
            p.terminate()
    clips = c1 - c0
    return dict(value=clips / (t1 - t0), clips=clips, seconds=t1 - t0, cores=procs_n, host_cores=cores)


class _SyntheticClips:
    """In-memory stand-in for the reference's dataset (run.py:151-168 minus decode): every item is
    a raw uint8 clip pushed through the transform in the worker, like LabeledVideoDataset does."""

    def __init__(self, shape, kwargs, length):
        self.shape, self.kwargs, self.length = tuple(shape), dict(kwargs), length
        self.clip = None
        self.transform = None

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        import torch

        if self.clip is None:  # first touch inside the worker
            g = torch.Generator().manual_seed(1234 + (i % 97))
            self.clip = torch.randint(0, 256, self.shape, dtype=torch.uint8, generator=g)
            self.transform = reference_cpu_transform(**self.kwargs)
        return self.transform({"video": self.clip, "label": i % 700})


class _UntilStopped:
    """Sampler (main process) that deals indices until told to stop, so that the loader below can be
    run for a DURATION and still wind down by itself: once it stops, the workers finish what they
    were handed and the iterator ends normally."""

    def __init__(self):
        self.stop = False

    def __iter__(self):
        i = 0
        while not self.stop:
            yield i
            i += 1

    def __len__(self):
        return 1 << 30


def cpu_dataloader_throughput(workload, seconds, warm_batches=2):
    """clips/s of the reference's execution model (run.py:170-176): a torch DataLoader with one
    worker per host core, batch_size = the workload's batch, default_collate, batches handed to the
    main process through shared memory."""
    from torch.utils.data import DataLoader

    wl = WORKLOADS[workload]
    workers = host_cores()
    ds = _SyntheticClips(wl["shape"], wl["kwargs"], 1 << 30)
    sampler = _UntilStopped()
    loader = DataLoader(ds, batch_size=wl["batch"], sampler=sampler, num_workers=workers, pin_memory=False,
                        multiprocessing_context="spawn", persistent_workers=False)
    it = iter(loader)
    for _ in range(warm_batches + -(-workers // wl["batch"])):
        next(it)
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        b = next(it)
        v = b["video"]
        assert (v[0] if isinstance(v, (list, tuple)) else v).shape[0] == wl["batch"]
        n += wl["batch"]
    t1 = time.perf_counter()
    sampler.stop = True
    for _ in it:  # the batches already dealt out: let the workers finish them and exit cleanly
        pass
    del it, loader
    return dict(value=n / (t1 - t0), clips=n, seconds=t1 - t0, workers=workers)


def cpu_baseline_block(workload, seconds, full=True):
    r = cpu_reference_throughput(workload, seconds)
    block = {
        "value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "cpu_model": cpu_model(),
        "sample": (f"{r['clips']} clips in {r['seconds']:.1f}s, oracle.pipeline.make_transform (restated "
                   f"reference transform), {r['cores']} single-thread processes of {r['host_cores']} host cores")}
    if full:
        one = cpu_reference_throughput(workload, max(4.0, seconds / 3), warm_seconds=1.0, procs=1)
        block["one_thread"] = {"value": one["value"], "unit": UNIT, "cores": 1,
                               "sample": f"{one['clips']} clips in {one['seconds']:.1f}s, 1 process x 1 thread"}
        try:
            dl = cpu_dataloader_throughput(workload, max(6.0, seconds / 2))
            block["dataloader"] = {
                "value": dl["value"], "unit": UNIT, "cores": dl["workers"],
                "sample": (f"{dl['clips']} clips in {dl['seconds']:.1f}s, torch DataLoader(num_workers={dl['workers']}, "
                           f"batch_size={WORKLOADS[workload]['batch']}, default_collate) over an in-memory dataset: "
                           "the reference's execution model, run.py:170-176")}
        except Exception as e:  # noqa: BLE001 -- the baseline must not take the bench line down
            block["dataloader"] = {"error": f"{type(e).__name__}: {e}"}
    return block


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # other ranks exit 0 without work
    wl = WORKLOADS[args.workload]
    # a "step" is one batch of wl['batch'] clips; the run is duration-bounded so that it ends
    # within minutes whatever K is
    est = 10.0  # clips/s order of magnitude per core-second is ~10
    want_clips = (args.steps + args.warmup) * wl["batch"]
    seconds = max(5.0, min(40.0, want_clips / est))
    r = cpu_reference_throughput(args.workload, seconds)
    ms_per_step = 1000.0 * wl["batch"] / r["value"]
    sample = (f"{r['clips']} clips of {wl['shape']} uint8 through oracle.pipeline.make_transform"
              f"({wl['kwargs']}) in {r['seconds']:.1f}s on {r['cores']} single-thread processes; "
              "duration-bounded sample, step = one batch equivalent")
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, wl, max(1, args.gpus)),
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": sample,
                         "cpu_model": cpu_model()},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------
# clocks: sampled by a SEPARATE process, so that the sampler never competes with the launch loop
# for the interpreter lock (an in-process NVML thread cost ~1.6 us per step of a 20-step region)
# --------------------------------------------------------------------------------------
_SAMPLER_CODE = r"""
import json, select, sys, time
idx, period = int(sys.argv[1]), float(sys.argv[2])
R = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
     0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}
out = {"ok": False, "samples": [], "max_mhz": None}
try:
    import pynvml as nv
    nv.nvmlInit()
    h = nv.nvmlDeviceGetHandleByIndex(idx)
    out["max_mhz"] = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
    out["ok"] = True
except Exception as e:
    out["error"] = repr(e)
print("ready", flush=True)
while True:
    if out["ok"]:
        try:
            t = time.time()
            mhz = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            out["samples"].append([t, mhz, [n for b, n in R.items() if mask & b]])
        except Exception:
            pass
    if select.select([sys.stdin], [], [], period)[0]:
        break
print(json.dumps(out), flush=True)
"""


class ClockSampler:
    """SM clock and throttle reasons of one GPU, sampled through NVML by a child process while the
    timed region runs.  ``mark()`` brackets the region on the shared wall clock; samples inside it
    (or, for a region shorter than the period, the two samples around it) are reported."""

    def __init__(self, index, period=0.002, enabled=True):
        self.proc = None
        if not enabled:
            self.t0 = self.t1 = None
            return
        try:
            self.proc = subprocess.Popen([sys.executable, "-c", _SAMPLER_CODE, str(index), str(period)],
                                         stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True)
            self.proc.stdout.readline()  # "ready": NVML is up and the loop is running
        except Exception:  # noqa: BLE001
            self.proc = None
        self.t0 = self.t1 = None

    def mark_start(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def finish(self):
        res = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return res
        try:
            out, _ = self.proc.communicate("stop\n", timeout=10)
            data = json.loads(out.strip().splitlines()[-1])
        except Exception:  # noqa: BLE001
            self.proc.kill()
            return res
        res["sm_max_mhz"] = data.get("max_mhz")
        samples = data.get("samples", [])
        inside = [s for s in samples if self.t0 is not None and self.t0 <= s[0] <= self.t1]
        if not inside and samples and self.t0 is not None:
            before = [s for s in samples if s[0] < self.t0][-1:]
            after = [s for s in samples if s[0] > self.t1][:1]
            inside = before + after
        if inside:
            res["sm_mhz"] = statistics.median(s[1] for s in inside)
            res["reasons"] = sorted({r for s in inside for r in s[2]})
            res["samples"] = len(inside)
        res["sampler"] = "NVML in a child process, 2 ms period; samples inside (or bracketing) the timed region"
        return res


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except (ValueError, IndexError):
            return local_rank
    return local_rank


# --------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------
def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        v = float(json.load(open(path))["hbm_gbs"])
        return v, "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def recorded_traffic(workload):
    """Steady-state DRAM bytes per launch of the dominant kernel, measured with ncu over consecutive
    launches on rotating buffers (profiles/traffic.json says how); None when not measured."""
    try:
        v = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(workload)
        return v if isinstance(v, (int, float)) else None
    except Exception:  # noqa: BLE001
        return None


def pin_to_rank_cores(local_rank, world):
    """Give each rank its own contiguous share of the host cores: the launch loop of one rank must
    not be descheduled in favour of another rank's helper threads inside a sub-millisecond timed
    region (N = 8 measured 0.94 of the 1-GPU rate without this)."""
    if world <= 1:
        return None
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = len(cores) // world
        if per < 1:
            return None
        mine = cores[local_rank * per:(local_rank + 1) * per]
        os.sched_setaffinity(0, mine)
        return mine
    except (AttributeError, OSError):
        return None


def run_b200_arm(args):
    import torch

    import pva_b200
    from pva_b200 import _lib

    rank, world, local_rank = pva_b200.rank_and_world()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this path has no CPU fallback)")
    pinned_cores = pin_to_rank_cores(local_rank, world)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        # stdout carries exactly one JSON line, but NCCL prints there while the communicator comes up
        # (its version banner under the image's NCCL_DEBUG=VERSION): point fd 1 at stderr meanwhile
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize(dev)
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    wl = dict(WORKLOADS[args.workload])
    if args.batch:
        wl["batch"] = args.batch  # experiments only; the named workloads define the reported metric
    if args.size:
        wl["kwargs"] = dict(wl["kwargs"], min_short_side_scale=args.size, max_short_side_scale=args.size)
    B, shape, kw = wl["batch"], wl["shape"], wl["kwargs"]
    t = pva_b200.make_transform(**kw, device=dev, out_dtype=torch.float16, channels_last=not args.planar,
                                norm_mode=args.norm_mode, kernel=args.kernel)
    esize = 2
    C, crop, n_fast, n_slow = shape[0], t.crop_size, t.num_frames, t.n_slow

    # ---- synthetic, device-resident, rotated so the working set exceeds L2 ----
    n_sets, in_bytes, out_bytes = rotation_sets(wl)
    gen = torch.Generator().manual_seed(pva_b200.rank_seed(1234, rank))
    torch.manual_seed(pva_b200.rank_seed(42, rank))
    sets, algo_bytes = [], []
    for s in range(n_sets):
        clips = torch.randint(0, 256, (B,) + tuple(shape), dtype=torch.uint8, generator=gen).to(dev)
        if args.float_clips:
            clips = clips.float() + 0.25  # float32 samples that are not whole bytes: nothing can be narrowed to uint8
        params = [t.sample_params(shape[2], shape[3]) for _ in range(B)]
        # resident and complete before the timed region; every set has its own outputs, so consecutive
        # launches never share an output buffer (what a prefetching input pipeline's ring looks like)
        sets.append(t.prepare_batch(clips, params, out=t.alloc_outputs(B, dev), clips_ready=True,
                                    outputs_disjoint=not args.serialize_stores, max_ctas=args.max_ctas))
        b = 0
        for p in params:
            fh, fw = _lib.source_footprint(shape[2], shape[3], p.oh, p.ow, p.top, p.left, crop)
            b += C * n_fast * fh * fw * (4 if args.float_clips else 1) + C * (n_fast + n_slow) * crop * crop * esize
        algo_bytes.append(b)
    stream = torch.cuda.current_stream(dev)
    raw_stream = stream.cuda_stream
    launchers = [s.launch_raw for s in sets]

    def step(i):
        launchers[i % n_sets](raw_stream)

    graph = None
    if args.graph:
        # one CUDA graph replays a full rotation over the buffer sets: n_sets launches per replay
        for i in range(n_sets):
            step(i)
        torch.cuda.synchronize(dev)
        cap = torch.cuda.Stream(dev)
        with torch.cuda.stream(cap):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=cap):
                for i in range(n_sets):
                    sets[i].launch_raw(cap.cuda_stream)
        torch.cuda.synchronize(dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    planned = {}

    def plan_steps(first, count):
        # the order of the plain launches is fixed (and the list built) before the clock starts
        key = (first % n_sets, count)
        if key not in planned:
            planned[key] = [launchers[(first + j) % n_sets] for j in range(count)]
        return planned[key]

    def run_steps(first, count):
        # exactly `count` steps: whole rotations come from the graph, the remainder are plain launches
        i = 0
        if graph is not None:
            while first % n_sets and i < count:  # align to a rotation boundary
                step(first)
                first, i = first + 1, i + 1
            while count - i >= n_sets:
                graph.replay()
                first, i = first + n_sets, i + n_sets
        for fn in plan_steps(first, count - i):
            fn(raw_stream)

    sampler = ClockSampler(physical_gpu_index(local_rank), period=args.clock_period, enabled=not args.no_clocks)
    # Warm-up: the W steps asked for, but never fewer than three rotations over the buffer sets, and then
    # one untimed rehearsal of the timed region itself (same barrier, same event pair, same K launches) when
    # K is small -- a 20-step region lasts 0.7 ms, and a host that runs it for the first time (cold
    # instruction and data caches on the launch path) adds ~20 us to it.
    warm_steps = max(args.warmup, 3 * n_sets)
    run_steps(0, warm_steps)
    rehearsed = 0
    if args.steps <= 256 and not args.no_rehearsal:
        for _ in range(2):
            r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            r0.record(stream)
            run_steps(args.warmup, args.steps)
            r1.record(stream)
            barrier()
            rehearsed += args.steps
    info = _lib.last_launch_info()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if graph is None:
        plan_steps(args.warmup, args.steps)
    barrier()
    sampler.mark_start()
    e0.record(stream)
    run_steps(args.warmup, args.steps)
    e1.record(stream)
    barrier()
    sampler.mark_end()
    clocks = sampler.finish()
    own_ms = e0.elapsed_time(e1)
    elapsed_ms = own_ms
    rank_stats = None
    if dist is not None:
        mine = torch.tensor([own_ms], device=dev, dtype=torch.float64)
        every = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        per_rank = [float(v.item()) / args.steps for v in every]
        elapsed_ms = max(per_rank) * args.steps
        rank_stats = {"min": min(per_rank), "median": statistics.median(per_rank), "max": max(per_rank),
                      "slowest_rank": per_rank.index(max(per_rank))}
    ms_per_step = elapsed_ms / args.steps
    value = world * B * args.steps / (elapsed_ms / 1000.0)
    used = [(args.warmup + i) % n_sets for i in range(args.steps)]
    mean_algo = sum(algo_bytes[u] for u in used) / len(used)

    # ---- host cost of one call: K calls into an empty queue, wall clock (nothing blocks: the
    # launch queue is far deeper than K)
    barrier()
    k_host = min(max(args.steps, 10), 100)
    todo = [launchers[j % n_sets] for j in range(k_host)]
    h0 = time.perf_counter()
    for fn in todo:
        fn(raw_stream)
    host_us = (time.perf_counter() - h0) * 1e6 / k_host
    torch.cuda.synchronize(dev)

    # ---- one launch in isolation: queued behind a filler (a plain device memset long enough to
    # cover the host's enqueue time), so both events and the launch are already in the queue when
    # the filler ends and no host time sits between the events
    filler = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    iso = []
    for j in range(50):
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        filler.zero_()
        a0.record(stream)
        launchers[j % n_sets](raw_stream)
        a1.record(stream)
        torch.cuda.synchronize(dev)
        iso.append(a0.elapsed_time(a1) * 1000.0)
    del filler
    launch_us_isolated = statistics.median(iso)

    peak, peak_src = measured_peak()
    achieved = mean_algo / (ms_per_step / 1000.0) / 1e9
    traffic = recorded_traffic(args.workload) if not (args.batch or args.size or args.planar or args.float_clips) else None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src, "frac_of_nominal_8TBs": achieved / NOMINAL_HBM_GBS,
                "algorithmic_bytes_per_launch": mean_algo,
                "kernel": {1: "pva_generic_kernel", 2: "pva_tiled_kernel"}.get(info["kernel"], "?"),
                "time_basis": "steady-state time per launch (K back-to-back launches / K; consecutive launches overlap)",
                "isolated_launch_GBps": mean_algo / (launch_us_isolated * 1e-6) / 1e9}

    # ---- informational: the same K steps issued round-robin on two streams.  Consecutive batches are
    # independent, so the block scheduler starts the next launch's CTAs on SMs the previous launch has
    # already left: the pipeline fill of one launch overlaps the drain of the other.
    overlapped = None
    if args.overlap_streams > 1:
        side = [torch.cuda.Stream(dev) for _ in range(args.overlap_streams)]
        fork, joins = torch.cuda.Event(), [torch.cuda.Event() for _ in side]

        def run_overlapped(first, count):
            fork.record(stream)
            for st in side:
                st.wait_event(fork)
            for i in range(count):
                sets[(first + i) % n_sets].launch_raw(side[i % len(side)].cuda_stream)
            for st, ev in zip(side, joins):
                ev.record(st)
                stream.wait_event(ev)

        run_overlapped(0, args.warmup)
        barrier()
        o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        o0.record(stream)
        run_overlapped(args.warmup, args.steps)
        o1.record(stream)
        barrier()
        o_ms = o0.elapsed_time(o1)
        if dist is not None:
            tt = torch.tensor([o_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            o_ms = float(tt.item())
        overlapped = {"streams": args.overlap_streams, "value": world * B * args.steps / (o_ms / 1000.0), "unit": UNIT,
                      "ms_per_step": o_ms / args.steps,
                      "algorithmic_GBps": mean_algo / (o_ms / args.steps / 1000.0) / 1e9}

    # ---- e2e: HOST buffers through the public API, copies inside the timed region ----------
    e2e = run_e2e(torch, pva_b200, t, wl, dev, dist, world, rank, args)

    config = workload_config(args.workload, wl, world)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": config,
        "impl_detail": {"out": "fp16 CTHW (the reference's memory order)" if args.planar else "fp16 THWC (channels_last_3d)",
                        "norm_mode": args.norm_mode, "launch": info,
                        "untimed_steps_before_the_timed_region": warm_steps + rehearsed,
                        "flags": "PVA_FLAG_CLIPS_READY" + ("" if args.serialize_stores else " | PVA_FLAG_OUTPUTS_DISJOINT"), "cuda_graph": bool(args.graph),
                        "pinned_cores": pinned_cores},
        "roofline": roofline, "launch_us_isolated": launch_us_isolated, "host_us_per_call": host_us,
        "e2e": e2e, "gpu_launches": args.steps * info["launches"], "clocks": clocks,
    }
    if rank_stats is not None:
        line["rank_ms_per_step"] = rank_stats
    if overlapped is not None:
        line["overlapped_launches"] = overlapped
    if world == 1 and rank == 0 and not args.no_cpu_baseline:
        del sets, launchers
        torch.cuda.empty_cache()
        line["cpu_baseline"] = cpu_baseline_block(args.workload, args.cpu_seconds, full=not args.cpu_baseline_short)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


AUTO_E2E_UPLOAD = "inplace"


def upload_desc(h2d_bytes, B, shape, t):
    whole = B * shape[0] * shape[1] * shape[2] * shape[3]
    if h2d_bytes == whole:
        return "H2D copy of whole clips"
    if h2d_bytes == whole // shape[1] * t.num_frames:
        return "H2D of the selected frames (strided DMA, pva_upload_selected_frames)"
    return "no H2D copy: the kernel reads the selected frames' rows from the pinned batch over PCIe (h2d bytes = rows touched x full width)"


def copy_only_ceiling(torch, dev, dist, world, B, steps, h2d_bytes, d2h_bytes):
    """What the host link allows for this step's bytes with NO preprocessing: every step moves
    ``h2d_bytes`` from pinned host memory to the device and ``d2h_bytes`` back into pinned host memory
    with plain asynchronous copies, three slots, both directions in flight at once (PCIe is full
    duplex), all ranks concurrently.  The end-to-end figure cannot exceed this on the same box."""
    depth = 3
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    h_in = [torch.empty(max(h2d_bytes, 1), dtype=torch.uint8, pin_memory=True) for _ in range(depth)]
    d_in = [torch.empty(max(h2d_bytes, 1), dtype=torch.uint8, device=dev) for _ in range(depth)]
    d_out = [torch.empty(max(d2h_bytes, 1), dtype=torch.uint8, device=dev) for _ in range(depth)]
    h_out = [torch.empty(max(d2h_bytes, 1), dtype=torch.uint8, pin_memory=True) for _ in range(depth)]

    def run(n):
        for i in range(n):
            k = i % depth
            with torch.cuda.stream(s_in):
                d_in[k].copy_(h_in[k], non_blocking=True)
            with torch.cuda.stream(s_out):
                h_out[k].copy_(d_out[k], non_blocking=True)
        s_in.synchronize()
        s_out.synchronize()

    run(3)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    run(steps)
    ms = (time.perf_counter() - t0) * 1000.0
    if dist is not None:
        tt = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    return world * B * steps / (ms / 1000.0), ms / steps


def run_e2e(torch, pva_b200, t, wl, dev, dist, world, rank, args):
    """Public-API path with HOST buffers.  Every step: worker-side draws (SampleParams) and
    collate of raw uint8 clips into pinned memory were done up front (they are CPU-worker work
    in the reference too); inside the timed region each step gets that step's pinned clips to
    the device, runs the fused kernel, and copies [slow, fast] back into pinned host memory.
    StagedPreprocess overlaps the phases of consecutive steps (PCIe is full duplex)."""
    B, shape = wl["batch"], wl["shape"]
    steps = max(5, min(args.steps, args.e2e_steps))
    gen = torch.Generator().manual_seed(99 + rank)
    sp = pva_b200.SampleParams(t)
    host_batches = []
    for _ in range(3):
        samples = [sp({"video": torch.randint(0, 256, shape, dtype=torch.uint8, generator=gen), "label": 0})
                   for _ in range(B)]
        b = pva_b200.collate_raw_clips(samples)
        b["video"] = b["video"].pin_memory()
        b["params"] = [pva_b200.ClipParams.from_row(r) for r in b["pva_params"].tolist()]
        host_batches.append(b)

    def measure(host_out):
        # How the clips get to the device (--e2e-upload):
        #   all       one flat copy of the whole pinned batch (134 MB)
        #   selected  strided DMA of the frames the transform selects (67 MB)
        #   inplace   no copy: the kernel reads the selected frames' rows from the pinned batch over
        #             PCIe (~59 MB), on a few SMs
        mode = args.e2e_upload
        if mode == "auto":
            mode = AUTO_E2E_UPLOAD
        if args.upload_all_frames:
            mode = "all"
        staged = pva_b200.StagedPreprocess(t, B, shape, dev, depth=3, host_out=host_out,
                                           upload_selected_frames=(mode == "selected"),
                                           read_in_place=(mode == "inplace"), max_ctas=args.e2e_ctas)

        def run(n):
            tickets = []
            for i in range(n):
                hb = host_batches[i % 3]
                tickets.append(staged.submit(hb["video"], hb["params"]))
                if len(tickets) > 2:
                    staged.result(tickets[-3])  # consume results in order, two batches in flight
            for tk in tickets[-2:]:
                staged.result(tk)

        run(4)
        staged.drain()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        run(steps)
        staged.drain()
        torch.cuda.synchronize(dev)
        ms = (time.perf_counter() - t0) * 1000.0  # wall clock around fully drained work: includes every copy
        if dist is not None:
            tt = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        return ms, staged.h2d_bytes, staged.d2h_bytes

    ms, h2d, d2h = measure(host_out=True)
    ms_dev, h2d_dev, _ = measure(host_out=False)
    value = world * B * steps / (ms / 1000.0)
    ceil_value, ceil_ms = copy_only_ceiling(torch, dev, dist, world, B, steps, int(h2d), int(d2h))
    return {"value": value, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "steps": steps, "ms_per_step": ms / steps,
            "path": "SampleParams -> collate_raw_clips (pinned) -> StagedPreprocess: " + upload_desc(h2d, B, shape, t)
                    + " | fused kernel | D2H of [slow, fast] to pinned memory, 3 slots",
            # the same bytes per step and direction as plain pinned copies, no preprocessing: the
            # host link's own limit on this box at this N
            "ceiling": {"value": ceil_value, "unit": UNIT, "ms_per_step": ceil_ms,
                        "what": "copy-only: h2d_bytes_per_step H2D + d2h_bytes_per_step D2H per step, pinned, "
                                "3 slots, both directions concurrent, all ranks at once"},
            "frac_of_ceiling": value / ceil_value if ceil_value else None,
            # informational: the same pipeline when the result stays on the device for the model
            # (what a training step consumes), i.e. without the D2H leg
            "device_resident_result": {"value": world * B * steps / (ms_dev / 1000.0), "unit": UNIT,
                                       "ms_per_step": ms_dev / steps, "h2d_bytes_per_step": int(h2d_dev),
                                       "d2h_bytes_per_step": 0,
                                       "path": upload_desc(h2d_dev, B, shape, t) + " | fused kernel"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="slowfast_train_b8")
    ap.add_argument("--norm-mode", dest="norm_mode", choices=["affine", "exact"], default="affine")
    ap.add_argument("--kernel", choices=["auto", "generic", "tiled"], default="auto")
    ap.add_argument("--e2e-steps", dest="e2e_steps", type=int, default=40)
    ap.add_argument("--cpu-seconds", dest="cpu_seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-baseline-short", action="store_true",
                    help="only the one-process-per-core CPU figure (skip the one-thread and DataLoader modes)")
    ap.add_argument("--overlap-streams", dest="overlap_streams", type=int, default=2,
                    help="also time the K steps issued round-robin on this many streams (informational; 0/1 = skip)")
    ap.add_argument("--e2e-ctas", dest="e2e_ctas", type=int, default=4,
                    help="grid cap of launches that read pinned host clips in place (PvaPlan.max_ctas)")
    ap.add_argument("--e2e-upload", dest="e2e_upload", choices=["auto", "all", "selected", "inplace"], default="auto",
                    help="e2e H2D leg: whole clips, only the selected frames, or auto (see run_e2e)")
    ap.add_argument("--upload-all-frames", action="store_true",
                    help="e2e: copy whole clips to the device instead of only the frames the transform selects")
    ap.add_argument("--max-ctas", dest="max_ctas", type=int, default=0,
                    help="PvaPlan.max_ctas of the device-resident launches (0 = one CTA per SM)")
    ap.add_argument("--no-rehearsal", action="store_true",
                    help="skip the untimed rehearsal of the timed region that follows the warm-up steps")
    ap.add_argument("--no-clocks", action="store_true", help="do not sample clocks (experiments)")
    ap.add_argument("--clock-period", dest="clock_period", type=float, default=0.002)
    ap.add_argument("--serialize-stores", action="store_true",
                    help="do not set PVA_FLAG_OUTPUTS_DISJOINT: every launch waits for the previous one before its first store")
    ap.add_argument("--float-clips", dest="float_clips", action="store_true",
                    help="device-resident region on float32 clips (experiments; the e2e leg still uses uint8)")
    ap.add_argument("--planar", action="store_true",
                    help="write the reference's contiguous C x T x S x S order instead of channels-last (experiments)")
    ap.add_argument("--batch", type=int, default=0, help="override the workload's batch (experiments)")
    ap.add_argument("--size", type=int, default=0, help="fix the short-side scale (bandwidth sweep)")
    ap.add_argument("--graph", dest="graph", action="store_true", default=False,
                    help="replay the steps from a captured CUDA graph (4 launches per replay)")
    ap.add_argument("--no-graph", dest="graph", action="store_false",
                    help="plain launches through the C ABI (default: every launch is chained to the previous "
                         "one by programmatic dependent launch; a graph replay boundary is a full barrier)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
