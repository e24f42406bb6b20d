#!/usr/bin/env python
"""Where the time of a SHORT timed region goes (the driver's protocol is --steps 20 --warmup 5).

Times K back-to-back launches (same protocol as bench.py: sync, event, K plain launches through the
C ABI, event, sync) for several K, many trials each, and fits  T(K) = fixed + K * steady.  Also
reports the host cost of one pva_preprocess call (wall clock around K calls into an empty queue).

    python scripts/steps_fit.py [--workload NAME] [--trials 30]
"""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402
import pva_b200  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="slowfast_train_b8")
    ap.add_argument("--trials", type=int, default=30)
    ap.add_argument("--ks", default="1,2,5,10,20,40,80,200,1000")
    ap.add_argument("--tag", default="")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    wl = bench.WORKLOADS[args.workload]
    B, shape, kw = wl["batch"], wl["shape"], wl["kwargs"]
    t = pva_b200.make_transform(**kw, device=dev)
    n_sets = 4
    gen = torch.Generator().manual_seed(1234)
    torch.manual_seed(42)
    sets = []
    for _ in range(n_sets):
        clips = torch.randint(0, 256, (B,) + tuple(shape), dtype=torch.uint8, generator=gen).to(dev)
        params = [t.sample_params(shape[2], shape[3]) for _ in range(B)]
        sets.append(t.prepare_batch(clips, params, out=t.alloc_outputs(B, dev), clips_ready=True,
                                    outputs_disjoint=not os.environ.get("PVA_SERIALIZE_STORES")))
    stream = torch.cuda.current_stream(dev)
    raw = stream.cuda_stream

    def timed(k, first=0):
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h0 = time.perf_counter()
        e0.record(stream)
        for i in range(k):
            sets[(first + i) % n_sets].launch_raw(raw)
        e1.record(stream)
        h1 = time.perf_counter()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) * 1000.0, (h1 - h0) * 1e6  # us, us

    for _ in range(5):
        timed(20)
    out = {"workload": args.workload, "rows": []}
    for k in [int(v) for v in args.ks.split(",")]:
        dev_us, host_us = [], []
        for tr in range(args.trials if k <= 200 else 5):
            d, h = timed(k, first=tr)
            dev_us.append(d)
            host_us.append(h)
        row = {"K": k, "dev_us_median": statistics.median(dev_us), "dev_us_min": min(dev_us), "dev_us_max": max(dev_us),
               "per_step_us_median": statistics.median(dev_us) / k, "host_issue_us_per_call": statistics.median(host_us) / k}
        out["rows"].append(row)
        print(json.dumps(row), flush=True)
    pts = [(r["K"], r["dev_us_median"]) for r in out["rows"] if 5 <= r["K"] <= 200]
    n = len(pts)
    if n >= 2:
        sx = sum(p[0] for p in pts); sy = sum(p[1] for p in pts)
        sxx = sum(p[0] ** 2 for p in pts); sxy = sum(p[0] * p[1] for p in pts)
        steady = (n * sxy - sx * sy) / (n * sxx - sx * sx)
        fixed = (sy - steady * sx) / n
        out["fit"] = {"fixed_us": fixed, "steady_us_per_launch": steady}
        print(json.dumps(out["fit"]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"steps_fit_{args.workload}{args.tag}.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
