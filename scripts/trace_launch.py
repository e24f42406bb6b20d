#!/usr/bin/env python
"""Per-CTA time stamps of consecutive tiled-kernel launches (experiments build only).

    make -C pytorchvideo-accelerate_b200/csrc experiments
    PVA_LIBRARY=scripts/libpva_preproc_exp.so python scripts/trace_launch.py [--launches 12]

The experiments library stamps %globaltimer at: 0 kernel entry, 1 set-up done, 2 first tile landed,
3 first strip computed, 4 PDL wait returned (previous kernel complete), 5 last tile handed to the store
engine, 6 CTA exit, 7 producer issued its last load.  The script prints, per launch, the median / max
over CTAs of each stamp relative to the launch's first CTA entry, and how launches overlap.
"""
import argparse
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402
import pva_b200  # noqa: E402

KMAXGRID = 160


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="slowfast_train_b8")
    ap.add_argument("--launches", type=int, default=12)
    ap.add_argument("--batch", type=int, default=0)
    args = ap.parse_args()
    assert "exp" in pva_b200.library_path(), "point PVA_LIBRARY at scripts/libpva_preproc_exp.so"
    dev = torch.device("cuda", 0)
    wl = dict(bench.WORKLOADS[args.workload])
    if args.batch:
        wl["batch"] = args.batch
    B, shape, kw = wl["batch"], wl["shape"], wl["kwargs"]
    t = pva_b200.make_transform(**kw, device=dev)
    gen = torch.Generator().manual_seed(1234)
    torch.manual_seed(42)
    sets = []
    for _ in range(4):
        clips = torch.randint(0, 256, (B,) + tuple(shape), dtype=torch.uint8, generator=gen).to(dev)
        params = [t.sample_params(shape[2], shape[3]) for _ in range(B)]
        sets.append(t.prepare_batch(clips, params, out=t.alloc_outputs(B, dev), clips_ready=True,
                                    outputs_disjoint=not os.environ.get("PVA_SERIALIZE_STORES")))
    raw = torch.cuda.current_stream(dev).cuda_stream
    for i in range(8):
        sets[i % 4].launch_raw(raw)
    torch.cuda.synchronize()
    n = args.launches
    trace = torch.zeros((n, KMAXGRID, 128), dtype=torch.int64, device=dev)
    os.environ["PVA_TILED_TRACE"] = hex(trace.data_ptr())
    os.environ["PVA_TILED_TRACE_SLOTS"] = str(n)
    torch.cuda.synchronize()
    for i in range(n):
        sets[i % 4].launch_raw(raw)
    torch.cuda.synchronize()
    del os.environ["PVA_TILED_TRACE"]
    tr = trace.cpu()
    grid = int((tr[0, :, 0] != 0).sum())
    # the static launch counter started before this script's traced launches: find the rotation
    order = sorted(range(n), key=lambda j: int(tr[j, :grid, 0].min()))
    base = int(tr[order[0], :grid, 0].min())
    names = ["entry", "setup", "tile0", "strip0", "pdlwait", "laststore", "exit", "loads_done"]
    print(f"grid {grid}; times in us relative to the first traced launch's first CTA entry")
    prev_exit = None
    for j in order:
        r = tr[j, :grid].double()
        e0 = float(r[:, 0].min())
        line = [f"launch@{(e0 - base) / 1e3:8.2f}"]
        for k, nm in enumerate(names):
            col = (r[:, k] - e0) / 1e3
            line.append(f"{nm} {statistics.median(col.tolist()):6.2f}/{float(col.max()):6.2f}")
        if prev_exit is not None:
            line.append(f"| entry-prev_last_exit {(e0 - prev_exit) / 1e3:6.2f}")
        prev_exit = float(r[:, 6].max())
        print("  ".join(line))
    # per-tile phases of each group's first thread: medians over CTAs and the group's tiles 4..15, LAST launch
    r = tr[order[-1], :grid, 8:128].reshape(grid, 2, 12, 5).double()
    names2 = ["drain wait (top -> slot free)", "tile wait (slot free -> landed)", "compute (landed -> strip done)",
              "hand-over (strip done -> store issued)", "loop (store issued -> next top)"]
    for g in range(2):
        rg = r[:, g]
        ok = (rg[:, 0:11, :] > 0).all(dim=2) & (rg[:, 1:12, 0] > 0)
        print(f"group {g}: per group-tile phases, us (median / 90th percentile), last launch:")
        for j in range(4):
            d = (rg[:, 0:11, j + 1] - rg[:, 0:11, j])[ok] / 1e3
            if d.numel():
                print(f"  {names2[j]:42s} {float(d.median()):6.2f} / {float(d.quantile(0.9)):6.2f}   mean {float(d.mean()):5.2f}")
        d = (rg[:, 1:12, 0] - rg[:, 0:11, 4])[ok] / 1e3
        print(f"  {names2[4]:42s} {float(d.median()):6.2f} / {float(d.quantile(0.9)):6.2f}   mean {float(d.mean()):5.2f}")
        per = (rg[:, 1:12, 0] - rg[:, 0:11, 0])[ok] / 1e3
        print(f"  group-tile period                          {float(per.median()):6.2f} / {float(per.quantile(0.9)):6.2f}   mean {float(per.mean()):5.2f}")
    # phase of group 1 relative to group 0: (start of group 1's compute - start of group 0's compute) / period
    both = (r[:, 0, 0:11, 2] > 0) & (r[:, 1, 0:11, 2] > 0) & (r[:, 0, 1:12, 2] > 0)
    off = ((r[:, 1, 0:11, 2] - r[:, 0, 0:11, 2]) / (r[:, 0, 1:12, 2] - r[:, 0, 0:11, 2]))[both]
    if off.numel():
        hist = torch.histc(off.clamp(0, 1).float(), bins=10, min=0, max=1)
        print("offset of group 1's compute start within group 0's period (deciles):", [int(v) for v in hist.tolist()])
    torch.save(tr, os.path.join(ROOT, "gpurun_out", "trace.pt"))


if __name__ == "__main__":
    main()
