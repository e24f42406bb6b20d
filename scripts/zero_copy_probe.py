#!/usr/bin/env python
"""Can the fused kernel read its clips straight from pinned host memory (no H2D copy)?

The bulk copies of the tiled kernel take any global-space address; under unified addressing a
pinned host allocation is one.  This probe points the clip descriptors of a prepared batch at the
pinned host copy of the same clips, checks the result bit for bit against the device-resident run,
and times both: the H2D copy + kernel, and the kernel reading over PCIe (only the selected frames'
band rows ever cross the bus).  Experiment only.
"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pva_b200  # noqa: E402
from pva_b200 import _lib  # noqa: E402

dev = torch.device("cuda", 0)
B, shape = 8, (3, 64, 256, 340)
t = pva_b200.make_transform(num_frames=32, slowfast_alpha=4, training=True, device=dev)
torch.manual_seed(1)
host = torch.randint(0, 256, (B,) + shape, dtype=torch.uint8).pin_memory()
params = [t.sample_params(shape[2], shape[3]) for _ in range(B)]
devclips = host.to(dev)
pb = t.prepare_batch(devclips, params, clips_ready=True)
ref = [o.clone() for o in pb.launch()]
torch.cuda.synchronize()

clip_bytes = host[0].numel()
for i in range(B):
    pb.descs[i].src = host.data_ptr() + i * clip_bytes  # the pinned host copy
for colcrop, grid_env in (("0", None), ("0", "16"), ("0", "8"), ("0", "4"), ("1", None), ("1", "16"), ("1", "8")):
    os.environ["PVA_TILED_COLCROP"] = colcrop
    if grid_env is None:
        os.environ.pop("PVA_TILED_MAXGRID", None)
    else:
        os.environ["PVA_TILED_MAXGRID"] = grid_env
    got = pb.launch()
    torch.cuda.synchronize()
    same = all(torch.equal(a, b) for a, b in zip(ref, got))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record()
    for _ in range(n):
        pb.launch()
    e1.record()
    torch.cuda.synchronize()
    print(f"zero-copy source, column crop {colcrop}, grid cap {grid_env}: identical={same}  "
          f"{e0.elapsed_time(e1) / n:.3f} ms per batch (grid {_lib.last_launch_info()['grid']})", flush=True)
os.environ.pop("PVA_TILED_MAXGRID", None)
os.environ["PVA_TILED_COLCROP"] = "0"

# the copy + kernel it would replace
for i in range(B):
    pb.descs[i].src = devclips.data_ptr() + i * clip_bytes
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    devclips.copy_(host, non_blocking=True)
    pb.launch()
e1.record()
torch.cuda.synchronize()
print(f"H2D copy of whole clips + kernel: {e0.elapsed_time(e1) / 10:.3f} ms per batch")

# ---- outputs written in place as well: [slow, fast] in pinned host memory, one kernel does everything
os.environ.pop("PVA_TILED_MAXGRID", None)
host_out = [torch.empty(r.shape, dtype=r.dtype).pin_memory() for r in pb.raw_out]
for ctas in (8, 16, 32, 148):
    pbh = t.prepare_batch(host, params, out=host_out, clips_ready=True, max_ctas=ctas, device=dev)
    pbh.launch()
    torch.cuda.synchronize()
    same = all(torch.equal(a.cpu(), t._logical(b)) for a, b in zip(ref, host_out))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        pbh.launch()
    e1.record()
    torch.cuda.synchronize()
    print(f"host clips in, host result out, one kernel, {ctas} CTAs: identical={same}  {e0.elapsed_time(e1) / 10:.3f} ms per batch")
