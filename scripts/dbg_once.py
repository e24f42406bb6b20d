import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, pva_b200
from pva_b200 import _lib
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
sync_each = int(sys.argv[3]) if len(sys.argv) > 3 else 1
t = pva_b200.make_transform(num_frames=32, slowfast_alpha=4, training=True)
g = torch.Generator().manual_seed(1234)
torch.manual_seed(42)
sets = []
for s in range(4):
    clips = torch.randint(0, 256, (B, 3, 64, 256, 340), dtype=torch.uint8, generator=g).cuda()
    params = [t.sample_params(256, 340) for _ in range(B)]
    sets.append(t.prepare_batch(clips, params))
for i in range(reps):
    sets[i % 4].launch()
    if sync_each:
        torch.cuda.synchronize()
        print("launch", i, "set", i % 4, "ok", _lib.last_launch_info()["smem_bytes"], flush=True)
torch.cuda.synchronize()
print("all ok")
