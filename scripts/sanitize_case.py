"""Small invocation of both kernels for compute-sanitizer (memcheck): exact-size tensors from
cudaMalloc-like torch allocations, ragged clips, both load paths."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import pva_b200  # noqa: E402
from pva_b200.geometry import ClipParams  # noqa: E402

torch.cuda.set_device(0)
for kernel in ("generic", "tiled"):
    for training, shape in ((True, (3, 9, 130, 172)), (False, (3, 8, 256, 340)), (True, (3, 5, 144, 131))):
        crop = 128
        t = pva_b200.make_transform(min_short_side_scale=128, max_short_side_scale=160, crop_size=crop, num_frames=4,
                                    slowfast_alpha=2, training=training, kernel=kernel)
        g = torch.Generator().manual_seed(1)
        clips = [torch.randint(0, 256, shape, dtype=torch.uint8, generator=g).cuda() for _ in range(3)]
        torch.manual_seed(2)
        params = [t.sample_params(shape[2], shape[3]) for _ in clips]
        # one window pinned to the bottom-right corner: the last bytes of the allocation are read
        oh, ow = params[0].oh, params[0].ow
        params[0] = ClipParams(params[0].size, oh, ow, oh - crop, ow - crop, False)
        slow, fast = t.run_batch(clips, params)
        torch.cuda.synchronize()
        print(kernel, training, shape, float(fast.float().abs().mean()))
print("done")
