"""Calibration of what HBM sustains for this path's traffic mix (torch kernels only; not part of the product).
Pure write (fill_), pure read (sum), copy (1:1) and a 1:2 read:write mix, on buffers rotated past L2."""
import torch, time
dev = torch.device("cuda")
MB = 1 << 20
def timeit(fn, n=200):
    for _ in range(10): fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3  # us
nbuf = 6
w = [torch.empty(126 * MB, dtype=torch.uint8, device=dev) for _ in range(nbuf)]
r = [torch.empty(58 * MB, dtype=torch.uint8, device=dev) for _ in range(nbuf)]
r2 = [torch.empty(63 * MB, dtype=torch.uint8, device=dev) for _ in range(nbuf)]
t = timeit(lambda i: w[i % nbuf].fill_(1)); print(f"fill 126 MB: {t:.1f} us  {126*MB/t/1e6:.2f} TB/s write")
t = timeit(lambda i: r[i % nbuf].view(torch.int32).sum()); print(f"sum 58 MB: {t:.1f} us  {58*MB/t/1e6:.2f} TB/s read")
t = timeit(lambda i: w[i % nbuf][:63*MB].copy_(r2[i % nbuf])); print(f"copy 63->63 MB: {t:.1f} us  {126*MB/t/1e6:.2f} TB/s total")
def mix(i):
    w[i % nbuf][:58*MB].copy_(r[i % nbuf]); w[i % nbuf][58*MB:].fill_(1)
t = timeit(mix); print(f"copy 58 + fill 68 MB (2 kernels): {t:.1f} us  {(58+126)*MB/t/1e6:.2f} TB/s total")
