"""Experiment: HBM throughput of plain torch streams by read:write mix on this B200 (ceiling for the write-heavy kernels)."""
import torch
n = 64 * 1024 * 1024
a = torch.empty(n, dtype=torch.float32, device="cuda").normal_()
b = torch.empty(n, dtype=torch.float32, device="cuda")
c = torch.empty((n, 3), dtype=torch.float32, device="cuda")


def t(fn, iters=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


for name, fn, byts in (("write only (fill_)", lambda: c.fill_(1.0), c.numel() * 4),
                       ("read only (sum)", lambda: a.sum(), n * 4),
                       ("1 read : 1 write (copy_)", lambda: b.copy_(a), 2 * n * 4),
                       ("1 read : 3 write (expand copy)", lambda: c.copy_(a.view(n, 1).expand(n, 3)), 4 * n * 4)):
    s = t(fn)
    print(f"{name:34s} {byts / s / 1e9:7.0f} GB/s")
