"""Pre-process kernel alone at the geometries BASELINE.md §4 sets targets for, every arithmetic mode (32 frames per launch)."""
import sys
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from trackiellm_b200 import vision as V

PEAK = 6548.8


def time_us(fn, iters=50):
    for _ in range(5):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


print("| input | output | arith | batch | us/launch | frames/s | algorithmic GB/s | % of 6548.8 |")
print("|---|---|---|---|---|---|---|---|")
for name, w, h, fmt in (("640x480 RGB8", 640, 480, V.PIXEL_RGB8), ("1080p RGB8", 1920, 1080, V.PIXEL_RGB8),
                        ("4K RGB8", 3840, 2160, V.PIXEL_RGB8), ("4K I420", 3840, 2160, V.PIXEL_I420)):
    n = 64 if w <= 1920 else 32
    fb = w * h * 3 if fmt == V.PIXEL_RGB8 else w * h * 3 // 2
    fr = torch.randint(0, 256, (n, fb), dtype=torch.uint8, device="cuda")
    for dt in (torch.float32, torch.float16):
        out = torch.empty((n, 3, 640, 640), dtype=dt, device="cuda")
        for an, ar in (("EXACT", V.ARITH_EXACT), ("FAST", V.ARITH_FAST), ("HALF", V.ARITH_HALF)):
            if ar == V.ARITH_HALF and (dt != torch.float16 or fmt == V.PIXEL_I420):
                continue
            us = time_us(lambda: V.preprocess_batch(fr, 640, 640, width=w, height=h, fmt=fmt, out_dtype=dt, arith=ar, out=out))
            alg = n * (fb + 640 * 640 * 3 * out.element_size())
            print(f"| {name} | 640x640 {'fp16' if dt == torch.float16 else 'fp32'} | {an} | {n} | {us:.1f} | {n / us * 1e6:.0f} | "
                  f"{alg / us / 1e3:.0f} | {alg / us / 1e3 / PEAK * 100:.1f} |")
