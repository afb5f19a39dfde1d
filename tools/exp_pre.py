"""Experiment: pre-process kernel time per arithmetic mode at the bench workload (64 x 1080p -> 640^2 fp16)."""
import sys, torch, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from trackiellm_b200 import vision as V
fr = torch.randint(0, 256, (64, 1080, 1920, 3), dtype=torch.uint8, device="cuda")
out = torch.empty((64, 3, 640, 640), dtype=torch.float16, device="cuda")
for name, ar in (("EXACT", V.ARITH_EXACT), ("FAST", V.ARITH_FAST), ("HALF", V.ARITH_HALF)):
    for _ in range(5):
        V.preprocess_batch(fr, 640, 640, out_dtype=torch.float16, arith=ar, out=out)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(100):
        V.preprocess_batch(fr, 640, 640, out_dtype=torch.float16, arith=ar, out=out)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 10
    print(f"{name}: {us:.1f} us/launch, {64*8678400/us/1e3:.0f} GB/s algorithmic ({64*8678400/us/1e3/6538*100:.1f}% of 6538)")
