"""How does the one-kernel post-process (tk_kernels_postprocess_fused, layout 0) compare with decode_kernel + nms_kernel
when frames are CROWDED (more than 512 candidates each: its NMS tail then works in a global area, nms_kernel in shared
memory)?  64 x 25200x85, thresholds chosen for ~480 / ~1500 / ~3000 candidates per frame."""
import sys

import numpy as np
import torch

sys.path.insert(0, "."); sys.path.insert(0, "tests")
import synth  # noqa: E402
from trackiellm_b200 import vision as V  # noqa: E402

B = 64
dev = torch.device("cuda", 0)


def time_us(fn, reps=30):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


for objects, dups, thr in ((40, 12, 0.5), (120, 12, 0.5), (250, 12, 0.5)):
    logits = torch.from_numpy(np.stack([synth.logits_tensor(i, objects=objects, dups=dups) for i in range(B)])).to(dev)
    c, n = V.decode_detections(logits, thr)
    torch.cuda.synchronize()
    two = time_us(lambda: V.nms_classwise(*V.decode_detections(logits, thr), 0.5, 500))
    one = time_us(lambda: V.postprocess_fused(logits, thr, 0.5, 500, layout=0))
    print(f"candidates per frame ~{int(n.float().mean())} (max {int(n.max())}): decode + nms {two:.1f} us, fused {one:.1f} us")
