"""decode_kernel + nms_kernel alone at the bench workload (64 x 25200x85), for ncu captures and A/B of decode variants."""
import sys
import numpy as np
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import synth
from trackiellm_b200 import vision as V

B = 64
logits = torch.from_numpy(np.stack([synth.logits_tensor(i, adversarial=(i == 0)) for i in range(B)])).cuda()


def time_us(fn, iters=100):
    for _ in range(5):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


c, n = V.decode_detections(logits, 0.5)
print(f"decode_kernel : {time_us(lambda: V.decode_detections(logits, 0.5, candidates=c, counts=n)):.1f} us")
d, m = V.nms_classwise(c, n, 0.5, 500)
print(f"nms_kernel    : {time_us(lambda: V.nms_classwise(c, n, 0.5, 500, detections=d, num_detections=m)):.1f} us")
print(f"decode + nms  : {time_us(lambda: V.nms_classwise(*V.decode_detections(logits, 0.5, candidates=c, counts=n), 0.5, 500, detections=d, num_detections=m)):.1f} us")
print("candidates per frame:", n.cpu().numpy()[:8], "... detections:", m.cpu().numpy()[:8])
