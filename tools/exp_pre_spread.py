"""Spread of the pre-process kernel's CTA end times (static row partition): %globaltimer stamps of every CTA
(tk_b200_debug_trace), 64 x 1080p -> 640^2 fp16, alone and inside the pipeline step."""
import ctypes as C
import sys

import numpy as np
import torch

sys.path.insert(0, "."); sys.path.insert(0, "tests")
import synth  # noqa: E402
from trackiellm_b200 import _capi as capi  # noqa: E402
from trackiellm_b200 import vision as V  # noqa: E402

B = 64
dev = torch.device("cuda", 0)
frames = torch.randint(0, 256, (B, 1080, 1920, 3), dtype=torch.uint8, device=dev)
logits = torch.from_numpy(np.stack([synth.logits_tensor(i, adversarial=(i == 0)) for i in range(B)])).to(dev)
out = torch.empty((B, 3, 640, 640), dtype=torch.float16, device=dev)
lib = capi.lib()
lib.tk_b200_debug_trace.argtypes = [C.c_void_p, C.c_size_t]
lib.tk_b200_debug_trace.restype = None
buf = torch.zeros((4096 + 1) * 8, dtype=torch.int64, device=dev)


def traced(fn, name):
    buf.zero_()
    torch.cuda.synchronize()
    lib.tk_b200_debug_trace(C.c_void_p(buf.data_ptr()), C.c_size_t(buf.numel() * 8))
    fn()
    torch.cuda.synchronize()
    lib.tk_b200_debug_trace(None, 0)
    h = buf.cpu().numpy().view(np.uint64)
    n = int(h[0])
    rec = h[8:8 + 8 * n].reshape(n, 8)
    kid = (rec[:, 0] >> np.uint64(32)).astype(np.int64)
    smid = (rec[:, 1] >> np.uint64(32)).astype(np.int64)
    t = rec[:, 2:].astype(np.int64)
    p = kid == 1
    t0 = t[p, 0].min()
    s, e = (t[p, 0] - t0) / 1e3, (t[p, 1] - t0) / 1e3
    d = e - s
    print(f"{name}: {p.sum()} CTAs on {len(set(smid[p]))} SMs; start {s.min():.1f}..{s.max():.1f} us (median {np.median(s):.1f}); "
          f"end min {e.min():.1f} / p10 {np.percentile(e, 10):.1f} / median {np.median(e):.1f} / p90 {np.percentile(e, 90):.1f} / max {e.max():.1f} us; "
          f"duration median {np.median(d):.1f}, max {d.max():.1f}")


for arith, nm in ((V.ARITH_EXACT, "EXACT"), (V.ARITH_HALF, "HALF")):
    def pre():
        V.preprocess_batch(frames, 640, 640, out_dtype=torch.float16, arith=arith, out=out)
    for _ in range(5):
        pre()
    traced(pre, f"pre-process alone, {nm}")
    with V.Pipeline(max_batch=B, width=1920, height=1080, out_dtype=torch.float16, arith=arith) as pipe:
        dt, dd, dc = pipe.alloc_outputs(B)
        for _ in range(5):
            pipe.process_device(frames, logits, dt, dd, dc)

        def step2():
            pipe.process_device(frames, logits, dt, dd, dc)
            pipe.process_device(frames, logits, dt, dd, dc)
        traced(step2, f"two pipeline steps, {nm} (both steps' CTAs; the second step's start under the first one's NMS)")
