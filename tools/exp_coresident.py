"""Experiment: does the fused post-process kernel (post.cu) run BESIDE the pre-process kernel, and what does each pay?

Times, at the bench workload (64 x 1080p -> 640^2 fp16, 64 x 25200x85 head tensors):
  pre-process alone, decode + NMS (two kernels) alone, post_kernel alone,
  post_kernel and pre-process on two streams (post first), and the pipeline's own fork/join step,
then takes a %globaltimer / %smid trace (tk_b200_debug_trace) of one concurrent step and writes it to
gpurun_out/<tag>_trace.npy together with a text summary: per-CTA start / end of the pre-process kernel, per-frame
phase stamps of the post-process kernel, and how many SMs host both.
"""
import ctypes as C
import sys

import numpy as np
import torch

sys.path.insert(0, "."); sys.path.insert(0, "tests")
import synth  # noqa: E402
from trackiellm_b200 import _capi as capi  # noqa: E402
from trackiellm_b200 import vision as V  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else "r2_cores"
arith = {"exact": V.ARITH_EXACT, "half": V.ARITH_HALF}[sys.argv[2] if len(sys.argv) > 2 else "exact"]
B = 64
dev = torch.device("cuda", 0)
frames = torch.randint(0, 256, (B, 1080, 1920, 3), dtype=torch.uint8, device=dev)
logits = torch.from_numpy(np.stack([synth.logits_tensor(i, adversarial=(i == 0)) for i in range(B)])).to(dev)
out = torch.empty((B, 3, 640, 640), dtype=torch.float16, device=dev)
lib = capi.lib()
lib.tk_b200_debug_trace.argtypes = [C.c_void_p, C.c_size_t]
lib.tk_b200_debug_trace.restype = None


def pre():
    V.preprocess_batch(frames, 640, 640, out_dtype=torch.float16, arith=arith, out=out)


def post():
    return V.postprocess_fused(logits, 0.5, 0.5, 500)


def two():
    c, n = V.decode_detections(logits, 0.5)
    return V.nms_classwise(c, n, 0.5, 500)


def time_us(fn, reps=50):
    for _ in range(5):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


side = torch.cuda.Stream(priority=-1)
main = torch.cuda.current_stream()


def both():
    side.wait_stream(main)
    with torch.cuda.stream(side):
        post()
    pre()
    main.wait_stream(side)


print(f"pre-process alone      : {time_us(pre):7.1f} us")
print(f"decode + nms (2 kernels): {time_us(two):7.1f} us")
print(f"post_kernel alone      : {time_us(post):7.1f} us")
print(f"post || pre (2 streams): {time_us(both):7.1f} us")

pipes = {}
for mode in (1, 2):
    with V.Pipeline(max_batch=B, width=1920, height=1080, out_dtype=torch.float16, arith=arith, fuse_postprocess=mode) as pipe:
        dt, dd, dc = pipe.alloc_outputs(B)
        us = time_us(lambda: pipe.process_device(frames, logits, dt, dd, dc))
        print(f"pipeline step, fuse={mode} : {us:7.1f} us")

# ---- trace of one concurrent step ---------------------------------------------------------------
cap = 4096
buf = torch.zeros((cap + 1) * 8, dtype=torch.int64, device=dev)


def traced(fn, name):
    buf.zero_()
    torch.cuda.synchronize()
    lib.tk_b200_debug_trace(C.c_void_p(buf.data_ptr()), C.c_size_t(buf.numel() * 8))
    fn()
    torch.cuda.synchronize()
    lib.tk_b200_debug_trace(None, 0)
    h = buf.cpu().numpy().view(np.uint64)
    n = int(h[0])
    rec = h[8:8 + 8 * n].reshape(n, 8).copy()
    np.save(f"gpurun_out/{tag}_{name}_trace.npy", rec)
    kid = (rec[:, 0] >> np.uint64(32)).astype(np.int64)
    smid = (rec[:, 1] >> np.uint64(32)).astype(np.int64)
    t = rec[:, 2:].astype(np.int64)
    t0 = t[:, 0][t[:, 0] > 0].min()
    print(f"--- trace '{name}': {n} records")
    p = kid == 1
    if p.any():
        s, e = (t[p, 0] - t0) / 1e3, (t[p, 1] - t0) / 1e3
        print(f"pre-process CTAs: {p.sum()}  start {s.min():.1f}..{s.max():.1f} us  end {e.min():.1f}..{e.max():.1f} us "
              f"(median {np.median(e):.1f}); on {len(set(smid[p]))} SMs")
    q = kid == 2
    if q.any():
        tq = (t[q, :5] - t0) / 1e3
        ph = np.diff(tq, axis=1)
        print(f"post frames: {q.sum()}  start {tq[:, 0].min():.1f}..{tq[:, 0].max():.1f} us  end {tq[:, 4].min():.1f}..{tq[:, 4].max():.1f} us; "
              f"on {len(set(smid[q]))} SMs")
        for j, nm in enumerate(("pass 1", "prefix + expand", "pass 2", "NMS")):
            print(f"   {nm:16s}: median {np.median(ph[:, j]):6.1f} us  max {ph[:, j].max():6.1f} us")
        if p.any():
            both_sm = set(smid[p]) & set(smid[q])
            e_with = np.array([(t[p, 1] - t0)[i] / 1e3 for i in range(p.sum()) if smid[p][i] in set(smid[q])])
            e_wo = np.array([(t[p, 1] - t0)[i] / 1e3 for i in range(p.sum()) if smid[p][i] not in set(smid[q])])
            print(f"   SMs hosting both kernels: {len(both_sm)}; pre-process CTA end on those: median {np.median(e_with) if len(e_with) else -1:.1f} us, "
                  f"elsewhere: median {np.median(e_wo) if len(e_wo) else -1:.1f} us")
            from collections import Counter
            cnt = Counter(smid[q])
            print(f"   post CTAs per SM: max {max(cnt.values())}")


traced(pre, "pre_alone")
traced(post, "post_alone")
traced(both, "both")
with V.Pipeline(max_batch=B, width=1920, height=1080, out_dtype=torch.float16, arith=arith, fuse_postprocess=2) as pipe:
    dt, dd, dc = pipe.alloc_outputs(B)
    for _ in range(3):
        pipe.process_device(frames, logits, dt, dd, dc)
    traced(lambda: pipe.process_device(frames, logits, dt, dd, dc), "pipeline")
