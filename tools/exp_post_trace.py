"""Where do decode_kernel and nms_kernel spend their time?  %globaltimer phase stamps (TRACE build: make -C
trackiellm_b200/csrc trace; run with TK_B200_LIB=trackiellm_b200/libtk_b200_trace.so) of one decode + NMS of the bench
batch (64 x 25200x85), relative to the first stamp of each kernel."""
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
logits = torch.from_numpy(np.stack([synth.logits_tensor(i, adversarial=(i == 0)) for i in range(B)])).to(dev)
lib = capi.lib()
lib.tk_b200_debug_trace.argtypes = [C.c_void_p, C.c_size_t]
lib.tk_b200_debug_trace.restype = None
cap = 4096
buf = torch.zeros((cap + 1) * 8, dtype=torch.int64, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


FUSED = len(sys.argv) > 1 and sys.argv[1] == "fused"     # tk_kernels_postprocess_fused layout 0: one kernel, NMS in the cluster's first CTA


def two():
    if FUSED:
        return V.postprocess_fused(logits, 0.5, 0.5, 500, layout=0)
    c, n = V.decode_detections(logits, 0.5)
    return V.nms_classwise(c, n, 0.5, 500)


for _ in range(5):
    two()
for rep in range(3):
    flush.zero_()                                   # cold L2, as inside the step
    buf.zero_()
    torch.cuda.synchronize()
    lib.tk_b200_debug_trace(C.c_void_p(buf.data_ptr()), C.c_size_t(buf.numel() * 8))
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    if FUSED:
        V.postprocess_fused(logits, 0.5, 0.5, 500, layout=0)
        e1.record()
    else:
        c, n = V.decode_detections(logits, 0.5)
        e1.record()
        V.nms_classwise(c, n, 0.5, 500)
    e2.record()
    torch.cuda.synchronize()
    lib.tk_b200_debug_trace(None, 0)
    h = buf.cpu().numpy().view(np.uint64)
    nrec = int(h[0])
    rec = h[8:8 + 8 * nrec].reshape(nrec, 8)
    kid = (rec[:, 0] >> np.uint64(32)).astype(np.int64)
    t = rec[:, 2:].astype(np.int64)
    print(f"=== run {rep}: events decode {e0.elapsed_time(e1) * 1e3:.1f} us, nms {e1.elapsed_time(e2) * 1e3:.1f} us; {nrec} records")
    for k, name, phases in ((3, "decode_kernel CTAs", ("pass 1 (scan)", "cluster wait + prefix", "expand", "pass 2 (rows)", "cluster wait + write-out")),
                            (4, "nms frames", ("two sorts", "segment scan", "pair masks", "resolve", "compaction"))):
        m = kid == k
        if not m.any():
            continue
        tk = t[m]
        t0 = t[kid == 3][:, 0].min() if FUSED else tk[:, 0].min()     # fused: NMS stamps on the decode CTAs' clock
        rel = (tk - t0) / 1e3
        print(f"{name}: {m.sum()}  start {rel[:, 0].min():.1f}..{rel[:, 0].max():.1f} us, end {rel[:, 5].min():.1f}..{rel[:, 5].max():.1f} us (median {np.median(rel[:, 5]):.1f})")
        ph = np.diff(rel, axis=1)
        if k == 3 and not FUSED:
            # scan end per CTA, grouped by cluster (8 consecutive CTAs = one frame): is the spread inside clusters or between them?
            blk = (rec[m][:, 0] & np.uint64(0xffffffff)).astype(np.int64)
            order = np.argsort(blk)
            se = rel[order, 1].reshape(-1, 8)
            print(f"   scan end: all CTAs min {se.min():.1f} / median {np.median(se):.1f} / max {se.max():.1f}; per cluster: mean of (max - min) {np.mean(se.max(1) - se.min(1)):.2f} us, "
                  f"cluster means min {se.mean(1).min():.1f} / max {se.mean(1).max():.1f}, cluster max: min {se.max(1).min():.1f} / median {np.median(se.max(1)):.1f} / max {se.max(1).max():.1f}")
            smid = (rec[m][:, 1] >> np.uint64(32)).astype(np.int64)[order].reshape(-1, 8)
            print(f"   SMs of cluster 0: {smid[0].tolist()}, rank-wise mean scan end: {np.round(se.mean(0), 1).tolist()}")
        for j, nm in enumerate(phases):
            print(f"   {nm:26s}: median {np.median(ph[:, j]):6.2f} us  p90 {np.percentile(ph[:, j], 90):6.2f}  max {ph[:, j].max():6.2f}   (phase ends at median {np.median(rel[:, j + 1]):6.1f} us)")
