"""Experiment: decode reading the logits in place from pinned host memory (zero-copy over PCIe)
vs cudaMemcpyAsync + decode from HBM."""
import ctypes as C
import sys
import time

import numpy as np
import torch

sys.path.insert(0, "."); sys.path.insert(0, "tests")
import synth
from trackiellm_b200 import _capi as capi
from trackiellm_b200 import vision as V

B = 32
lg = np.stack([synth.logits_tensor(i) for i in range(8)])
lg = np.concatenate([lg] * (B // 8))
h = torch.from_numpy(lg).pin_memory()
d = torch.empty_like(h, device="cuda")
cands = torch.empty((B, 4096, 7), dtype=torch.int32, device="cuda")
counts = torch.empty((B,), dtype=torch.int32, device="cuda")
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)


def dec(ptr):
    prm = capi.DecodeParams(ptr, B, 25200, 80, 0.5, 0, cands.data_ptr(), counts.data_ptr(), 4096)
    capi.check(capi.lib().tk_kernels_decode_detections(C.byref(prm), st), "decode")


def timeit(fn, n=10):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


def copy_then_decode():
    d.copy_(h, non_blocking=True)
    dec(d.data_ptr())


t_copy = timeit(lambda: d.copy_(h, non_blocking=True))
t_cd = timeit(copy_then_decode)
ref = counts.cpu().clone(); refc = cands.cpu().clone()
t_zc = timeit(lambda: dec(h.data_ptr()))
ok = torch.equal(counts.cpu(), ref) and torch.equal(cands.cpu()[:, :400], refc[:, :400])
print(f"B={B}: H2D copy {t_copy:.3f} ms ({h.numel()*4/t_copy/1e6:.1f} GB/s), copy+decode {t_cd:.3f} ms, zero-copy decode {t_zc:.3f} ms, same={ok}")
