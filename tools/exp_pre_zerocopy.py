"""Experiment: pre-process kernel reading page-locked HOST frames in place (TMA rows over PCIe) vs copy + kernel.
Pays when the geometry is sparse (y_ratio >= 2: only two source rows per output row are ever fetched)."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from trackiellm_b200 import _capi as capi
from trackiellm_b200 import vision as V


def run(name, w, h, fmt, bpp_bytes, n):
    fb = int(w * h * bpp_bytes)
    host = torch.from_numpy(np.random.default_rng(0).integers(0, 256, (n, fb), dtype=np.uint8)).pin_memory()
    dev = torch.empty_like(host, device="cuda")
    out = torch.empty((n, 3, 640, 640), dtype=torch.float16, device="cuda")
    out2 = torch.empty_like(out)

    def launch(ptr, o):
        p = capi.PreprocessBatchParams(
            d_input=ptr, d_input_ptrs=None, input_frame_pitch_bytes=fb, num_frames=n, input_width=w, input_height=h,
            input_stride_bytes=0, input_format=fmt, d_output=o.data_ptr(), output_width=640, output_height=640,
            output_type=V.OUT_F16, resize_mode=V.RESIZE_STRETCH, pad_value=114.0, arith=V.ARITH_HALF,
            mean=V._f3(V.IMAGENET_MEAN), std_dev=V._f3(V.IMAGENET_STD))
        capi.check(capi.lib().tk_b200_preprocess_batch(C.byref(p), V._stream()), "pre")

    def copy_then_kernel():
        dev.copy_(host, non_blocking=True)
        launch(dev.data_ptr(), out)

    def in_place():
        launch(host.data_ptr(), out2)

    res = {}
    for nm, fn in (("copy + kernel", copy_then_kernel), ("in place", in_place)):
        for _ in range(2): fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5): fn()
        torch.cuda.synchronize()
        res[nm] = (time.perf_counter() - t0) / 5 * 1e3
    same = torch.equal(out, out2)
    print(f"{name}: copy + kernel {res['copy + kernel']:.2f} ms, in place {res['in place']:.2f} ms, identical={same}, "
          f"{n * fb / 1e6:.0f} MB per batch")


run("32 x 4K I420 -> 640x640", 3840, 2160, V.PIXEL_I420, 1.5, 32)
run("32 x 4K RGB8 -> 640x640", 3840, 2160, V.PIXEL_RGB8, 3, 32)
run("64 x 1080p RGB8 -> 640x640", 1920, 1080, V.PIXEL_RGB8, 3, 64)
