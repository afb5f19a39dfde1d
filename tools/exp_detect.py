"""Where does one tk_object_detector_detect call (640x480 host frame, BASELINE configs[0]) spend its time?
TK_B200_DETECT_TRACE=1 makes the library synchronise after every phase and print the split; without it the script
prints the wall clock per blocking call (median of 300)."""
import ctypes as C
import sys
import time

import torch

sys.path.insert(0, "."); sys.path.insert(0, "tests")
import synth  # noqa: E402
from trackiellm_b200 import vision as V  # noqa: E402

dev = torch.device("cuda", 0)
cu = C.CDLL("libcuda.so.1")
frame = synth.rgb_frame(640, 480, 0)
lg = torch.from_numpy(synth.logits_tensor(0, adversarial=True)).to(dev)
P = lg.shape[0]


def net(d_in, n, iw, ih, d_out, p, l, stream):
    return cu.cuMemcpyDtoDAsync_v2(C.c_uint64(d_out), C.c_uint64(lg.data_ptr()), C.c_size_t(n * p * l * 4), C.c_void_p(stream))


with V.ObjectDetector(device_id=0, class_labels=[f"c{i}" for i in range(80)], conf_thr=0.5, iou_thr=0.5) as det:
    det.bind_network(net, P)
    for _ in range(20):
        r = det.detect(frame)
    ts = []
    for _ in range(300):
        t0 = time.perf_counter()
        r = det.detect(frame)
        ts.append((time.perf_counter() - t0) * 1e6)
ts.sort()
print(f"detect: median {ts[150]:.1f} us, p10 {ts[30]:.1f}, p90 {ts[270]:.1f}; {len(r)} detections")
