"""Batch / geometry sweep of the device-resident step and of the pre-process kernel alone
(BASELINE.json configs 4 and 5).  Prints a markdown table; run on the GPU box."""
import sys, time
import numpy as np, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import synth
from trackiellm_b200 import vision as V

PEAK = 6538.0


def time_us(fn, iters):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def pre_geometry():
    print("| input | output | batch | us/launch | frames/s | algorithmic GB/s | % of 6538 |")
    print("|---|---|---|---|---|---|---|")
    for name, w, h, fmt, dt in (("1080p RGB8", 1920, 1080, V.PIXEL_RGB8, torch.float16), ("1080p RGB8", 1920, 1080, V.PIXEL_RGB8, torch.float32),
                                ("640x480 RGB8", 640, 480, V.PIXEL_RGB8, torch.float32), ("4K RGB8", 3840, 2160, V.PIXEL_RGB8, torch.float16),
                                ("4K I420", 3840, 2160, V.PIXEL_I420, torch.float16), ("4K I420", 3840, 2160, V.PIXEL_I420, torch.float32)):
        n = 32
        fb = w * h * 3 if fmt == V.PIXEL_RGB8 else w * h * 3 // 2
        fr = torch.randint(0, 256, (n, fb), dtype=torch.uint8, device="cuda")
        out = torch.empty((n, 3, 640, 640), dtype=dt, device="cuda")
        us = time_us(lambda: V.preprocess_batch(fr, 640, 640, width=w, height=h, fmt=fmt, out_dtype=dt, out=out), 50)
        alg = n * (fb + 640 * 640 * 3 * out.element_size())
        print(f"| {name} | 640x640 {'fp16' if dt == torch.float16 else 'fp32 (EXACT)'} | {n} | {us:.1f} | {n / us * 1e6:.0f} | {alg / us / 1e3:.0f} | {alg / us / 1e3 / PEAK * 100:.1f} |")


def batch_sweep():
    print("\n| batch | step us | frames/s | pre us | decode us | nms us |")
    print("|---|---|---|---|---|---|")
    base_f = torch.randint(0, 256, (64, 1080, 1920, 3), dtype=torch.uint8, device="cuda")
    lg = np.stack([synth.logits_tensor(i) for i in range(8)])
    base_l = torch.from_numpy(np.concatenate([lg] * 8)).cuda()
    for b in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024):
        reps = (b + 63) // 64
        fr = base_f.repeat(reps, 1, 1, 1)[:b].contiguous() if b > 64 else base_f[:b]
        ll = base_l.repeat(reps, 1, 1)[:b].contiguous() if b > 64 else base_l[:b]
        pipe = V.Pipeline(max_batch=b)
        t, d, c = pipe.alloc_outputs(b)
        iters = 100 if b <= 64 else 20
        for _ in range(3):
            pipe.process_device(fr, ll, t, d, c)
        torch.cuda.synchronize()
        pipe.stage_timing(iters)
        us = time_us(lambda: pipe.process_device(fr, ll, t, d, c), iters)
        ms, n = pipe.stage_times()
        pipe.stage_timing(0)
        print(f"| {b} | {us:.1f} | {b / us * 1e6:.0f} | {ms[0] / n * 1e3:.1f} | {ms[1] / n * 1e3:.1f} | {ms[2] / n * 1e3:.1f} |")
        pipe.close(); del t, d, c, fr, ll
        torch.cuda.empty_cache()


def other_kernels():
    """HBM-bound secondary kernels: algorithmic bytes = every input byte once + every
    output byte once; buffers of >= 256 MB per launch so every iteration streams from HBM."""
    print("\n| kernel | workload | us/launch | algorithmic GB/s | % of 6538 |")
    print("|---|---|---|---|---|")
    rows = []
    n, h, w = 1024, 256, 256                                   # 1024 MiDaS maps of 256x256: 268 MB in, 268 MB out
    inv = torch.rand((n, h, w), device="cuda") * 900 + 3
    out = torch.empty_like(inv)
    us = time_us(lambda: V.inverse_depth_to_metric(inv, out=out), 30)
    rows.append(("inverse_depth_metric_kernel", f"{n} maps {h}x{w} f32 (min/max + convert, 8 B/element)", us, 8 * inv.numel()))
    n2, h2, w2 = 256, 480, 640
    inv2 = torch.rand((n2, h2, w2), device="cuda") * 900 + 3
    out2 = torch.empty_like(inv2)
    us = time_us(lambda: V.inverse_depth_to_metric(inv2, out=out2), 30)
    rows.append(("inverse_depth_metric_kernel", f"{n2} maps {h2}x{w2} f32", us, 8 * inv2.numel()))
    big = inv.view(n * h, w)                                    # one 262144 x 256 map for the single-map reference launchers
    us = time_us(lambda: V.kernels_postprocess_depth_map(big, 0.37, 1.5), 30)
    rows.append(("depth_scale_kernel", "67 M elements f32 (8 B/element)", us, 8 * big.numel()))
    us = time_us(lambda: V.kernels_depth_to_point_cloud(big, 300.0, 310.0, 128.0, 128.0), 30)
    rows.append(("depth_points_kernel", "67 M elements f32 -> xyz (16 B/element)", us, 16 * big.numel()))
    x = torch.randn((4 * 1024 * 1024, 80), device="cuda")
    us = time_us(lambda: V.kernels_softmax(x), 20)
    rows.append(("softmax_tile_kernel", "4 M rows x 80 f32 (8 B/element)", us, 8 * x.numel()))
    for name, wl, us, alg in rows:
        print(f"| {name} | {wl} | {us:.1f} | {alg / us / 1e3:.0f} | {alg / us / 1e3 / PEAK * 100:.1f} |")
    # latency-bound: per-detection distance over the pipeline's output shape
    nf, md = 64, 500
    depth = torch.rand((nf, 256, 256), device="cuda") * 20 + 0.2
    rng = np.random.default_rng(0)
    d = np.zeros((nf, md), V.CAND_DTYPE)
    d["x"] = rng.integers(0, 500, (nf, md)); d["y"] = rng.integers(0, 500, (nf, md))
    d["w"] = rng.integers(16, 320, (nf, md)); d["h"] = rng.integers(16, 320, (nf, md))
    dets = torch.from_numpy(d.view(np.int32).reshape(nf, md, 7)).cuda()
    cnt = torch.full((nf,), 56, dtype=torch.int32, device="cuda")
    us = time_us(lambda: V.detection_distance(dets, cnt, depth, 640, 640), 20)
    print(f"\ndetection_distance_kernel: {nf} frames x 56 detections (boxes 16..320 px of 640, depth maps 256x256): "
          f"{us:.1f} us/launch = {nf * 56 / us:.2f} detections/us (latency-bound, shared-memory sort)")


def detect_latency():
    """BASELINE config 0 through the reference's own call: tk_object_detector_detect on one HOST 640x480 frame
    (upload + pre-process + bound network + decode + NMS + results), wall clock per call.  The network is a stand-in
    (device-to-device copy of a prepared 25200x85 head tensor, enqueued from a Python ctypes callback)."""
    import ctypes as C
    cu = C.CDLL("libcuda.so.1")
    print("\n| frame | detect() wall us (median of 200) | p90 us | detections |")
    print("|---|---|---|---|")
    for (w, h) in ((640, 480), (1920, 1080), (3840, 2160)):
        frame = synth.rgb_frame(w, h, 0)
        lg = torch.from_numpy(synth.logits_tensor(0)).cuda()

        def net(d_in, n, iw, ih, d_out, p, l, stream):
            return cu.cuMemcpyDtoDAsync_v2(C.c_uint64(d_out), C.c_uint64(lg.data_ptr()), C.c_size_t(n * p * l * 4), C.c_void_p(stream))

        with V.ObjectDetector(class_labels=[f"c{i}" for i in range(80)]) as det:
            det.bind_network(net, 25200)
            for _ in range(10):
                r = det.detect(frame)
            ts = []
            for _ in range(200):
                t0 = time.perf_counter(); r = det.detect(frame); ts.append((time.perf_counter() - t0) * 1e6)
            ts.sort()
            print(f"| {w}x{h} RGB8 (pageable host memory) | {ts[100]:.0f} | {ts[180]:.0f} | {len(r)} |")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "detect":
        detect_latency()
    elif len(sys.argv) > 1 and sys.argv[1] == "others":
        other_kernels()
    elif len(sys.argv) > 1 and sys.argv[1] == "batch":
        batch_sweep()
    else:
        pre_geometry()
        batch_sweep()
        other_kernels()
        detect_latency()
