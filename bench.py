#!/usr/bin/env python
"""bench.py — frames/s of the vision hot path (pre-process + decode + class-wise NMS, 1080p -> 640x640).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one pass of the hot path over one batch of synthetic input per GPU
(BASELINE.json configs[1] geometry: 64 frames of 1920x1080 RGB8 -> 640x640 fp16 NCHW, plus
decode + class-wise NMS of 64 detector-head tensors of 25200x85 fp32, configs[2] shape).
Frames shard by frame across ranks with no collective: weak scaling, every rank
owns its own 64-frame batch; torch.distributed (gloo) is used only for the barrier and the max.

  value     whole-job frames/s with inputs resident in HBM, CUDA-event timed, max over ranks
  e2e       the same metric through tk_b200_pipeline_process_host: pinned HOST frames + logits in,
            HOST detections out, every host<->device copy inside the timed region
  roofline  the dominant kernel's algorithmic bytes / its CUDA-event time vs the measured HBM peak
  cpu_baseline / --impl reference
            the reference's own CPU pre-processor (oracle/_ref/libtkref.so, compiled from
            /root/reference/src/vision/tk_image_preprocessor.c) + the CPU restatement of
            decode/NMS (the reference's postprocess is a stub) on all host cores

The oracle is imported only by cpu_baseline() / the reference arm.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "frames/s (pre+decode+NMS, 1080p->640^2)"
SRC_W, SRC_H = 1920, 1080
DST = 640
P, NC = 25200, 80
FRAME_BYTES = SRC_W * SRC_H * 3
TENSOR_BYTES_F16 = DST * DST * 3 * 2
LOGIT_BYTES = P * (5 + NC) * 4
# every source byte once + every output byte once; decode+NMS: every logit byte once.
ALG_PRE = FRAME_BYTES + TENSOR_BYTES_F16          # 8 678 400 B / frame
ALG_POST = LOGIT_BYTES                            # 8 568 000 B / frame
CONF_THR, IOU_THR, MAX_DET = 0.5, 0.5, 500        # tests/tk_vision_full_test.c:67, tk_vision_pipeline.c:376-377
I420 = False
SRC_NAME = "1080p RGB8"


def frame_bytes_over_pcie() -> tuple[int, bool]:
    """Bytes of one frame that cross PCIe in the host-buffer path, and whether the frames are read in place: sparse
    geometries (y_ratio >= 2) with 16-byte rows are pulled row by row by the pre-process kernel, which only ever
    stages the two source rows of each output row (src/vision/tk_image_preprocessor.c:49,51) — otherwise the whole
    frame is copied.  Mirrors tkb::preprocess_reads_rows_sparsely (csrc/preprocess.cu)."""
    row = SRC_W * (1 if I420 else 3)
    y_ratio = np.float32(np.float32(SRC_H) - np.float32(1)) / np.float32(DST)
    if y_ratio < 2 or row % 16 or (I420 and (SRC_W // 2) % 16):
        return FRAME_BYTES, False
    ys = np.minimum((y_ratio * np.arange(DST, dtype=np.float32)).astype(np.int64), SRC_H - 2)
    luma = np.unique(np.concatenate([ys, ys + 1]))
    total = len(luma) * row
    if I420:
        total += 2 * len(np.unique(np.minimum(luma >> 1, SRC_H // 2 - 1))) * (SRC_W // 2)
    return int(total), True


def set_workload(name: str) -> None:
    """Default = BASELINE.json configs[1]+[2] (the metric's configuration).  "4k-i420" = configs[3]: 4K planar
    YUV420 camera frames (src/ffi/c_api/tk_ffi_api.c:917-931) through the full pre + post pipeline."""
    global SRC_W, SRC_H, FRAME_BYTES, ALG_PRE, I420, METRIC, SRC_NAME
    if name == "4k-i420":
        SRC_W, SRC_H, I420, SRC_NAME = 3840, 2160, True, "4K I420"
        FRAME_BYTES = SRC_W * SRC_H * 3 // 2
        METRIC = "frames/s (pre+decode+NMS, 4K I420->640^2)"
    ALG_PRE = FRAME_BYTES + TENSOR_BYTES_F16


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def make_inputs(batch: int, seed0: int):
    """Seeded synthetic batch: `batch` random 1080p RGB8 frames + `batch` detector-head tensors."""
    import synth
    frames = np.empty((batch, FRAME_BYTES), np.uint8) if I420 else np.empty((batch, SRC_H, SRC_W, 3), np.uint8)
    logits = np.empty((batch, P, 5 + NC), np.float32)
    for i in range(batch):
        frames[i] = synth.i420_frame(SRC_W, SRC_H, seed0 + i) if I420 else synth.rgb_frame(SRC_W, SRC_H, seed0 + i)
        logits[i] = synth.logits_tensor(seed0 + i, adversarial=(i == 0))
    return frames, logits


# --------------------------------------------------------------------------- #
# CPU arm: the reference's own code on the host cores                           #
# --------------------------------------------------------------------------- #
class CpuPath:
    """Reference CPU path for one frame: tk_preprocessor_resize_and_normalize_to_chw (fp32 CHW, the
    reference has no fp16 output) + decode + class-wise NMS.  One frame per thread — the reference
    function is re-entrant."""

    def __init__(self, threads: int):
        from oracle import oracle as O
        self.O = O
        self.threads = threads
        self.kind = "reference" if O.have_ref() else "port"
        self.lib = O.lib()
        self.ref = O.ref_lib() if O.have_ref() else None
        self.mean = (C.c_float * 3)(*O.IMAGENET_MEAN)
        self.std = (C.c_float * 3)(*O.IMAGENET_STD)
        self.tls = threading.local()

    def _bufs(self):
        t = self.tls
        if not hasattr(t, "out"):
            t.out = np.empty((3, DST, DST), np.float32)
            t.cand = np.zeros(4096, self.O.CAND_DTYPE)
            t.kept = np.zeros(MAX_DET, self.O.CAND_DTYPE)
            t.rgb = np.empty((SRC_H, SRC_W, 3), np.uint8) if I420 else None
        return t

    def one(self, frame: np.ndarray, logits: np.ndarray) -> int:
        t = self._bufs()
        O = self.O
        if I420:      # the reference has no YUV arithmetic: the oracle's BT.601 conversion, then its pre-processor
            self.lib.oracle_i420_to_rgb8(frame.ctypes.data_as(C.c_void_p), C.c_uint32(SRC_W), C.c_uint32(SRC_H),
                                         t.rgb.ctypes.data_as(C.c_void_p))
            frame = t.rgb
        if self.ref is not None:
            vf = O._VideoFrame(SRC_W, SRC_H, SRC_W * 3, 0, frame.ctypes.data)
            rc = self.ref.tk_preprocessor_resize_and_normalize_to_chw(
                C.byref(vf), t.out.ctypes.data_as(C.c_void_p), C.c_uint32(DST), C.c_uint32(DST), self.mean, self.std)
        else:
            rc = self.lib.oracle_preprocess_rgb8(frame.ctypes.data_as(C.c_void_p), C.c_uint32(SRC_W), C.c_uint32(SRC_H),
                                                 C.c_uint32(SRC_W * 3), t.out.ctypes.data_as(C.c_void_p), C.c_uint32(DST),
                                                 C.c_uint32(DST), self.mean, self.std, C.c_int(0), C.c_int(0), C.c_float(0))
        assert rc == 0
        total = self.lib.oracle_decode(logits.ctypes.data_as(C.c_void_p), C.c_uint32(P), C.c_uint32(NC),
                                       C.c_float(CONF_THR), C.c_int(0), t.cand.ctypes.data_as(C.c_void_p), C.c_uint32(4096))
        n = self.lib.oracle_nms_classwise(t.cand.ctypes.data_as(C.c_void_p), C.c_uint32(min(total, 4096)),
                                          C.c_float(IOU_THR), C.c_uint32(MAX_DET), t.kept.ctypes.data_as(C.c_void_p))
        return int(n)

    def run_batch(self, frames: np.ndarray, logits: np.ndarray, pool) -> float:
        """All frames of the batch through the pool; returns seconds."""
        t0 = time.perf_counter()
        list(pool.map(lambda i: self.one(frames[i], logits[i]), range(frames.shape[0])))
        return time.perf_counter() - t0

    def describe(self, frames_done: int) -> str:
        pre = ("reference tk_preprocessor_resize_and_normalize_to_chw compiled from the reference tree (oracle/_ref, "
               "gcc -O2 -mavx, fp32 CHW out)" if self.kind == "reference"
               else "C restatement of the reference pre-processor (oracle/, fp32 CHW out)")
        return (f"{frames_done} frames of the bench workload ({SRC_NAME} -> 640x640 + 25200x85 decode + class-wise NMS), "
                f"{pre} + CPU restatement of decode/NMS (the reference's postprocess_detections is a stub), "
                f"one frame per thread, {self.threads} threads")


def cpu_baseline(frames: np.ndarray, logits: np.ndarray, target_cpu_seconds: float = 16.0) -> dict:
    """Bounded sample of the same workload on the host cores (rank 0, N=1 only)."""
    from concurrent.futures import ThreadPoolExecutor
    threads = os.cpu_count() or 1
    cpu = CpuPath(threads)
    with ThreadPoolExecutor(threads) as pool:
        sub = min(frames.shape[0], max(threads, 16))
        t_probe = cpu.run_batch(frames[:sub], logits[:sub], pool)            # warm-up + cost probe
        per_frame_core_s = t_probe * min(threads, sub) / sub
        reps = max(1, int(target_cpu_seconds / max(per_frame_core_s * frames.shape[0], 1e-9)))
        reps = min(reps, 16)
        t = 0.0
        for _ in range(reps):
            t += cpu.run_batch(frames, logits, pool)
    done = reps * frames.shape[0]
    return {"value": done / t, "unit": "frames/s", "cores": threads, "kind": cpu.kind, "sample": cpu.describe(done)}


# --------------------------------------------------------------------------- #
# Clock sampling during the timed region (NVML)                                 #
# --------------------------------------------------------------------------- #
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, device_index: int):
        self.samples, self.mask, self.ok = [], 0, False
        self._stop = threading.Event()
        self._on = threading.Event()
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.nv = pynvml
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # pragma: no cover
            log(f"[clocks] NVML unavailable: {e}")
            self.max_mhz = None
        self.thread = threading.Thread(target=self._run, daemon=True)
        if self.ok:
            self.thread.start()

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            if self._on.is_set():
                try:
                    self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                    self.mask |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                except Exception:
                    pass
            time.sleep(0.002)

    def resume(self):
        self._on.set()

    def pause(self):
        self._on.clear()

    def result(self) -> dict:
        self._stop.set()
        reasons = [n for b, n in self.REASONS.items() if self.mask & b and n != "gpu_idle"]
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": reasons, "samples": len(self.samples)}


# --------------------------------------------------------------------------- #
# Reference arm                                                                 #
# --------------------------------------------------------------------------- #
def config_dict(n_gpus: int, batch: int) -> dict:
    return {"workload": f"{SRC_NAME} x{batch} per GPU -> 640x640 fp16 NCHW (stretch-bilinear, ImageNet mean/std) + "
                        f"decode(raw, conf>{CONF_THR}) + class-wise NMS(IoU>{IOU_THR}, max {MAX_DET}) of {batch} tensors 25200x85 fp32",
            "frames_per_gpu": batch, "global_batch": batch * n_gpus, "src": [SRC_W, SRC_H], "dst": [DST, DST],
            "proposals": P, "classes": NC, "sharding": f"frames x{n_gpus}, no collective",
            "l2": "inputs larger than L2: %.0f MB of frames + logits read per step per GPU" % (batch * (FRAME_BYTES + LOGIT_BYTES) / 1e6)}


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return                                   # rank 0 alone runs the CPU arm
    from concurrent.futures import ThreadPoolExecutor
    frames, logits = make_inputs(args.batch, 0)
    threads = os.cpu_count() or 1
    cpu = CpuPath(threads)
    with ThreadPoolExecutor(threads) as pool:
        for _ in range(args.warmup):
            cpu.run_batch(frames, logits, pool)
        t = 0.0
        for _ in range(args.steps):
            t += cpu.run_batch(frames, logits, pool)
    fps = args.steps * args.batch / t
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(args.gpus, args.batch),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": cpu.kind,
                             "sample": cpu.describe(args.steps * args.batch) + f"; each step = one {args.batch}-frame batch"},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- #
# B200 arm                                                                      #
# --------------------------------------------------------------------------- #
def load_traffic() -> dict:
    """ncu dram bytes per launch of each kernel at this workload, written from a --set full capture
    (profiles/traffic.json); absent -> null."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def run_b200(args) -> None:
    import torch
    import torch.distributed as dist
    from trackiellm_b200 import _capi as capi
    from trackiellm_b200 import vision as V

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libtk_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        # frames shard with no data-path collective; gloo only carries the barrier and the max
        dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = torch.device("cuda", local)
    B, K, W = args.batch, args.steps, args.warmup

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    frames_np, logits_np = make_inputs(B, rank * B)
    h_frames = torch.from_numpy(frames_np).pin_memory()
    h_logits = torch.from_numpy(logits_np).pin_memory()
    d_frames = h_frames.to(dev)
    d_logits = h_logits.to(dev)

    pipe = V.Pipeline(device_id=local, max_batch=B, chunk_frames=args.chunk, width=SRC_W, height=SRC_H,
                      fmt=V.PIXEL_I420 if I420 else V.PIXEL_RGB8,
                      tensor_w=DST, tensor_h=DST, out_dtype=torch.float16, num_proposals=P, num_classes=NC,
                      act=V.ACT_RAW, conf_thr=CONF_THR, iou_thr=IOU_THR, max_detections=MAX_DET,
                      stage_host_logits=args.logits_mode)
    d_tensor, d_dets, d_counts = pipe.alloc_outputs(B)
    _, h_dets, h_counts = pipe.alloc_outputs(B, pinned=True)
    lib = capi.lib()
    lib.tk_b200_launch_count.restype = C.c_ulonglong
    sampler = ClockSampler(local)

    # ---------------- device-resident: `value` ----------------
    for _ in range(max(W, 3)):
        pipe.process_device(d_frames, d_logits, d_tensor, d_dets, d_counts)
    barrier()
    # The timed region is the call a user makes: K steps with nothing between the kernels.  The per-kernel
    # durations come from a second K-step pass right behind it, same inputs, with CUDA events recorded on the
    # launching stream between the kernels; the four records serialise kernel tail and next launch and cost
    # about 12 us of a 151 us step on B200 (tools/exp_pdl.py), which is why they are not inside `value`.
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = int(lib.tk_b200_launch_count())
    sampler.resume()
    e0.record()
    for _ in range(K):
        pipe.process_device(d_frames, d_logits, d_tensor, d_dets, d_counts)
    e1.record()
    barrier()
    launches = int(lib.tk_b200_launch_count()) - launches0
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    pipe.stage_timing(K)
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for _ in range(K):
        pipe.process_device(d_frames, d_logits, d_tensor, d_dets, d_counts)
    e3.record()
    barrier()
    sampler.pause()
    ms_total_staged = max_over_ranks(e2.elapsed_time(e3))
    stage_ms, calls = pipe.stage_times()
    pipe.stage_timing(0)
    dev_counts = d_counts.cpu().numpy().copy()
    dev_cand_counts = (logits_np[:, :, 4] > CONF_THR).sum(axis=1)      # rows that pass the threshold (read in full by decode)
    dev_dets = d_dets.cpu().numpy().copy()

    # ---------------- end to end through the host-buffer C-ABI call: `e2e` ----------------
    Ke = max(2, min(K, args.e2e_steps))
    for _ in range(2):
        pipe.process_host(h_frames, h_logits, h_dets, h_counts)
    barrier()
    sampler.resume()
    t0 = time.perf_counter()
    for _ in range(Ke):
        pipe.process_host(h_frames, h_logits, h_dets, h_counts)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    sampler.pause()
    barrier()
    # same, with the logits left on the device by a co-located detector (informative)
    t0 = time.perf_counter()
    for _ in range(Ke):
        pipe.process_host(h_frames, d_logits, h_dets, h_counts, logits_on_device=True)
    torch.cuda.synchronize()
    e2e_dl_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    # the host path must give what the device path gave
    hc = h_counts.numpy()
    same = bool(np.array_equal(hc, dev_counts)) and all(
        np.array_equal(h_dets.numpy()[i, :hc[i]], dev_dets[i, :hc[i]]) for i in range(B))
    if not same:
        raise SystemExit("bench.py: host-buffer path and device-resident path disagree")
    clocks = sampler.result()

    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "MEASURED_PEAKS.json hbm_gbs (burst copy)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
        names = ["preprocess_kernel", "decode_kernel", "nms_kernel"]
        alg = [ALG_PRE * B, ALG_POST * B, 0]
        per_launch_ms = [m / max(calls, 1) for m in stage_ms]
        traffic = {} if I420 else load_traffic()          # the ncu capture is of the default workload
        stages = []
        for nme, a, m in zip(names, alg, per_launch_ms):
            ach = a / (m * 1e-3) / 1e9 if m > 0 and a else None
            stages.append({"kernel": nme, "ms_per_launch": m, "algorithmic_bytes_per_launch": a,
                           "achieved": ach, "frac": (ach / peak) if ach else None, "traffic": traffic.get(nme)})
        dom = max(range(3), key=lambda i: per_launch_ms[i])
        d = stages[dom]
        roofline = {"bound": "hbm", "kernel": d["kernel"], "achieved": d["achieved"], "peak": peak, "unit": "GB/s",
                    "frac": d["frac"], "traffic": d["traffic"], "peak_source": peak_src,
                    "ms_per_launch": d["ms_per_launch"], "algorithmic_bytes_per_launch": d["algorithmic_bytes_per_launch"],
                    "share_of_step": d["ms_per_launch"] / max(sum(per_launch_ms), 1e-12),
                    "timed_by": "CUDA events on the launching stream between the kernels, in a second K-step pass right after "
                                "the timed region (same inputs); ms_per_step_with_stage_events is that pass's step time",
                    "ms_per_step_with_stage_events": ms_total_staged / K}
        frame_pcie, frames_in_place = frame_bytes_over_pcie()
        step_ms = ms_total / K
        step_ach = (ALG_PRE + ALG_POST) * B / (step_ms * 1e-3) / 1e9
        line = {"metric": METRIC, "value": world * B * K / (ms_total * 1e-3), "unit": "frames/s", "n_gpus": world,
                "steps": K, "warmup": max(W, 3), "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": ("i32 BT.601 + f32 lerp -> f16 tensor (pre-process); f32 / i32 (decode, NMS, bit-exact)" if I420 else
                          "f16 lerp -> f16 tensor (pre-process, within the 1e-2 fp16 bar); f32 / i32 (decode, NMS, bit-exact)"),
                "data": "synthetic", "config": config_dict(world, B),
                "roofline": roofline, "roofline_stages": stages,
                "roofline_step": {"algorithmic_bytes_per_frame": ALG_PRE + ALG_POST, "achieved": step_ach, "peak": peak,
                                  "unit": "GB/s", "frac": step_ach / peak},
                "e2e": {"value": world * B * Ke / e2e_s, "unit": "frames/s",
                        "h2d_bytes_per_step": B * ((0 if frames_in_place else FRAME_BYTES)
                                                   + (LOGIT_BYTES, P * 4, 0)[(1, 0, 2).index(args.logits_mode)]),
                        "host_bytes_read_in_place_per_step": (B * frame_pcie if frames_in_place else 0) + (
                            0 if args.logits_mode == 1 else int(np.sum(np.minimum(dev_cand_counts, 4096))) * (5 + NC) * 4
                            + (B * P * 64 if args.logits_mode == 2 else 0)),
                        "d2h_bytes_per_step": B * (MAX_DET * 28 + 4), "steps": Ke, "ms_per_step": 1e3 * e2e_s / Ke,
                        "api": "tk_b200_pipeline_process_host (pinned host frames + logits in, host detections + counts out; "
                               "the fp16 tensors stay on the device for the detector); "
                               + ("the pinned frames are read in place by the pre-process kernel's row ring (only the two source "
                                  "rows of each output row cross PCIe: %d of %d bytes per frame), " % (frame_pcie, FRAME_BYTES)
                                  if frames_in_place else "frames are copied H2D in chunks, ")
                               + {1: "logits are copied H2D in chunks",
                                  2: "the pinned logits are read in place over PCIe by decode_kernel (objectness word of every "
                                     "row, 64 B per request, + surviving rows), not copied",
                                  0: "the pinned logits stay in host memory: their objectness column (4 of every 340 bytes) is "
                                     "packed by 4 host threads while the frame copies run and uploaded as one dense array; "
                                     "decode_kernel reads only the surviving rows in place over PCIe"}[args.logits_mode]},
                "e2e_device_logits": {"value": world * B * Ke / e2e_dl_s, "unit": "frames/s",
                                      "h2d_bytes_per_step": 0 if frames_in_place else B * FRAME_BYTES,
                                      "host_bytes_read_in_place_per_step": B * frame_pcie if frames_in_place else 0,
                                      "d2h_bytes_per_step": B * (MAX_DET * 28 + 4),
                                      "api": "tk_b200_pipeline_process_host_frames (logits already in HBM)"},
                "gpu_launches": launches, "clocks": clocks,
                "detections_per_frame": float(np.mean(dev_counts))}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(frames_np, logits_np)
        print(json.dumps(line), flush=True)
    pipe.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="1080p-rgb", choices=["1080p-rgb", "4k-i420"],
                    help="1080p-rgb: BASELINE configs[1]+[2], the metric's configuration (default); 4k-i420: configs[3]")
    ap.add_argument("--batch", type=int, default=0, help="frames per GPU per step (default 64; 32 for 4k-i420)")
    ap.add_argument("--chunk", type=int, default=8, help="frames per in-flight chunk of the host-buffer call")
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--logits-mode", type=int, default=0, choices=[0, 1, 2],
                    help="e2e, host logits: 0 objectness column packed on the host + surviving rows in place (default), "
                         "1 whole tensor copied to HBM, 2 objectness and rows read in place over PCIe")
    args = ap.parse_args()
    set_workload(args.workload)
    if args.batch <= 0:
        args.batch = 32 if I420 else 64
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
