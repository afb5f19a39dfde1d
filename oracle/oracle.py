"""ctypes front-end of the CPU oracle.  TEST INFRASTRUCTURE ONLY.

Loads oracle/libtk_oracle.so (the C restatement, tk_oracle.c) and, when
present, oracle/_ref/libtkref*.so (the reference's OWN C files compiled in
place by oracle/Makefile).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ACT_RAW, ACT_SIGMOID, ACT_SOFTMAX = 0, 1, 2
NORM_AVX, NORM_SCALAR = 0, 1
RESIZE_STRETCH, RESIZE_LETTERBOX = 0, 1
IMAGENET_MEAN = (0.485, 0.456, 0.406)   # src/vision/tk_object_detector.c:248
IMAGENET_STD = (0.229, 0.224, 0.225)    # src/vision/tk_object_detector.c:249
COLOR_NAMES = ("red", "yellow", "green", "cyan", "blue", "magenta", "black", "white", "gray")

CAND_DTYPE = np.dtype([("index", "<i4"), ("x", "<i4"), ("y", "<i4"), ("w", "<i4"), ("h", "<i4"),
                       ("confidence", "<f4"), ("class_id", "<u4")])
assert CAND_DTYPE.itemsize == 28


def build(force: bool = False) -> None:
    """Compile the oracle (and oracle/_ref when /root/reference is present)."""
    if force or not os.path.exists(os.path.join(HERE, "libtk_oracle.so")) \
            or os.path.exists("/root/reference/src/vision/tk_image_preprocessor.c"):
        subprocess.run(["make", "-s", "-C", HERE, "all"], check=True)


_lib = None
_ref = None
_ref_sm = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "libtk_oracle.so")
        if not os.path.exists(path):
            build()
        _lib = C.CDLL(path)
        _lib.oracle_decode.restype = C.c_uint32
        _lib.oracle_nms_classwise.restype = C.c_uint32
        _lib.oracle_iou.restype = C.c_float
        _lib.oracle_f32_to_f16.restype = C.c_uint16
        _lib.oracle_det_expf.restype = C.c_float
        _lib.oracle_det_expf.argtypes = [C.c_float]
        _lib.oracle_det_sigmoidf.restype = C.c_float
        _lib.oracle_det_sigmoidf.argtypes = [C.c_float]
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


def preprocess_rgb8(frame: np.ndarray, dst_w: int, dst_h: int, mean=IMAGENET_MEAN, std=IMAGENET_STD,
                    norm=NORM_AVX, resize=RESIZE_STRETCH, pad_value: float = 114.0,
                    stride: int | None = None, width: int | None = None) -> np.ndarray:
    """frame: (H, W, 3) u8 (or (H, stride) u8 with width given). Returns (3, dst_h, dst_w) f32."""
    frame = np.ascontiguousarray(frame, dtype=np.uint8)
    if frame.ndim == 3:
        h, w = frame.shape[:2]
        stride = w * 3
    else:
        h, w = frame.shape[0], int(width)
        stride = frame.shape[1]
    out = np.empty((3, dst_h, dst_w), dtype=np.float32)
    rc = lib().oracle_preprocess_rgb8(_p(frame), C.c_uint32(w), C.c_uint32(h), C.c_uint32(stride), _p(out),
                                      C.c_uint32(dst_w), C.c_uint32(dst_h), _f3(mean), _f3(std),
                                      C.c_int(norm), C.c_int(resize), C.c_float(pad_value))
    if rc != 0:
        raise ValueError("oracle_preprocess_rgb8: invalid arguments")
    return out


def i420_to_rgb8(i420: np.ndarray, w: int, h: int) -> np.ndarray:
    i420 = np.ascontiguousarray(i420, dtype=np.uint8).reshape(-1)
    assert i420.size >= w * h + 2 * (w // 2) * (h // 2)
    rgb = np.empty((h, w, 3), dtype=np.uint8)
    lib().oracle_i420_to_rgb8(_p(i420), C.c_uint32(w), C.c_uint32(h), _p(rgb))
    return rgb


def f32_to_f16(a: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.float32)
    out = np.empty(a.shape, dtype=np.uint16)
    lib().oracle_f32_to_f16_array(_p(a), _p(out), C.c_size_t(a.size))
    return out.view(np.float16)


def decode(logits: np.ndarray, conf_thr: float, act: int = ACT_RAW, cap: int | None = None):
    """logits (P, 5+C) f32 -> (candidates[min(total,cap)], total)."""
    logits = np.ascontiguousarray(logits, dtype=np.float32)
    p, l = logits.shape
    cap = p if cap is None else int(cap)
    out = np.zeros(max(cap, 1), dtype=CAND_DTYPE)
    total = lib().oracle_decode(_p(logits), C.c_uint32(p), C.c_uint32(l - 5), C.c_float(conf_thr),
                                C.c_int(act), _p(out), C.c_uint32(cap))
    return out[:min(total, cap)].copy(), int(total)


def iou(a, b) -> float:
    aa = np.array([a], dtype=CAND_DTYPE)
    bb = np.array([b], dtype=CAND_DTYPE)
    return float(lib().oracle_iou(_p(aa), _p(bb)))


def nms_classwise(cands: np.ndarray, iou_thr: float, max_det: int) -> np.ndarray:
    cands = np.ascontiguousarray(cands, dtype=CAND_DTYPE)
    kept = np.zeros(max(max_det, 1), dtype=CAND_DTYPE)
    n = lib().oracle_nms_classwise(_p(cands), C.c_uint32(len(cands)), C.c_float(iou_thr),
                                   C.c_uint32(max_det), _p(kept))
    return kept[:n].copy()


def restore_coords(dets: np.ndarray, src_w, src_h, in_w, in_h) -> np.ndarray:
    d = np.ascontiguousarray(dets, dtype=CAND_DTYPE).copy()
    lib().oracle_restore_coords(_p(d), C.c_uint32(len(d)), C.c_uint32(src_w), C.c_uint32(src_h),
                                C.c_uint32(in_w), C.c_uint32(in_h))
    return d


def softmax_rows(x: np.ndarray, det_exp: bool = False) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty_like(x)
    lib().oracle_softmax_rows(_p(x), _p(out), C.c_uint32(x.shape[0]), C.c_uint32(x.shape[1]), C.c_int(int(det_exp)))
    return out


def argmax_rows(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.shape[0], dtype=np.uint32)
    lib().oracle_argmax_rows(_p(x), _p(out), C.c_uint32(x.shape[0]), C.c_uint32(x.shape[1]))
    return out


def dominant_color(frame: np.ndarray, bbox) -> tuple[int, np.ndarray]:
    frame = np.ascontiguousarray(frame, dtype=np.uint8)
    h, w = frame.shape[:2]
    bins = np.zeros(9, dtype=np.int64)
    k = lib().oracle_dominant_color(_p(frame), C.c_uint32(w), C.c_uint32(h), C.c_uint32(w * 3),
                                    C.c_int(bbox[0]), C.c_int(bbox[1]), C.c_int(bbox[2]), C.c_int(bbox[3]), _p(bins))
    return int(k), bins


def door_state(frame: np.ndarray, bbox) -> tuple[int, int]:
    frame = np.ascontiguousarray(frame, dtype=np.uint8)
    h, w = frame.shape[:2]
    cnt = C.c_int32(0)
    k = lib().oracle_door_state(_p(frame), C.c_uint32(w), C.c_uint32(h), C.c_uint32(w * 3),
                                C.c_int(bbox[0]), C.c_int(bbox[1]), C.c_int(bbox[2]), C.c_int(bbox[3]), C.byref(cnt))
    return int(k), int(cnt.value)


def inverse_depth_to_metric(inv: np.ndarray, min_depth: float = 0.1, max_depth: float = 10.0) -> np.ndarray:
    """One map; src/vision/tk_depth_midas.c:473-499 with its defaults :46-47."""
    inv = np.ascontiguousarray(inv, np.float32)
    out = np.empty_like(inv)
    lib().oracle_inverse_depth_to_metric(_p(inv), _p(out), C.c_size_t(inv.size), C.c_float(min_depth), C.c_float(max_depth))
    return out


def detection_distance(bbox, depth: np.ndarray, frame_w: int, frame_h: int) -> float:
    """src/vision/src/object_analysis.rs:227-279 for one bbox (x, y, w, h) and one depth map [H][W]."""
    depth = np.ascontiguousarray(depth, np.float32)
    scratch = np.empty(depth.size, np.float32)
    fn = lib().oracle_detection_distance
    fn.restype = C.c_float
    x, y, w, h = (int(v) for v in bbox)
    return float(fn(C.c_int(x), C.c_int(y), C.c_int(w), C.c_int(h), _p(depth), C.c_uint32(depth.shape[1]),
                    C.c_uint32(depth.shape[0]), C.c_uint32(frame_w), C.c_uint32(frame_h), _p(scratch)))


def det_expf(x: float) -> float:
    return float(lib().oracle_det_expf(C.c_float(x)))


# --------------------------------------------------------------------------- #
# The reference's own compiled code (oracle/_ref), when built.                 #
# --------------------------------------------------------------------------- #
class _VideoFrame(C.Structure):      # src/cortex/tk_cortex_main.h:98-104
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("stride", C.c_uint32),
                ("format", C.c_int), ("data", C.c_void_p)]


class _Rect(C.Structure):            # src/vision/tk_vision_pipeline.h:158-163
    _fields_ = [("x", C.c_int), ("y", C.c_int), ("w", C.c_int), ("h", C.c_int)]


def have_ref() -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libtkref.so"))


def ref_lib() -> C.CDLL:
    global _ref
    if _ref is None:
        _ref = C.CDLL(os.path.join(HERE, "_ref", "libtkref.so"))
    return _ref


def ref_preprocess(frame: np.ndarray, dst_w: int, dst_h: int, mean=IMAGENET_MEAN, std=IMAGENET_STD) -> np.ndarray:
    """The reference's tk_preprocessor_resize_and_normalize_to_chw itself."""
    frame = np.ascontiguousarray(frame, dtype=np.uint8)
    h, w = frame.shape[:2]
    vf = _VideoFrame(w, h, w * 3, 0, frame.ctypes.data)
    out = np.empty((3, dst_h, dst_w), dtype=np.float32)
    rc = ref_lib().tk_preprocessor_resize_and_normalize_to_chw(C.byref(vf), _p(out), C.c_uint32(dst_w),
                                                               C.c_uint32(dst_h), _f3(mean), _f3(std))
    if rc != 0:
        raise RuntimeError(f"reference preprocessor returned {rc}")
    return out


def _ref_string_call(fn, frame: np.ndarray, bbox) -> str:
    frame = np.ascontiguousarray(frame, dtype=np.uint8)
    h, w = frame.shape[:2]
    vf = _VideoFrame(w, h, w * 3, 0, frame.ctypes.data)
    r = _Rect(*[int(v) for v in bbox])
    s = C.c_void_p()
    rc = fn(C.byref(vf), C.byref(r), C.byref(s))
    if rc != 0:
        raise RuntimeError(f"reference classifier returned {rc}")
    name = C.string_at(s.value).decode()
    C.CDLL(None).free(s)
    return name


def ref_dominant_color(frame, bbox) -> str:
    return _ref_string_call(ref_lib().tk_classify_dominant_color, frame, bbox)


def ref_door_state(frame, bbox) -> str:
    return _ref_string_call(ref_lib().tk_classify_door_state, frame, bbox)


def have_ref_softmax() -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libtkref_softmax.so"))


def ref_softmax(x: np.ndarray) -> np.ndarray:
    """The reference's softmax_cpu (tests/tk_gpu_softmax_test.cpp:24-51) itself."""
    global _ref_sm
    if _ref_sm is None:
        _ref_sm = C.CDLL(os.path.join(HERE, "_ref", "libtkref_softmax.so"), mode=os.RTLD_LAZY)
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty_like(x)
    _ref_sm.ref_softmax_cpu(_p(x), _p(out), C.c_int(x.shape[0]), C.c_int(x.shape[1]))
    return out
