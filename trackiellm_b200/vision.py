"""Host-side Python mirror of the reference's vision/GPU interface for the hot path.

Every function here is a thin argument-marshalling layer over the C-ABI of
libtk_b200.so (include/tk_b200_vision.h); torch is used only for device memory
and streams.  Names follow the reference (`tk_kernels_preprocess_image`,
`tk_preprocessor_resize_and_normalize_to_chw`, ...) with the `tk_` prefix
dropped.  Nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _capi as capi
from ._capi import (ACT_RAW, ACT_SIGMOID, ACT_SOFTMAX, ARITH_EXACT, ARITH_FAST, ARITH_HALF, OUT_F16, OUT_F32,
                    PIXEL_BGR8, PIXEL_I420, PIXEL_RGB8, PIXEL_RGBA8, RESIZE_LETTERBOX, RESIZE_STRETCH)

IMAGENET_MEAN = (0.485, 0.456, 0.406)   # src/vision/tk_object_detector.c:248
IMAGENET_STD = (0.229, 0.224, 0.225)    # src/vision/tk_object_detector.c:249
COLOR_NAMES = ("red", "yellow", "green", "cyan", "blue", "magenta", "black", "white", "gray")

CAND_DTYPE = np.dtype([("index", "<i4"), ("x", "<i4"), ("y", "<i4"), ("w", "<i4"), ("h", "<i4"),
                       ("confidence", "<f4"), ("class_id", "<u4")])
CAND_WORDS = 7


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f3(v) -> capi.TkFloat3:
    return capi.TkFloat3(float(v[0]), float(v[1]), float(v[2]))


def _require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (there is no CPU path)")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")


# --------------------------------------------------------------------------- #
# (1) pre-processing                                                            #
# --------------------------------------------------------------------------- #
def preprocess_batch(frames: torch.Tensor, out_w: int, out_h: int, *, width: int | None = None,
                     height: int | None = None, fmt: int = PIXEL_RGB8, stride: int = 0,
                     out_dtype: torch.dtype = torch.float32, arith: int | None = None,
                     resize: int = RESIZE_STRETCH, pad_value: float = 114.0,
                     mean=IMAGENET_MEAN, std=IMAGENET_STD, out: torch.Tensor | None = None) -> torch.Tensor:
    """tk_b200_preprocess_batch.  frames: u8 CUDA tensor, [N,H,W,3] (RGB8), [N,H,W,4] (RGBA8) or
    [N, bytes] with width/height given (I420 or strided).  Returns [N,3,out_h,out_w]."""
    _require_cuda(frames, "frames")
    if frames.dtype != torch.uint8:
        raise ValueError("frames must be uint8")
    n = frames.shape[0]
    if frames.dim() == 4:
        height, width = frames.shape[1], frames.shape[2]
    elif width is None or height is None:
        raise ValueError("width/height required for flat frame buffers")
    pitch = frames.stride(0) * frames.element_size()
    f16 = out_dtype == torch.float16
    if arith is None:
        arith = ARITH_EXACT                            # the reference's arithmetic; HALF / FAST are opt-in
    if out is None:
        out = torch.empty((n, 3, out_h, out_w), dtype=out_dtype, device=frames.device)
    p = capi.PreprocessBatchParams(
        d_input=frames.data_ptr(), d_input_ptrs=None, input_frame_pitch_bytes=pitch, num_frames=n,
        input_width=width, input_height=height, input_stride_bytes=stride, input_format=fmt,
        d_output=out.data_ptr(), output_width=out_w, output_height=out_h,
        output_type=OUT_F16 if f16 else OUT_F32, resize_mode=resize, pad_value=pad_value, arith=arith,
        mean=_f3(mean), std_dev=_f3(std))
    with torch.cuda.device(frames.device):
        capi.check(capi.lib().tk_b200_preprocess_batch(C.byref(p), _stream()), "tk_b200_preprocess_batch")
    return out


def kernels_preprocess_image(frame: torch.Tensor, out_w: int, out_h: int, mean=IMAGENET_MEAN, std=IMAGENET_STD,
                             stride: int = 0) -> torch.Tensor:
    """tk_kernels_preprocess_image (reference signature, src/gpu/cuda/tk_cuda_kernels.h:51)."""
    _require_cuda(frame, "frame")
    h, w = frame.shape[0], frame.shape[1]
    out = torch.empty((3, out_h, out_w), dtype=torch.float32, device=frame.device)
    p = capi.PreprocessParams(frame.data_ptr(), w, h, stride or w * 3, out.data_ptr(), out_w, out_h, _f3(mean), _f3(std))
    with torch.cuda.device(frame.device):
        capi.check(capi.lib().tk_kernels_preprocess_image(C.byref(p), _stream()), "tk_kernels_preprocess_image")
    return out


def preprocessor_resize_and_normalize_to_chw(frame: np.ndarray, out_w: int, out_h: int,
                                             mean=IMAGENET_MEAN, std=IMAGENET_STD) -> np.ndarray:
    """tk_preprocessor_resize_and_normalize_to_chw (src/vision/tk_image_preprocessor.h:49-56):
    HOST frame in, HOST fp32 CHW tensor out; the GPU does the work."""
    frame = np.ascontiguousarray(frame, dtype=np.uint8)
    h, w = frame.shape[:2]
    vf = capi.VideoFrame(w, h, w * 3, PIXEL_RGB8, frame.ctypes.data)
    out = np.empty((3, out_h, out_w), dtype=np.float32)
    m = (C.c_float * 3)(*mean)
    s = (C.c_float * 3)(*std)
    capi.check(capi.lib().tk_preprocessor_resize_and_normalize_to_chw(
        C.byref(vf), out.ctypes.data_as(C.c_void_p), C.c_uint32(out_w), C.c_uint32(out_h), m, s),
        "tk_preprocessor_resize_and_normalize_to_chw")
    return out


# --------------------------------------------------------------------------- #
# (2)(3) decode + NMS                                                           #
# --------------------------------------------------------------------------- #
def decode_detections(logits: torch.Tensor, conf_thr: float, *, act: int = ACT_RAW, max_candidates: int = 4096,
                      candidates: torch.Tensor | None = None, counts: torch.Tensor | None = None):
    """tk_kernels_decode_detections.  logits [N,P,5+C] f32 CUDA ->
    (candidates int32 [N,max_candidates,7] (view as CAND_DTYPE on host), counts u32-as-int32 [N])."""
    _require_cuda(logits, "logits")
    n, p, l = logits.shape
    if candidates is None:
        candidates = torch.empty((n, max_candidates, CAND_WORDS), dtype=torch.int32, device=logits.device)
    if counts is None:
        counts = torch.empty((n,), dtype=torch.int32, device=logits.device)
    prm = capi.DecodeParams(logits.data_ptr(), n, p, l - 5, conf_thr, act, candidates.data_ptr(),
                            counts.data_ptr(), max_candidates)
    with torch.cuda.device(logits.device):
        capi.check(capi.lib().tk_kernels_decode_detections(C.byref(prm), _stream()), "tk_kernels_decode_detections")
    return candidates, counts


def nms_classwise(candidates: torch.Tensor, counts: torch.Tensor, iou_thr: float, max_detections: int, *,
                  class_agnostic: bool = False, detections: torch.Tensor | None = None,
                  num_detections: torch.Tensor | None = None, restore: tuple[int, int, int, int] | None = None):
    """tk_kernels_nms_classwise -> (detections int32 [N,max_detections,7], num_detections int32 [N])."""
    _require_cuda(candidates, "candidates")
    n, cap = candidates.shape[0], candidates.shape[1]
    if detections is None:
        detections = torch.empty((n, max_detections, CAND_WORDS), dtype=torch.int32, device=candidates.device)
    if num_detections is None:
        num_detections = torch.empty((n,), dtype=torch.int32, device=candidates.device)
    prm = capi.NmsParams(candidates.data_ptr(), counts.data_ptr(), n, cap, iou_thr, max_detections,
                         int(class_agnostic), detections.data_ptr(), num_detections.data_ptr(),
                         *(restore or (0, 0, 0, 0)))   # restore = (src_w, src_h, in_w, in_h)
    with torch.cuda.device(candidates.device):
        capi.check(capi.lib().tk_kernels_nms_classwise(C.byref(prm), _stream()), "tk_kernels_nms_classwise")
    return detections, num_detections


def nms_classwise_static(candidates: torch.Tensor, counts: torch.Tensor, iou_thr: float, detections: torch.Tensor,
                         num_detections: torch.Tensor, *, class_agnostic: bool = False,
                         restore: tuple[int, int, int, int] | None = None) -> None:
    """tk_b200_nms_classwise_static: tk_kernels_nms_classwise with its scratch in a static device array — the library
    allocates nothing, and neither does this wrapper: the caller passes the outputs (detections int32
    [N,max_detections,7], num_detections int32 [N]).  At most nms_static_max_candidates() candidates per frame; calls
    that overlap on the GPU race (they share the array)."""
    _require_cuda(candidates, "candidates")
    n, cap = candidates.shape[0], candidates.shape[1]
    prm = capi.NmsParams(candidates.data_ptr(), counts.data_ptr(), n, cap, iou_thr, detections.shape[1],
                         int(class_agnostic), detections.data_ptr(), num_detections.data_ptr(),
                         *(restore or (0, 0, 0, 0)))
    with torch.cuda.device(candidates.device):
        capi.check(capi.lib().tk_b200_nms_classwise_static(C.byref(prm), _stream()), "tk_b200_nms_classwise_static")


def nms_static_max_candidates() -> int:
    return int(capi.lib().tk_b200_nms_classwise_static_max_candidates())


def postprocess_fused(logits: torch.Tensor, conf_thr: float, iou_thr: float, max_detections: int, *, act: int = ACT_RAW,
                      max_candidates: int = 4096, class_agnostic: bool = False,
                      restore: tuple[int, int, int, int] | None = None, layout: int = 0):
    """tk_kernels_postprocess_fused: decode + class-wise NMS per frame in one kernel (layout 0: a cluster of 8 CTAs
    per frame; 1: persistent co-resident CTAs).
    -> (candidates, counts, detections, num_detections) as decode_detections / nms_classwise return them."""
    _require_cuda(logits, "logits")
    n, p, l = logits.shape
    dev = logits.device
    cands = torch.empty((n, max_candidates, CAND_WORDS), dtype=torch.int32, device=dev)
    counts = torch.empty((n,), dtype=torch.int32, device=dev)
    dets = torch.empty((n, max_detections, CAND_WORDS), dtype=torch.int32, device=dev)
    ndets = torch.empty((n,), dtype=torch.int32, device=dev)
    L = capi.lib()
    L.tk_b200_postprocess_fused_workspace_bytes.restype = C.c_size_t
    with torch.cuda.device(dev):
        need = int(L.tk_b200_postprocess_fused_workspace_bytes(C.c_uint32(max_candidates), C.c_uint32(n), C.c_int(layout)))
        ws = torch.empty((need,), dtype=torch.uint8, device=dev) if need else None
        prm = capi.PostParams(logits.data_ptr(), n, p, l - 5, conf_thr, act, cands.data_ptr(), counts.data_ptr(),
                              max_candidates, iou_thr, max_detections, int(class_agnostic), dets.data_ptr(),
                              ndets.data_ptr(), *(restore or (0, 0, 0, 0)), ws.data_ptr() if ws is not None else None, need, layout)
        capi.check(L.tk_kernels_postprocess_fused(C.byref(prm), _stream()), "tk_kernels_postprocess_fused")
        if ws is not None:
            ws.record_stream(torch.cuda.current_stream())
    return cands, counts, dets, ndets


def candidates_to_numpy(cands: torch.Tensor, counts: torch.Tensor) -> list[np.ndarray]:
    """Device candidate/detection lists -> per-frame structured host arrays."""
    c = cands.cpu().numpy()
    k = counts.cpu().numpy()
    cap = c.shape[1]
    return [np.ascontiguousarray(c[i, :min(int(k[i]), cap)]).view(CAND_DTYPE).reshape(-1) for i in range(c.shape[0])]


# --------------------------------------------------------------------------- #
# (4) softmax / argmax / dominant colour                                        #
# --------------------------------------------------------------------------- #
def kernels_softmax(x: torch.Tensor) -> torch.Tensor:
    """tk_kernels_softmax (src/gpu/cuda/tk_cuda_kernels.h:92)."""
    _require_cuda(x, "x")
    out = torch.empty_like(x)
    p = capi.SoftmaxParams(x.data_ptr(), out.data_ptr(), x.shape[0], x.shape[1])
    with torch.cuda.device(x.device):
        capi.check(capi.lib().tk_kernels_softmax(C.byref(p), _stream()), "tk_kernels_softmax")
    return out


def argmax_rows(x: torch.Tensor, want_softmax: bool = True):
    """tk_kernels_argmax_rows -> (softmax or None, argmax int32 [rows])."""
    _require_cuda(x, "x")
    sm = torch.empty_like(x) if want_softmax else None
    am = torch.empty((x.shape[0],), dtype=torch.int32, device=x.device)
    p = capi.ArgmaxParams(x.data_ptr(), sm.data_ptr() if sm is not None else None, am.data_ptr(), x.shape[0], x.shape[1])
    with torch.cuda.device(x.device):
        capi.check(capi.lib().tk_kernels_argmax_rows(C.byref(p), _stream()), "tk_kernels_argmax_rows")
    return sm, am


def dominant_color(frame: torch.Tensor, boxes: torch.Tensor, want_bins: bool = False):
    """tk_kernels_dominant_color.  frame [H,W,3] u8 CUDA, boxes int32 [M,4] (x,y,w,h) CUDA."""
    _require_cuda(frame, "frame")
    _require_cuda(boxes, "boxes")
    h, w = frame.shape[0], frame.shape[1]
    m = boxes.shape[0]
    ids = torch.empty((m,), dtype=torch.int32, device=frame.device)
    bins = torch.empty((m, 9), dtype=torch.int64, device=frame.device) if want_bins else None
    p = capi.ColorParams(frame.data_ptr(), w, h, w * 3, boxes.data_ptr(), m, ids.data_ptr(),
                         bins.data_ptr() if bins is not None else None)
    with torch.cuda.device(frame.device):
        capi.check(capi.lib().tk_kernels_dominant_color(C.byref(p), _stream()), "tk_kernels_dominant_color")
    return ids, bins


def door_state(frame: torch.Tensor, boxes: torch.Tensor, want_edges: bool = False):
    """tk_kernels_door_state.  Returns (states int32 [M] (1 = closed), edge counts int32 [M] or None)."""
    _require_cuda(frame, "frame")
    _require_cuda(boxes, "boxes")
    h, w = frame.shape[0], frame.shape[1]
    m = boxes.shape[0]
    st = torch.empty((m,), dtype=torch.int32, device=frame.device)
    ed = torch.empty((m,), dtype=torch.int32, device=frame.device) if want_edges else None
    p = capi.DoorParams(frame.data_ptr(), w, h, w * 3, boxes.data_ptr(), m, st.data_ptr(),
                        ed.data_ptr() if ed is not None else None)
    with torch.cuda.device(frame.device):
        capi.check(capi.lib().tk_kernels_door_state(C.byref(p), _stream()), "tk_kernels_door_state")
    return st, ed


def _host_frame(frame: np.ndarray) -> tuple[capi.VideoFrame, np.ndarray]:
    frame = np.ascontiguousarray(frame, dtype=np.uint8)
    h, w = frame.shape[:2]
    return capi.VideoFrame(w, h, w * 3, PIXEL_RGB8, frame.ctypes.data), frame


def _classify_host(fn_name: str, frame: np.ndarray, bbox) -> str:
    vf, keep = _host_frame(frame)
    r = capi.Rect(*[int(v) for v in bbox])
    s = C.c_void_p()
    capi.check(getattr(capi.lib(), fn_name)(C.byref(vf), C.byref(r), C.byref(s)), fn_name)
    name = C.string_at(s.value).decode()
    C.CDLL(None).free(s)
    return name


def classify_dominant_color(frame: np.ndarray, bbox) -> str:
    """tk_classify_dominant_color (src/vision/tk_attribute_classifier.h): HOST frame, returns the colour name."""
    return _classify_host("tk_classify_dominant_color", frame, bbox)


def classify_door_state(frame: np.ndarray, bbox) -> str:
    """tk_classify_door_state: HOST frame, returns "open" or "closed"."""
    return _classify_host("tk_classify_door_state", frame, bbox)


def kernels_postprocess_depth_map(raw: torch.Tensor, scale: float, shift: float) -> torch.Tensor:
    """tk_kernels_postprocess_depth_map (src/gpu/cuda/tk_cuda_kernels.h:64)."""
    _require_cuda(raw, "raw")
    out = torch.empty_like(raw)
    p = capi.DepthPostParams(raw.data_ptr(), raw.shape[1], raw.shape[0], out.data_ptr(), scale, shift)
    with torch.cuda.device(raw.device):
        capi.check(capi.lib().tk_kernels_postprocess_depth_map(C.byref(p), _stream()), "tk_kernels_postprocess_depth_map")
    return out


def kernels_depth_to_point_cloud(depth: torch.Tensor, fx: float, fy: float, cx: float, cy: float) -> torch.Tensor:
    """tk_kernels_depth_to_point_cloud (src/gpu/cuda/tk_cuda_kernels.h:78) -> [H,W,3] f32."""
    _require_cuda(depth, "depth")
    h, w = depth.shape
    out = torch.empty((h, w, 3), dtype=torch.float32, device=depth.device)
    p = capi.DepthPointsParams(depth.data_ptr(), w, h, out.data_ptr(), fx, fy, cx, cy)
    with torch.cuda.device(depth.device):
        capi.check(capi.lib().tk_kernels_depth_to_point_cloud(C.byref(p), _stream()), "tk_kernels_depth_to_point_cloud")
    return out


def inverse_depth_to_metric(inv: torch.Tensor, min_depth: float = 0.1, max_depth: float = 10.0,
                            out: torch.Tensor | None = None) -> torch.Tensor:
    """tk_kernels_inverse_depth_to_metric: [N,H,W] (or [H,W]) MiDaS inverse depth -> metric depth, each map
    normalised by its own min/max (src/vision/tk_depth_midas.c:473-499, defaults :46-47)."""
    _require_cuda(inv, "inv")
    n = 1 if inv.dim() == 2 else inv.shape[0]
    h, w = inv.shape[-2:]
    if out is None:
        out = torch.empty_like(inv)
    p = capi.DepthMetricParams(inv.data_ptr(), out.data_ptr(), n, w, h, min_depth, max_depth)
    with torch.cuda.device(inv.device):
        capi.check(capi.lib().tk_kernels_inverse_depth_to_metric(C.byref(p), _stream()), "tk_kernels_inverse_depth_to_metric")
    return out


def detection_distance(dets: torch.Tensor, counts: torch.Tensor | None, depth: torch.Tensor, frame_w: int,
                       frame_h: int) -> torch.Tensor:
    """tk_kernels_detection_distance: dets [N,max_det,7] i32 view of tk_b200_candidate_t, counts [N] (or None),
    depth [N,H,W] f32 -> [N,max_det] f32 distances, -1 where there is no estimate
    (src/vision/src/object_analysis.rs:227-279)."""
    _require_cuda(dets, "dets")
    _require_cuda(depth, "depth")
    n, max_det = dets.shape[0], dets.shape[1]
    h, w = depth.shape[-2:]
    out = torch.empty((n, max_det), dtype=torch.float32, device=dets.device)
    need = int(capi.lib().tk_b200_detection_distance_workspace_bytes(w, h))
    ws = torch.empty((need,), dtype=torch.uint8, device=dets.device) if need else None
    p = capi.DistanceParams(dets.data_ptr(), counts.data_ptr() if counts is not None else None, n, max_det,
                            depth.data_ptr(), w, h, frame_w, frame_h, out.data_ptr(),
                            ws.data_ptr() if ws is not None else None, need)
    with torch.cuda.device(dets.device):
        capi.check(capi.lib().tk_kernels_detection_distance(C.byref(p), _stream()), "tk_kernels_detection_distance")
        if ws is not None:
            ws.record_stream(torch.cuda.current_stream())
    return out


class CudaDispatcher:
    """tk_cuda_dispatch_* (src/gpu/cuda/tk_cuda_dispatch.h:72-171) as a context object."""

    def __init__(self, device_id: int = 0):
        self._h = C.c_void_p()
        cfg = capi.DispatcherConfig(device_id)
        capi.check(capi.lib().tk_cuda_dispatch_create(C.byref(self._h), C.byref(cfg)), "tk_cuda_dispatch_create")

    def close(self) -> None:
        if self._h:
            capi.lib().tk_cuda_dispatch_destroy(C.byref(self._h))

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def malloc(self, size: int) -> C.c_void_p:
        b = C.c_void_p()
        capi.check(capi.lib().tk_cuda_dispatch_malloc(self._h, C.byref(b), C.c_size_t(size)), "tk_cuda_dispatch_malloc")
        return b

    def free(self, buf: C.c_void_p) -> None:
        capi.lib().tk_cuda_dispatch_free(self._h, C.byref(buf))

    def upload_async(self, buf, host: np.ndarray) -> None:
        capi.check(capi.lib().tk_cuda_dispatch_upload_async(self._h, buf, host.ctypes.data_as(C.c_void_p),
                                                            C.c_size_t(host.nbytes)), "tk_cuda_dispatch_upload_async")

    def download_async(self, host: np.ndarray, buf) -> None:
        capi.check(capi.lib().tk_cuda_dispatch_download_async(self._h, host.ctypes.data_as(C.c_void_p), buf,
                                                              C.c_size_t(host.nbytes)), "tk_cuda_dispatch_download_async")

    def synchronize(self) -> None:
        capi.check(capi.lib().tk_cuda_dispatch_synchronize(self._h), "tk_cuda_dispatch_synchronize")

    def device_ptr(self, buf) -> int:
        return int(capi.lib().tk_b200_buffer_device_ptr(buf) or 0)

    def preprocess_image(self, d_in, w, h, d_out, out_w, out_h, mean=IMAGENET_MEAN, std=IMAGENET_STD, stride=0) -> None:
        p = capi.PreprocessParams(self.device_ptr(d_in), w, h, stride or w * 3, self.device_ptr(d_out), out_w, out_h,
                                  _f3(mean), _f3(std))
        capi.check(capi.lib().tk_cuda_dispatch_preprocess_image(self._h, C.byref(p)), "tk_cuda_dispatch_preprocess_image")

    def tensor_nms(self, d_boxes, d_scores, num_boxes: int, d_indices, max_detections: int, iou_thr: float,
                   score_thr: float) -> int:
        """tk_cuda_tensor_nms (src/gpu/extensions/cuda/tk_cuda_tensor_ops.h:513-522)."""
        tb = capi.TensorDescriptor(buffer=d_boxes, type=0, layout=2, dimensions=3)
        tb.shape[0], tb.shape[1], tb.shape[2] = 1, num_boxes, 4
        ts = capi.TensorDescriptor(buffer=d_scores, type=0, layout=2, dimensions=2)
        ts.shape[0], ts.shape[1] = 1, num_boxes
        n = C.c_uint32(0)
        capi.check(capi.lib().tk_cuda_tensor_nms(self._h, C.byref(tb), C.byref(ts), d_indices, C.byref(n),
                                                 C.c_uint32(max_detections), C.c_float(iou_thr), C.c_float(score_thr)),
                   "tk_cuda_tensor_nms")
        return int(n.value)


# --------------------------------------------------------------------------- #
# Batched pipeline (pre-process + decode + class-wise NMS)                      #
# --------------------------------------------------------------------------- #
class Pipeline:
    """tk_b200_pipeline_* — the batched form of what tk_object_detector_detect does per frame
    (src/vision/tk_object_detector.c:182-203), minus the detector network itself."""

    def __init__(self, *, device_id: int = 0, max_batch: int = 64, chunk_frames: int = 0,
                 width: int = 1920, height: int = 1080, stride: int = 0, fmt: int = PIXEL_RGB8,
                 tensor_w: int = 640, tensor_h: int = 640, out_dtype: torch.dtype = torch.float16,
                 resize: int = RESIZE_STRETCH, arith: int | None = None, pad_value: float = 114.0,
                 mean=IMAGENET_MEAN, std=IMAGENET_STD, num_proposals: int = 25200, num_classes: int = 80,
                 act: int = ACT_RAW, conf_thr: float = 0.5, iou_thr: float = 0.5, max_candidates: int = 4096,
                 max_detections: int = 500, class_agnostic: bool = False, restore_coordinates: bool = False,
                 stage_host_logits: int = 0, fuse_postprocess: int = 0):
        self.cfg = self._make_config(
            device_id=device_id, max_batch=max_batch, chunk_frames=chunk_frames, width=width, height=height, stride=stride,
            fmt=fmt, tensor_w=tensor_w, tensor_h=tensor_h, out_dtype=out_dtype, resize=resize, arith=arith,
            pad_value=pad_value, mean=mean, std=std, num_proposals=num_proposals, num_classes=num_classes, act=act,
            conf_thr=conf_thr, iou_thr=iou_thr, max_candidates=max_candidates, max_detections=max_detections,
            class_agnostic=class_agnostic, restore_coordinates=restore_coordinates, stage_host_logits=stage_host_logits,
            fuse_postprocess=fuse_postprocess)
        self.out_dtype = out_dtype
        self.device = torch.device("cuda", device_id)
        self._h = C.c_void_p()
        capi.check(capi.lib().tk_b200_pipeline_create(C.byref(self._h), C.byref(self.cfg)), "tk_b200_pipeline_create")
        L = capi.lib()
        L.tk_b200_pipeline_frame_bytes.restype = C.c_size_t
        L.tk_b200_pipeline_frame_bytes.argtypes = [C.c_void_p]
        L.tk_b200_pipeline_tensor_bytes.restype = C.c_size_t
        L.tk_b200_pipeline_tensor_bytes.argtypes = [C.c_void_p]
        self.frame_bytes = int(L.tk_b200_pipeline_frame_bytes(self._h))
        self.tensor_bytes = int(L.tk_b200_pipeline_tensor_bytes(self._h))
        self.max_detections = max_detections or 500

    def _make_config(self, *, device_id: int = 0, max_batch: int = 64, chunk_frames: int = 0,
                     width: int = 1920, height: int = 1080, stride: int = 0, fmt: int = PIXEL_RGB8,
                     tensor_w: int = 640, tensor_h: int = 640, out_dtype: torch.dtype = torch.float16,
                     resize: int = RESIZE_STRETCH, arith: int | None = None, pad_value: float = 114.0,
                     mean=IMAGENET_MEAN, std=IMAGENET_STD, num_proposals: int = 25200, num_classes: int = 80,
                     act: int = ACT_RAW, conf_thr: float = 0.5, iou_thr: float = 0.5, max_candidates: int = 4096,
                     max_detections: int = 500, class_agnostic: bool = False, restore_coordinates: bool = False,
                     stage_host_logits: int = 0, fuse_postprocess: int = 0) -> "capi.PipelineConfig":
        f16 = out_dtype == torch.float16
        if arith is None:
            arith = ARITH_EXACT                        # the reference's arithmetic; HALF / FAST are opt-in
        return capi.PipelineConfig(
            device_id=device_id, max_batch=max_batch, chunk_frames=chunk_frames, input_width=width,
            input_height=height, input_stride_bytes=stride, input_format=fmt, tensor_width=tensor_w,
            tensor_height=tensor_h, tensor_type=OUT_F16 if f16 else OUT_F32, resize_mode=resize, arith=arith,
            pad_value=pad_value, mean=_f3(mean), std_dev=_f3(std), num_proposals=num_proposals,
            num_classes=num_classes, activation=act, confidence_threshold=conf_thr, iou_threshold=iou_thr,
            max_candidates=max_candidates, max_detections=max_detections, class_agnostic=int(class_agnostic),
            restore_coordinates=int(restore_coordinates), stage_host_logits=int(stage_host_logits),
            fuse_postprocess=int(fuse_postprocess))

    def close(self) -> None:
        if self._h:
            capi.lib().tk_b200_pipeline_destroy(C.byref(self._h))

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def update_thresholds(self, conf_thr: float, iou_thr: float) -> None:
        capi.check(capi.lib().tk_b200_pipeline_update_thresholds(self._h, C.c_float(conf_thr), C.c_float(iou_thr)),
                   "tk_b200_pipeline_update_thresholds")

    def candidate_overflow(self) -> tuple[int, int]:
        """(frames of the last process_host call cut to max_candidates, largest count of rows over the threshold)."""
        a, b = C.c_uint32(0), C.c_uint32(0)
        capi.check(capi.lib().tk_b200_pipeline_candidate_overflow(self._h, C.byref(a), C.byref(b)),
                   "tk_b200_pipeline_candidate_overflow")
        return int(a.value), int(b.value)

    def stage_timing(self, max_calls: int) -> None:
        capi.check(capi.lib().tk_b200_pipeline_stage_timing(self._h, C.c_uint32(max_calls)), "tk_b200_pipeline_stage_timing")

    def stage_times(self) -> tuple[list[float], int]:
        """([pre ms, decode ms, nms ms] summed over the recorded calls, number of calls)."""
        ms = (C.c_double * 3)()
        n = C.c_uint32(0)
        capi.check(capi.lib().tk_b200_pipeline_stage_times(self._h, ms, C.byref(n)), "tk_b200_pipeline_stage_times")
        return [ms[0], ms[1], ms[2]], int(n.value)

    def alloc_outputs(self, n: int, pinned: bool = False):
        """(tensor, detections int32 [n,max_det,7], counts int32 [n]) on the device, or pinned host (no tensor)."""
        if pinned:
            return (None, torch.empty((n, self.max_detections, CAND_WORDS), dtype=torch.int32).pin_memory(),
                    torch.empty((n,), dtype=torch.int32).pin_memory())
        return (torch.empty((n, 3, self.cfg.tensor_height, self.cfg.tensor_width), dtype=self.out_dtype, device=self.device),
                torch.empty((n, self.max_detections, CAND_WORDS), dtype=torch.int32, device=self.device),
                torch.empty((n,), dtype=torch.int32, device=self.device))

    def process_device(self, frames: torch.Tensor | None, logits: torch.Tensor | None, tensor: torch.Tensor | None,
                       dets: torch.Tensor | None, counts: torch.Tensor | None) -> None:
        """tk_b200_pipeline_process_device on torch's current stream (asynchronous, 3 launches)."""
        n = frames.shape[0] if frames is not None else logits.shape[0]
        for t, name in ((frames, "frames"), (logits, "logits"), (tensor, "tensor"), (dets, "dets"), (counts, "counts")):
            if t is not None:
                _require_cuda(t, name)
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None   # noqa: E731
        with torch.cuda.device(self.device):
            capi.check(capi.lib().tk_b200_pipeline_process_device(self._h, ptr(frames), ptr(logits), C.c_uint32(n),
                                                                  ptr(tensor), ptr(dets), ptr(counts), _stream()),
                       "tk_b200_pipeline_process_device")

    def process_host(self, frames, logits, dets, counts, tensor=None, logits_on_device: bool = False) -> None:
        """tk_b200_pipeline_process_host[_frames]: HOST (ideally pinned) torch/numpy buffers in and out; blocks."""
        def hp(t):
            if t is None:
                return None
            return C.c_void_p(t.data_ptr() if isinstance(t, torch.Tensor) else t.ctypes.data)
        n = frames.shape[0] if frames is not None else logits.shape[0]
        fn = capi.lib().tk_b200_pipeline_process_host_frames if logits_on_device else capi.lib().tk_b200_pipeline_process_host
        capi.check(fn(self._h, hp(frames), hp(logits), C.c_uint32(n), hp(tensor), hp(dets), hp(counts)),
                   "tk_b200_pipeline_process_host")


class Shard:
    """tk_b200_shard_*: one process, one host thread + one pipeline per device, frames sharded by contiguous block
    (SURVEY §8e).  device_ids may repeat an id (two pipelines on one device).  Keyword arguments as Pipeline
    (max_batch = the whole job's batch)."""

    def __init__(self, device_ids, **pipeline_kwargs):
        # build the config through a throw-away Pipeline-like path without creating a pipeline
        probe = Pipeline.__new__(Pipeline)
        cfg = Pipeline._make_config(probe, **pipeline_kwargs)
        self.cfg = cfg
        self.max_detections = cfg.max_detections or 500
        self.out_dtype = pipeline_kwargs.get("out_dtype", torch.float16)
        ids = (C.c_int * len(device_ids))(*device_ids)
        self._h = C.c_void_p()
        capi.check(capi.lib().tk_b200_shard_create(C.byref(self._h), ids, C.c_uint32(len(device_ids)), C.byref(cfg)),
                   "tk_b200_shard_create")

    def close(self) -> None:
        if self._h:
            capi.lib().tk_b200_shard_destroy(C.byref(self._h))

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def update_thresholds(self, conf_thr: float, iou_thr: float) -> None:
        capi.check(capi.lib().tk_b200_shard_update_thresholds(self._h, C.c_float(conf_thr), C.c_float(iou_thr)),
                   "tk_b200_shard_update_thresholds")

    def process_host(self, frames, logits, dets, counts, tensor=None) -> None:
        def hp(t):
            if t is None:
                return None
            return C.c_void_p(t.data_ptr() if isinstance(t, torch.Tensor) else t.ctypes.data)
        n = frames.shape[0] if frames is not None else logits.shape[0]
        capi.check(capi.lib().tk_b200_shard_process_host(self._h, hp(frames), hp(logits), C.c_uint32(n), hp(tensor), hp(dets),
                                                         hp(counts)), "tk_b200_shard_process_host")

    def gather(self, dets, counts):
        """Final host gather -> (flat records [total,7] int32, offsets uint64[n+1])."""
        d = dets.numpy() if isinstance(dets, torch.Tensor) else dets
        c = (counts.numpy() if isinstance(counts, torch.Tensor) else counts).astype(np.uint32)
        n = d.shape[0]
        offs = np.zeros(n + 1, np.uint64)
        L = capi.lib()
        capi.check(L.tk_b200_shard_gather(None, c.ctypes.data_as(C.c_void_p), C.c_uint32(n), C.c_uint32(d.shape[1]), None,
                                          offs.ctypes.data_as(C.c_void_p)), "tk_b200_shard_gather")
        flat = np.zeros((int(offs[-1]), CAND_WORDS), np.int32)
        capi.check(L.tk_b200_shard_gather(np.ascontiguousarray(d).ctypes.data_as(C.c_void_p), c.ctypes.data_as(C.c_void_p), C.c_uint32(n),
                                          C.c_uint32(d.shape[1]), flat.ctypes.data_as(C.c_void_p), offs.ctypes.data_as(C.c_void_p)),
                   "tk_b200_shard_gather")
        return flat, offs


class ObjectDetector:
    """tk_object_detector_create / _detect / _update_thresholds / _destroy (src/vision/tk_object_detector.h:96-166)
    with the detector network bound as a device-side hook: `network(d_input_ptr, n, in_w, in_h, d_output_ptr,
    num_proposals, row_len, stream_ptr) -> int` enqueues the head tensor on the stream (0 = success)."""

    def __init__(self, *, device_id: int = 0, input_w: int = 640, input_h: int = 640, class_labels=None,
                 num_classes: int | None = None, conf_thr: float = 0.5, iou_thr: float = 0.5,
                 backend: int = capi.VISION_BACKEND_CUDA, model_path: bytes = b"model.onnx"):
        self._labels = None
        nlab = num_classes if num_classes is not None else (len(class_labels) if class_labels else 0)
        if class_labels:
            self._labels = (C.c_char_p * len(class_labels))(*[s.encode() for s in class_labels])
        self._path = C.create_string_buffer(model_path)       # opaque tk_path_t*: never dereferenced by the library
        cfg = capi.ObjectDetectorConfig(backend, device_id, C.cast(self._path, C.c_void_p), input_w, input_h,
                                        C.cast(self._labels, C.POINTER(C.c_char_p)) if self._labels else None,
                                        nlab, conf_thr, iou_thr)
        self._h = C.c_void_p()
        self._cb = None
        capi.check(capi.lib().tk_object_detector_create(C.byref(self._h), C.byref(cfg)), "tk_object_detector_create")

    def bind_network(self, network, num_proposals: int, act: int = ACT_RAW) -> None:
        def tramp(user, d_in, n, w, h, d_out, p, l, stream):
            try:
                return int(network(d_in, n, w, h, d_out, p, l, stream) or 0)
            except Exception:      # never unwind through C
                import traceback
                traceback.print_exc()
                return 1
        self._cb = capi.INFERENCE_FN(tramp)
        capi.check(capi.lib().tk_b200_object_detector_bind_inference(self._h, self._cb, None, C.c_uint32(num_proposals),
                                                                     C.c_int(act)), "tk_b200_object_detector_bind_inference")

    def update_thresholds(self, conf_thr: float, iou_thr: float) -> None:
        capi.lib().tk_object_detector_update_thresholds(self._h, conf_thr, iou_thr)

    @staticmethod
    def _records(res, cnt):
        return [(int(res[i].class_id), res[i].label.decode() if res[i].label else None, float(res[i].confidence),
                 (res[i].bbox.x, res[i].bbox.y, res[i].bbox.w, res[i].bbox.h)) for i in range(cnt)]

    def detect(self, frame: np.ndarray, fmt: int = PIXEL_RGB8, width: int | None = None, height: int | None = None,
               stride: int = 0):
        """tk_object_detector_detect on one HOST frame -> [(class_id, label, confidence, (x, y, w, h)), ...]."""
        frame = np.ascontiguousarray(frame, np.uint8)
        h, w = (height, width) if width else frame.shape[:2]
        vf = capi.VideoFrame(w, h, stride, fmt, frame.ctypes.data)
        res = C.POINTER(capi.DetectionResult)()
        cnt = C.c_size_t(0)
        capi.check(capi.lib().tk_object_detector_detect(self._h, C.byref(vf), C.byref(res), C.byref(cnt)),
                   "tk_object_detector_detect")
        out = self._records(res, cnt.value)
        capi.lib().tk_object_detector_free_results(C.byref(res))
        return out

    def detect_batch(self, frames: np.ndarray, fmt: int = PIXEL_RGB8):
        """tk_b200_object_detector_detect_batch on HOST frames [N,H,W,3|4]."""
        frames = np.ascontiguousarray(frames, np.uint8)
        n, h, w = frames.shape[:3]
        vfs = (capi.VideoFrame * n)(*[capi.VideoFrame(w, h, 0, fmt, frames[i].ctypes.data) for i in range(n)])
        res = (C.POINTER(capi.DetectionResult) * n)()
        cnt = (C.c_size_t * n)()
        capi.check(capi.lib().tk_b200_object_detector_detect_batch(self._h, vfs, C.c_uint32(n), res, cnt),
                   "tk_b200_object_detector_detect_batch")
        out = []
        for i in range(n):
            out.append(self._records(res[i], cnt[i]))
            p = res[i]
            capi.lib().tk_object_detector_free_results(C.byref(p))
        return out

    def close(self) -> None:
        if self._h:
            capi.lib().tk_object_detector_destroy(C.byref(self._h))

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def postprocess_detections(logits: np.ndarray, *, act: int = ACT_RAW, conf_thr: float = 0.5, iou_thr: float = 0.5,
                           max_detections: int = 500, class_labels: list[str] | None = None,
                           frame_wh=(0, 0), input_wh=(0, 0)):
    """tk_b200_postprocess_detections: HOST logits [P,5+C] -> list of (class_id, label, confidence, (x,y,w,h)),
    the reference's tk_detection_result_t records (src/vision/tk_object_detector.h:66-71)."""
    logits = np.ascontiguousarray(logits, dtype=np.float32)
    p, l = logits.shape
    labels = None
    nlab = 0
    if class_labels:
        nlab = len(class_labels)
        labels = (C.c_char_p * nlab)(*[s.encode() for s in class_labels])
    res = C.POINTER(capi.DetectionResult)()
    cnt = C.c_size_t(0)
    capi.check(capi.lib().tk_b200_postprocess_detections(
        logits.ctypes.data_as(C.c_void_p), C.c_uint32(p), C.c_uint32(l - 5), C.c_int(act), C.c_float(conf_thr),
        C.c_float(iou_thr), C.c_uint32(max_detections), labels, C.c_size_t(nlab), C.c_uint32(frame_wh[0]),
        C.c_uint32(frame_wh[1]), C.c_uint32(input_wh[0]), C.c_uint32(input_wh[1]), C.byref(res), C.byref(cnt)),
        "tk_b200_postprocess_detections")
    out = []
    for i in range(cnt.value):
        r = res[i]
        out.append((int(r.class_id), r.label.decode() if r.label else None, float(r.confidence),
                    (r.bbox.x, r.bbox.y, r.bbox.w, r.bbox.h)))
    if cnt.value:
        capi.lib().tk_object_detector_free_results(C.byref(res))
    return out
