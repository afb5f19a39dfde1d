"""ctypes binding of libtk_b200.so (the C-ABI declared in include/tk_b200_vision.h).

The library is the product; this module only loads it and mirrors its structs.
There is no fallback: if the shared object is missing, importing any compute
entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
# TK_B200_LIB: another build of the same library (the CHECKED build, tools/run_checked.sh); no other file is ever tried.
LIB_PATH = os.environ.get("TK_B200_LIB") or os.path.join(PKG_DIR, "libtk_b200.so")

TK_SUCCESS = 0
TK_ERROR_INVALID_ARGUMENT = 1001
TK_ERROR_NOT_IMPLEMENTED = 1003
TK_ERROR_BUFFER_TOO_SMALL = 1004
TK_ERROR_INVALID_DIMENSIONS = 5008

PIXEL_RGB8, PIXEL_RGBA8, PIXEL_I420, PIXEL_BGR8 = 0, 1, 2, 3
OUT_F32, OUT_F16 = 0, 1
RESIZE_STRETCH, RESIZE_LETTERBOX = 0, 1
ARITH_EXACT, ARITH_FAST, ARITH_HALF = 0, 1, 2
ACT_RAW, ACT_SIGMOID, ACT_SOFTMAX = 0, 1, 2


class TkFloat3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]


class VideoFrame(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("stride", C.c_uint32),
                ("format", C.c_int), ("data", C.c_void_p)]


class Rect(C.Structure):
    _fields_ = [("x", C.c_int), ("y", C.c_int), ("w", C.c_int), ("h", C.c_int)]


class DetectionResult(C.Structure):
    _fields_ = [("class_id", C.c_uint32), ("label", C.c_char_p), ("confidence", C.c_float), ("bbox", Rect)]


class PreprocessParams(C.Structure):       # tk_preprocess_params_t
    _fields_ = [("d_input_image", C.c_void_p), ("input_width", C.c_uint32), ("input_height", C.c_uint32),
                ("input_stride_bytes", C.c_uint32), ("d_output_tensor", C.c_void_p),
                ("output_width", C.c_uint32), ("output_height", C.c_uint32),
                ("mean", TkFloat3), ("std_dev", TkFloat3)]


class SoftmaxParams(C.Structure):          # tk_softmax_params_t
    _fields_ = [("d_input_tensor", C.c_void_p), ("d_output_tensor", C.c_void_p),
                ("num_rows", C.c_uint32), ("num_cols", C.c_uint32)]


class DepthPostParams(C.Structure):        # tk_postprocess_depth_params_t
    _fields_ = [("d_raw_depth_map", C.c_void_p), ("width", C.c_uint32), ("height", C.c_uint32),
                ("d_metric_depth_map", C.c_void_p), ("scale", C.c_float), ("shift", C.c_float)]


class DepthPointsParams(C.Structure):      # tk_depth_to_points_params_t
    _fields_ = [("d_metric_depth_map", C.c_void_p), ("width", C.c_uint32), ("height", C.c_uint32),
                ("d_point_cloud", C.c_void_p), ("fx", C.c_float), ("fy", C.c_float),
                ("cx", C.c_float), ("cy", C.c_float)]


class PreprocessBatchParams(C.Structure):  # tk_b200_preprocess_batch_params_t
    _fields_ = [("d_input", C.c_void_p), ("d_input_ptrs", C.c_void_p), ("input_frame_pitch_bytes", C.c_uint64),
                ("num_frames", C.c_uint32), ("input_width", C.c_uint32), ("input_height", C.c_uint32),
                ("input_stride_bytes", C.c_uint32), ("input_format", C.c_int), ("d_output", C.c_void_p),
                ("output_width", C.c_uint32), ("output_height", C.c_uint32), ("output_type", C.c_int),
                ("resize_mode", C.c_int), ("pad_value", C.c_float), ("arith", C.c_int),
                ("mean", TkFloat3), ("std_dev", TkFloat3)]


class DecodeParams(C.Structure):           # tk_b200_decode_params_t
    _fields_ = [("d_logits", C.c_void_p), ("num_frames", C.c_uint32), ("num_proposals", C.c_uint32),
                ("num_classes", C.c_uint32), ("confidence_threshold", C.c_float), ("activation", C.c_int),
                ("d_candidates", C.c_void_p), ("d_counts", C.c_void_p), ("max_candidates", C.c_uint32)]


class NmsParams(C.Structure):              # tk_b200_nms_params_t
    _fields_ = [("d_candidates", C.c_void_p), ("d_counts", C.c_void_p), ("num_frames", C.c_uint32),
                ("max_candidates", C.c_uint32), ("iou_threshold", C.c_float), ("max_detections", C.c_uint32),
                ("class_agnostic", C.c_int), ("d_detections", C.c_void_p), ("d_num_detections", C.c_void_p),
                ("restore_src_width", C.c_uint32), ("restore_src_height", C.c_uint32),
                ("restore_in_width", C.c_uint32), ("restore_in_height", C.c_uint32)]


class PipelineConfig(C.Structure):         # tk_b200_pipeline_config_t
    _fields_ = [("device_id", C.c_int), ("max_batch", C.c_uint32), ("chunk_frames", C.c_uint32),
                ("input_width", C.c_uint32), ("input_height", C.c_uint32), ("input_stride_bytes", C.c_uint32),
                ("input_format", C.c_int), ("tensor_width", C.c_uint32), ("tensor_height", C.c_uint32),
                ("tensor_type", C.c_int), ("resize_mode", C.c_int), ("arith", C.c_int), ("pad_value", C.c_float),
                ("mean", TkFloat3), ("std_dev", TkFloat3), ("num_proposals", C.c_uint32), ("num_classes", C.c_uint32),
                ("activation", C.c_int), ("confidence_threshold", C.c_float), ("iou_threshold", C.c_float),
                ("max_candidates", C.c_uint32), ("max_detections", C.c_uint32), ("class_agnostic", C.c_int),
                ("restore_coordinates", C.c_int), ("stage_host_logits", C.c_int), ("fuse_postprocess", C.c_int)]


class PostParams(C.Structure):             # tk_b200_post_params_t
    _fields_ = [("d_logits", C.c_void_p), ("num_frames", C.c_uint32), ("num_proposals", C.c_uint32),
                ("num_classes", C.c_uint32), ("confidence_threshold", C.c_float), ("activation", C.c_int),
                ("d_candidates", C.c_void_p), ("d_counts", C.c_void_p), ("max_candidates", C.c_uint32),
                ("iou_threshold", C.c_float), ("max_detections", C.c_uint32), ("class_agnostic", C.c_int),
                ("d_detections", C.c_void_p), ("d_num_detections", C.c_void_p),
                ("restore_src_width", C.c_uint32), ("restore_src_height", C.c_uint32),
                ("restore_in_width", C.c_uint32), ("restore_in_height", C.c_uint32),
                ("d_workspace", C.c_void_p), ("workspace_bytes", C.c_size_t), ("layout", C.c_int)]


class ArgmaxParams(C.Structure):           # tk_b200_argmax_params_t
    _fields_ = [("d_input", C.c_void_p), ("d_softmax", C.c_void_p), ("d_argmax", C.c_void_p),
                ("num_rows", C.c_uint32), ("num_cols", C.c_uint32)]


class ColorParams(C.Structure):            # tk_b200_color_params_t
    _fields_ = [("d_frame", C.c_void_p), ("width", C.c_uint32), ("height", C.c_uint32),
                ("stride_bytes", C.c_uint32), ("d_boxes", C.c_void_p), ("num_boxes", C.c_uint32),
                ("d_color_ids", C.c_void_p), ("d_bins", C.c_void_p)]


class DoorParams(C.Structure):             # tk_b200_door_params_t
    _fields_ = [("d_frame", C.c_void_p), ("width", C.c_uint32), ("height", C.c_uint32),
                ("stride_bytes", C.c_uint32), ("d_boxes", C.c_void_p), ("num_boxes", C.c_uint32),
                ("d_states", C.c_void_p), ("d_edge_pixels", C.c_void_p)]


class DepthMetricParams(C.Structure):      # tk_b200_depth_metric_params_t
    _fields_ = [("d_inverse_depth", C.c_void_p), ("d_metric_depth", C.c_void_p), ("num_maps", C.c_uint32),
                ("width", C.c_uint32), ("height", C.c_uint32), ("min_depth", C.c_float), ("max_depth", C.c_float)]


class DistanceParams(C.Structure):         # tk_b200_distance_params_t
    _fields_ = [("d_detections", C.c_void_p), ("d_num_detections", C.c_void_p), ("num_frames", C.c_uint32),
                ("max_detections", C.c_uint32), ("d_depth_maps", C.c_void_p), ("depth_width", C.c_uint32),
                ("depth_height", C.c_uint32), ("frame_width", C.c_uint32), ("frame_height", C.c_uint32),
                ("d_distances", C.c_void_p), ("d_workspace", C.c_void_p), ("workspace_bytes", C.c_size_t)]


class ObjectDetectorConfig(C.Structure):   # tk_object_detector_config_t (src/vision/tk_object_detector.h:42-58)
    _fields_ = [("backend", C.c_int), ("gpu_device_id", C.c_int), ("model_path", C.c_void_p),
                ("input_width", C.c_uint32), ("input_height", C.c_uint32), ("class_labels", C.POINTER(C.c_char_p)),
                ("class_count", C.c_size_t), ("confidence_threshold", C.c_float), ("iou_threshold", C.c_float)]


# tk_b200_inference_fn
INFERENCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p,
                           C.c_uint32, C.c_uint32, C.c_void_p)
VISION_BACKEND_CPU, VISION_BACKEND_CUDA, VISION_BACKEND_METAL, VISION_BACKEND_ROCM = 0, 1, 2, 3


class TensorDescriptor(C.Structure):       # tk_tensor_descriptor_t
    _fields_ = [("buffer", C.c_void_p), ("type", C.c_int), ("layout", C.c_int), ("dimensions", C.c_uint32),
                ("shape", C.c_uint32 * 8), ("stride", C.c_size_t * 8), ("data_size_bytes", C.c_size_t),
                ("batch_size", C.c_uint32)]


class DispatcherConfig(C.Structure):
    _fields_ = [("device_id", C.c_int)]


_lib = None


class TkError(RuntimeError):
    def __init__(self, code: int, what: str):
        self.code = code
        detail = ""
        try:
            detail = lib().tk_error_get_detail().decode()
            name = lib().tk_error_to_string(code).decode()
        except Exception:  # pragma: no cover
            name = "?"
        super().__init__(f"{what}: tk_error_code_t {code} ({name}) {detail}")


def lib() -> C.CDLL:
    """Load libtk_b200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(make -C trackiellm_b200/csrc). There is no CPU fallback.")
        _lib = C.CDLL(LIB_PATH)
        _lib.tk_error_get_detail.restype = C.c_char_p
        _lib.tk_error_to_string.restype = C.c_char_p
        _lib.tk_error_to_string.argtypes = [C.c_int]
        _lib.tk_b200_version.restype = C.c_char_p
        _lib.tk_b200_color_name.restype = C.c_char_p
        _lib.tk_b200_color_name.argtypes = [C.c_uint32]
        _lib.tk_b200_buffer_device_ptr.restype = C.c_void_p
        _lib.tk_b200_buffer_device_ptr.argtypes = [C.c_void_p]
        _lib.tk_b200_buffer_size.restype = C.c_size_t
        _lib.tk_b200_buffer_size.argtypes = [C.c_void_p]
        _lib.tk_b200_dispatch_stream.restype = C.c_void_p
        _lib.tk_b200_dispatch_stream.argtypes = [C.c_void_p]
        _lib.tk_b200_dispatch_device.argtypes = [C.c_void_p]
        _lib.tk_b200_detection_distance_workspace_bytes.restype = C.c_size_t
        _lib.tk_b200_detection_distance_workspace_bytes.argtypes = [C.c_uint32, C.c_uint32]
        _lib.tk_object_detector_update_thresholds.restype = None
        _lib.tk_object_detector_update_thresholds.argtypes = [C.c_void_p, C.c_float, C.c_float]
        _lib.tk_object_detector_destroy.restype = None
        _lib.tk_b200_nms_classwise_static_max_candidates.restype = C.c_uint32
        _lib.tk_b200_nms_classwise_static_max_candidates.argtypes = []
        _lib.tk_b200_host_free.argtypes = [C.c_void_p]
        _lib.tk_b200_host_free.restype = None
    return _lib


def check(code: int, what: str) -> None:
    if code != TK_SUCCESS:
        raise TkError(code, what)
