"""CPU tests of the drop-in boundary: libtk_b200.so loads, exports every symbol that
include/tk_b200_vision.h declares, the headers are valid C, and the ctypes mirrors
have the C layout.  No compute call is made (there is no GPU here)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from trackiellm_b200 import _capi as capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")


def declared_functions():
    names = set()
    for hdr in ("tk_b200_vision.h", "tk_b200_types.h"):
        src = open(os.path.join(INC, hdr)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        for m in re.finditer(r"\b(tk_[a-z0-9_]+)\s*\((?!\s*\*)", src):   # (?!..): not a function-pointer typedef
            names.add(m.group(1))
    return sorted(names)


def test_every_declared_symbol_is_exported():
    lib = capi.lib()
    names = declared_functions()
    assert len(names) >= 35
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/ but not exported: {missing}"


def test_reference_entry_points_present():
    lib = capi.lib()
    for n in ("tk_preprocessor_resize_and_normalize_to_chw", "tk_kernels_preprocess_image", "tk_kernels_softmax",
              "tk_kernels_postprocess_depth_map", "tk_kernels_depth_to_point_cloud", "tk_cuda_dispatch_create",
              "tk_cuda_dispatch_destroy", "tk_cuda_dispatch_malloc", "tk_cuda_dispatch_free",
              "tk_cuda_dispatch_upload_async", "tk_cuda_dispatch_download_async", "tk_cuda_dispatch_synchronize",
              "tk_cuda_dispatch_preprocess_image", "tk_cuda_dispatch_depth_to_point_cloud", "tk_cuda_tensor_nms",
              "tk_classify_dominant_color", "tk_classify_door_state", "tk_object_detector_free_results",
              "tk_object_detector_create", "tk_object_detector_destroy", "tk_object_detector_detect",
              "tk_object_detector_update_thresholds",
              "tk_error_to_string", "tk_error_set_detail", "tk_error_get_detail"):
        assert hasattr(lib, n), n


def test_library_carries_sm100a_only():
    out = subprocess.run(["cuobjdump", "-lelf", capi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_headers_are_plain_c_and_layouts_match(tmp_path):
    prog = tmp_path / "layout.c"
    structs = {
        "tk_video_frame_t": capi.VideoFrame, "tk_rect_t": capi.Rect, "tk_detection_result_t": capi.DetectionResult,
        "tk_preprocess_params_t": capi.PreprocessParams, "tk_softmax_params_t": capi.SoftmaxParams,
        "tk_postprocess_depth_params_t": capi.DepthPostParams, "tk_depth_to_points_params_t": capi.DepthPointsParams,
        "tk_b200_preprocess_batch_params_t": capi.PreprocessBatchParams, "tk_b200_decode_params_t": capi.DecodeParams,
        "tk_b200_nms_params_t": capi.NmsParams, "tk_b200_argmax_params_t": capi.ArgmaxParams,
        "tk_b200_color_params_t": capi.ColorParams, "tk_b200_door_params_t": capi.DoorParams,
        "tk_tensor_descriptor_t": capi.TensorDescriptor, "tk_cuda_dispatcher_config_t": capi.DispatcherConfig,
        "tk_b200_pipeline_config_t": capi.PipelineConfig,
        "tk_object_detector_config_t": capi.ObjectDetectorConfig,
        "tk_b200_depth_metric_params_t": capi.DepthMetricParams, "tk_b200_distance_params_t": capi.DistanceParams,
        "tk_b200_post_params_t": capi.PostParams,
    }
    body = "\n".join(f'    printf("{n} %zu\\n", sizeof({n}));' for n in structs)
    prog.write_text('#include <stdio.h>\n#include "tk_b200_vision.h"\nint main(void) {\n' + body +
                    '\n    printf("cand %zu\\n", sizeof(tk_b200_candidate_t));\n'
                    '    printf("codes %d %d %d\\n", (int)TK_ERROR_INVALID_ARGUMENT, (int)TK_ERROR_GPU_ERROR, (int)TK_ERROR_INVALID_DIMENSIONS);\n'
                    "    return 0;\n}\n")
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", "-I", INC, str(prog), "-o", str(exe)], check=True)
    out = dict(line.split(" ", 1) for line in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for n, ct in structs.items():
        assert int(out[n]) == C.sizeof(ct), (n, out[n], C.sizeof(ct))
    assert int(out["cand"]) == 28
    assert out["codes"].split() == [str(capi.TK_ERROR_INVALID_ARGUMENT), "5000", str(capi.TK_ERROR_INVALID_DIMENSIONS)]


def test_argument_validation_needs_no_gpu():
    lib = capi.lib()
    # NULL params are rejected before any CUDA call (reference convention, tk_cuda_kernels.cu:258-260)
    assert lib.tk_kernels_preprocess_image(None, None) == capi.TK_ERROR_INVALID_ARGUMENT
    assert lib.tk_kernels_softmax(None, None) == capi.TK_ERROR_INVALID_ARGUMENT
    assert lib.tk_b200_preprocess_batch(None, None) == capi.TK_ERROR_INVALID_ARGUMENT
    assert lib.tk_kernels_decode_detections(None, None) == capi.TK_ERROR_INVALID_ARGUMENT
    assert lib.tk_kernels_nms_classwise(None, None) == capi.TK_ERROR_INVALID_ARGUMENT
    assert lib.tk_b200_pipeline_create(None, None) == capi.TK_ERROR_INVALID_ARGUMENT
    assert lib.tk_b200_pipeline_process_host(None, None, None, 0, None, None, None) == capi.TK_ERROR_INVALID_ARGUMENT
    assert lib.tk_preprocessor_resize_and_normalize_to_chw(None, None, 0, 0, None, None) == capi.TK_ERROR_INVALID_ARGUMENT
    assert lib.tk_error_to_string(capi.TK_ERROR_INVALID_DIMENSIONS) == b"invalid dimensions"
    assert lib.tk_b200_color_name(0) == b"red" and lib.tk_b200_color_name(8) == b"gray"


def test_shard_block_and_gather_are_host_only():
    """tk_b200_shard_block tiles [0, n) with blocks that differ by at most one frame; tk_b200_shard_gather is the final
    host gather (prefix sum of the counts + memcpy) — neither needs a device."""
    import numpy as np
    lib = capi.lib()
    for n in (0, 1, 7, 64, 1000, 1023):
        for G in (1, 2, 3, 8):
            lo, cnt = C.c_uint32(), C.c_uint32()
            nxt, sizes = 0, []
            for g in range(G):
                lib.tk_b200_shard_block(C.c_uint32(n), C.c_uint32(g), C.c_uint32(G), C.byref(lo), C.byref(cnt))
                assert lo.value == nxt
                nxt += cnt.value; sizes.append(cnt.value)
            assert nxt == n and max(sizes) - min(sizes) <= 1
    rng = np.random.default_rng(0)
    n, md = 9, 5
    dets = rng.integers(0, 1 << 30, (n, md, 7), dtype=np.int32)
    counts = np.array([0, 5, 3, 9, 1, 0, 5, 2, 4], np.uint32)          # 9 > md: clipped to md
    offs = np.zeros(n + 1, np.uint64)
    assert lib.tk_b200_shard_gather(None, counts.ctypes.data_as(C.c_void_p), n, md, None, offs.ctypes.data_as(C.c_void_p)) == 0
    k = np.minimum(counts, md)
    assert np.array_equal(offs, np.concatenate([[0], np.cumsum(k)]).astype(np.uint64))
    flat = np.zeros((int(offs[-1]), 7), np.int32)
    assert lib.tk_b200_shard_gather(dets.ctypes.data_as(C.c_void_p), counts.ctypes.data_as(C.c_void_p), n, md,
                                    flat.ctypes.data_as(C.c_void_p), offs.ctypes.data_as(C.c_void_p)) == 0
    for f in range(n):
        assert np.array_equal(flat[int(offs[f]):int(offs[f + 1])], dets[f, :k[f]])
    assert lib.tk_b200_shard_create(None, None, 0, None) == capi.TK_ERROR_INVALID_ARGUMENT


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = capi.lib()
    assert lib.tk_b200_device_count() == 0
    # a fully valid call must fail loudly with a GPU error code, not compute on the host
    import numpy as np
    fr = np.zeros((8, 8, 3), np.uint8)
    vf = capi.VideoFrame(8, 8, 24, 0, fr.ctypes.data)
    out = np.full((3, 4, 4), 7.0, np.float32)
    m = (C.c_float * 3)(0.5, 0.5, 0.5)
    rc = lib.tk_preprocessor_resize_and_normalize_to_chw(C.byref(vf), out.ctypes.data_as(C.c_void_p), 4, 4, m, m)
    assert 5000 <= rc < 6000 and np.all(out == 7.0)


def test_product_does_not_reference_oracle():
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "trackiellm_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"(from|import)\s+oracle|oracle/|tk_oracle|libtk_oracle", txt):
                    bad.append(f)
    assert not bad, bad
