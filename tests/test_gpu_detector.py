"""GPU parity: the reference's object-detector entry points (src/vision/tk_object_detector.h:96-166) with
a FAKE detector network bound as the device-side hook (ONNX Runtime and the model file are not part
of the reference tree).  The fake checks the tensor it is handed against the oracle pre-processor
and answers with a seeded synthetic head tensor; detect() must then return exactly what the oracle
decode + NMS + coordinate restore give."""
import ctypes as C

import numpy as np
import pytest
import torch

import synth
from oracle import oracle as O
from trackiellm_b200 import _capi as capi
from trackiellm_b200 import vision as V

pytestmark = pytest.mark.gpu

LABELS = [f"class{i}" for i in range(80)]


class FakeNetwork:
    """Enqueues a device-to-device copy of prepared head tensors on the stream it is given (driver API:
    the hook gets raw pointers and a raw stream, like an I/O-bound ORT session would)."""

    def __init__(self, logits: np.ndarray):
        self.cu = C.CDLL("libcuda.so.1")
        self.d_logits = torch.from_numpy(np.ascontiguousarray(logits, np.float32)).cuda()
        self.calls = []
        self.seen_input = None
        self.fail = False

    def __call__(self, d_in, n, w, h, d_out, p, l, stream):
        if self.fail:
            return 7
        self.calls.append((n, w, h, p, l))
        nbytes = n * p * l * 4
        assert nbytes <= self.d_logits.numel() * 4
        rc = self.cu.cuMemcpyDtoDAsync_v2(C.c_uint64(d_out), C.c_uint64(self.d_logits.data_ptr()), C.c_size_t(nbytes),
                                          C.c_void_p(stream))
        # look at the tensor the network was handed (synchronising here is test-only)
        host = np.empty((n, 3, h, w), np.float32)
        self.cu.cuStreamSynchronize(C.c_void_p(stream))
        self.cu.cuMemcpyDtoH_v2(host.ctypes.data_as(C.c_void_p), C.c_uint64(d_in), C.c_size_t(host.nbytes))
        self.seen_input = host
        return int(rc)


def oracle_records(logits, fw, fh, conf=0.5, iou=0.5):
    c, _ = O.decode(logits, conf)
    kept = O.restore_coords(O.nms_classwise(c, iou, 500), fw, fh, 640, 640)
    return [(int(k["class_id"]), LABELS[int(k["class_id"])], fbits(k["confidence"]),
             (int(k["x"]), int(k["y"]), int(k["w"]), int(k["h"]))) for k in kept]


def fbits(x) -> int:
    return int(np.float32(x).view(np.uint32))


def norm(records):
    """confidence as its bit pattern (the adversarial tensor carries NaN scores; NaN != NaN)."""
    return [(c, l, fbits(s), b) for c, l, s, b in records]


def test_detect_single_frame_cfg0():
    """BASELINE config 0: one 640x480 RGB frame -> 640x640 fp32 NCHW -> 25200x85 head -> decode + NMS
    (tests/tk_vision_full_test.c:30-32,67: zero frame, thresholds 0.5 / 0.5)."""
    logits = synth.logits_tensor(0, adversarial=True)
    net = FakeNetwork(logits[None])
    for kind in ("zeros", "random"):
        frame = synth.rgb_frame(640, 480, 0, kind)
        with V.ObjectDetector(class_labels=LABELS, conf_thr=0.5, iou_thr=0.5) as det:
            det.bind_network(net, synth.NUM_PROPOSALS)
            got = det.detect(frame)
        assert net.calls[-1] == (1, 640, 640, synth.NUM_PROPOSALS, 85)
        assert np.array_equal(net.seen_input[0].view(np.uint32), O.preprocess_rgb8(frame, 640, 640).view(np.uint32))
        want = oracle_records(logits, 640, 480)
        assert len(got) == len(want) > 10
        assert norm(got) == want


def test_detect_batch_thresholds_and_formats():
    n = 5
    logits = np.stack([synth.logits_tensor(10 + i) for i in range(n)])
    frames = np.stack([synth.rgb_frame(1920, 1080, 20 + i) for i in range(n)])
    net = FakeNetwork(logits)
    with V.ObjectDetector(class_labels=LABELS, conf_thr=0.5, iou_thr=0.5) as det:
        det.bind_network(net, synth.NUM_PROPOSALS)
        got = det.detect_batch(frames)
        for i in range(n):
            assert np.array_equal(net.seen_input[i].view(np.uint32), O.preprocess_rgb8(frames[i], 640, 640).view(np.uint32))
            assert norm(got[i]) == oracle_records(logits[i], 1920, 1080), i
        # tk_object_detector_update_thresholds (src/vision/tk_object_detector.c:212-219)
        det.update_thresholds(0.7, 0.3)
        got = det.detect_batch(frames[:2])
        for i in range(2):
            assert norm(got[i]) == oracle_records(logits[i], 1920, 1080, 0.7, 0.3), i
        # BGR frames: the network sees the same tensor
        got = det.detect_batch(np.ascontiguousarray(frames[:1, :, :, ::-1]), fmt=V.PIXEL_BGR8)
        assert np.array_equal(net.seen_input[0].view(np.uint32), O.preprocess_rgb8(frames[0], 640, 640).view(np.uint32))
        # a frame with nothing above the threshold gives an empty, NULL result
        det.update_thresholds(2.0, 0.5)
        assert det.detect(frames[0]) == []


def test_detect_error_behaviour():
    lib = capi.lib()
    # reference argument checks: src/vision/tk_object_detector.c:84,189-191
    h = C.c_void_p()
    assert lib.tk_object_detector_create(C.byref(h), None) == capi.TK_ERROR_INVALID_ARGUMENT
    cfg = capi.ObjectDetectorConfig(capi.VISION_BACKEND_CUDA, 0, None, 640, 640, None, 80, 0.5, 0.5)
    assert lib.tk_object_detector_create(C.byref(h), C.byref(cfg)) == capi.TK_ERROR_INVALID_ARGUMENT     # model_path NULL
    with pytest.raises(capi.TkError) as e:
        V.ObjectDetector(backend=capi.VISION_BACKEND_CPU)
    assert e.value.code == capi.TK_ERROR_NOT_IMPLEMENTED                                                  # no CPU fallback
    frame = synth.rgb_frame(640, 480, 1)
    with V.ObjectDetector(class_labels=LABELS) as det:
        with pytest.raises(capi.TkError) as e:
            det.detect(frame)                                   # no network bound
        assert e.value.code == 4002                             # TK_ERROR_INFERENCE_FAILED
        net = FakeNetwork(synth.logits_tensor(3)[None])
        net.fail = True
        det.bind_network(net, synth.NUM_PROPOSALS)
        with pytest.raises(capi.TkError) as e:
            det.detect(frame)
        assert e.value.code == 4002
        res = C.POINTER(capi.DetectionResult)()
        cnt = C.c_size_t(5)
        assert lib.tk_object_detector_detect(det._h, None, C.byref(res), C.byref(cnt)) == capi.TK_ERROR_INVALID_ARGUMENT
    hh = C.c_void_p()
    lib.tk_object_detector_destroy(C.byref(hh))                 # NULL handle is a no-op
    lib.tk_object_detector_update_thresholds(None, 0.1, 0.1)


def test_detector_flow_from_c(tmp_path):
    """The same call sequence from plain C (tests/c/tk_detector_flow_test.c, built by __graft_entry__.build()):
    create -> bind network -> detect x2 -> free_results -> update_thresholds -> destroy, on a 640x480 frame."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(__file__), "c", "tk_detector_flow_test")
    if not os.path.exists(exe):
        pytest.fail("tests/c/tk_detector_flow_test is not built: run __graft_entry__.build()")
    logits = synth.logits_tensor(4)
    frame = synth.rgb_frame(640, 480, 4)
    lp, fp = tmp_path / "logits.f32", tmp_path / "frame.rgb"
    logits.tofile(lp)
    frame.tofile(fp)
    out = subprocess.run([exe, str(lp), str(synth.NUM_PROPOSALS), "80", str(fp), "640", "480"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    got = []
    for line in out.stdout.splitlines():
        c, label, bits, x, y, w, h = line.split()
        got.append((int(c), label, int(bits, 16), (int(x), int(y), int(w), int(h))))
    want = oracle_records(logits, 640, 480)
    assert len(want) > 10 and got == want
