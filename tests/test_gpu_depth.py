"""GPU parity: depth post-processing either side of the detector path:
MiDaS inverse depth -> metric (src/vision/tk_depth_midas.c:473-499) and the per-detection
distance statistic (src/vision/src/object_analysis.rs:227-279).  Bit-exact vs the oracle."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from trackiellm_b200 import vision as V

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


# --------------------------------------------------------------------------- #
# inverse depth -> metric
# --------------------------------------------------------------------------- #
@pytest.mark.parametrize("n,h,w", [(1, 256, 256), (5, 256, 256), (3, 384, 384), (2, 37, 53), (1, 1, 1), (2, 480, 640),
                                   (1, 1080, 1920)])
def test_inverse_depth_to_metric_bitexact(n, h, w):
    rng = np.random.default_rng(h * 1000 + w + n)
    inv = (rng.random((n, h, w), dtype=np.float32) * rng.uniform(1, 5000, (n, 1, 1)).astype(np.float32)
           + rng.uniform(-50, 50, (n, 1, 1)).astype(np.float32))
    got = V.inverse_depth_to_metric(torch.from_numpy(inv).cuda()).cpu().numpy()
    for i in range(n):
        assert np.array_equal(bits(got[i]), bits(O.inverse_depth_to_metric(inv[i]))), i


def test_inverse_depth_to_metric_edges():
    # flat map -> max_depth everywhere (:485-491); NaN inside is ignored by the scan; NaN seed sticks (:476-477)
    flat = np.full((2, 64, 64), 3.25, np.float32)
    flat[1] += np.float32(1e-7)
    got = V.inverse_depth_to_metric(torch.from_numpy(flat).cuda(), 0.1, 10.0).cpu().numpy()
    assert np.all(got == np.float32(10.0))
    rng = np.random.default_rng(1)
    a = rng.random((3, 64, 96), dtype=np.float32) * 100
    a[0, 5, 7] = np.nan                 # interior NaN: extrema unaffected, that element -> NaN
    a[1, 0, 0] = np.nan                 # NaN seed: everything NaN
    a[2, 3, 3] = np.inf                 # infinite range: reference arithmetic, whatever it gives
    got = V.inverse_depth_to_metric(torch.from_numpy(a).cuda()).cpu().numpy()
    for i in range(3):
        ref = O.inverse_depth_to_metric(a[i])
        assert np.array_equal(np.isnan(got[i]), np.isnan(ref)), i
        m = ~np.isnan(ref)
        assert np.array_equal(bits(got[i][m]), bits(ref[m])), i
    assert np.isnan(got[1]).all()


def test_inverse_depth_to_metric_in_place_and_custom_range():
    rng = np.random.default_rng(2)
    a = rng.normal(size=(4, 256, 256)).astype(np.float32)
    d = torch.from_numpy(a).cuda()
    V.inverse_depth_to_metric(d, 0.5, 80.0, out=d)
    got = d.cpu().numpy()
    for i in range(4):
        assert np.array_equal(bits(got[i]), bits(O.inverse_depth_to_metric(a[i], 0.5, 80.0)))


# --------------------------------------------------------------------------- #
# per-detection distance
# --------------------------------------------------------------------------- #
def make_dets(boxes):
    """boxes [N][K][4] ints -> candidate tensor [N,K,7] int32."""
    n, k = boxes.shape[:2]
    d = np.zeros((n, k), V.CAND_DTYPE)
    d["x"], d["y"], d["w"], d["h"] = boxes[..., 0], boxes[..., 1], boxes[..., 2], boxes[..., 3]
    d["confidence"] = 0.9
    return torch.from_numpy(d.view(np.int32).reshape(n, k, 7)).cuda()


def run_distance(boxes, counts, depth, fw, fh):
    dets = make_dets(boxes)
    cnt = torch.from_numpy(np.asarray(counts, np.int32)).cuda() if counts is not None else None
    out = V.detection_distance(dets, cnt, torch.from_numpy(depth).cuda(), fw, fh)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_distance_reference_unit_test_values():
    """object_analysis.rs:286-347: 100x100 map, 20x20 box at (10,10) over depth 10.0 -> 10.0; moved box over 12.0 -> 12.0."""
    d1 = np.zeros((100, 100), np.float32); d1[10:30, 10:30] = 10.0
    d2 = np.zeros((100, 100), np.float32); d2[11:31, 11:31] = 12.0
    depth = np.stack([d1, d2])
    boxes = np.array([[[10, 10, 20, 20]], [[11, 11, 20, 20]]], np.int32)
    got = run_distance(boxes, [1, 1], depth, 100, 100)
    assert got[0, 0] == np.float32(10.0) and got[1, 0] == np.float32(12.0)
    assert O.detection_distance(boxes[0, 0], d1, 100, 100) == 10.0
    assert O.detection_distance(boxes[1, 0], d2, 100, 100) == 12.0


@pytest.mark.parametrize("dw,dh,fw,fh", [(100, 100, 100, 100), (256, 256, 1920, 1080), (384, 384, 640, 640),
                                         (128, 96, 640, 480)])
def test_distance_bitexact_random(dw, dh, fw, fh):
    rng = np.random.default_rng(dw + fw)
    n, k = 3, 40
    depth = rng.gamma(2.0, 3.0, (n, dh, dw)).astype(np.float32)
    depth[rng.random((n, dh, dw)) < 0.1] = 0.0            # invalid samples
    depth[rng.random((n, dh, dw)) < 0.02] = 250.0         # beyond 100 m
    depth[0, : dh // 2] = np.round(depth[0, : dh // 2])   # many ties
    boxes = np.zeros((n, k, 4), np.int32)
    boxes[..., 0] = rng.integers(-20, fw, (n, k)); boxes[..., 1] = rng.integers(-20, fh, (n, k))
    boxes[..., 2] = rng.integers(0, fw // 2, (n, k)); boxes[..., 3] = rng.integers(0, fh // 2, (n, k))
    boxes[0, 0] = (0, 0, fw, fh)                          # x+w == frame_w -> depth_w-1: whole map
    boxes[0, 1] = (0, 0, fw + 1, fh)                      # rounds to depth_w-1 or beyond -> maybe -1
    boxes[0, 2] = (5, 5, 0, 10)                           # degenerate
    boxes[0, 3] = (fw // 3, fh // 3, 2, 2)                # tiny: fewer than 10 samples
    counts = [k, k - 7, 0]
    got = run_distance(boxes, counts, depth, fw, fh)
    for f in range(n):
        for j in range(k):
            want = O.detection_distance(boxes[f, j], depth[f], fw, fh) if j < counts[f] else -1.0
            assert bits(got[f, j]) == bits(np.float32(want)), (f, j, boxes[f, j], got[f, j], want)
    assert (got[0] > 0).sum() > 10


def test_distance_big_windows_use_workspace():
    """Windows with more valid samples than the shared-memory buffer (16384) sort through the global
    workspace; several at once so slots are contended."""
    rng = np.random.default_rng(5)
    dw = dh = 256
    n, k = 2, 24
    depth = (rng.random((n, dh, dw), dtype=np.float32) * 60 + 0.2).astype(np.float32)
    depth[1, 100:140] = 1000.0
    boxes = np.zeros((n, k, 4), np.int32)
    for j in range(k):
        boxes[:, j] = (j, 2 * j, 640 - 2 * j - (j % 3) * 150, 640 - 3 * j - (j % 4) * 100)
    got = run_distance(boxes, None, depth, 640, 640)
    sizes = []
    for f in range(n):
        for j in range(k):
            want = O.detection_distance(boxes[f, j], depth[f], 640, 640)
            assert bits(got[f, j]) == bits(np.float32(want)), (f, j, got[f, j], want)
    assert np.all(got > 0)


def test_distance_errors():
    from trackiellm_b200 import _capi as capi
    import ctypes as C
    p = capi.DistanceParams()
    assert capi.lib().tk_kernels_detection_distance(C.byref(p), None) == capi.TK_ERROR_INVALID_ARGUMENT
    dets = make_dets(np.zeros((1, 4, 4), np.int32))
    depth = torch.zeros((1, 256, 256), device="cuda")
    out = torch.empty((1, 4), device="cuda")
    p = capi.DistanceParams(dets.data_ptr(), None, 1, 4, depth.data_ptr(), 256, 256, 640, 640, out.data_ptr(), None, 0)
    assert capi.lib().tk_kernels_detection_distance(C.byref(p), None) == capi.TK_ERROR_BUFFER_TOO_SMALL
    p.depth_width = 0
    assert capi.lib().tk_kernels_detection_distance(C.byref(p), None) == capi.TK_ERROR_INVALID_DIMENSIONS
