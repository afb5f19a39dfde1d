"""Small GPU cases that reach every kernel template variant once, each checked against the oracle.

Three uses: (1) part of the ordinary `-m gpu` suite; (2) the core of what `tools/run_checked.sh` runs against the
CHECKED build of the library (device-side assertions + ring-content verification, csrc/tk_common.cuh) — the stand-in
for compute-sanitizer, which is closed on this GPU pool (profiles/r2_sanitizer.md); (3) the selection
`tools/run_sanitizer.sh` runs under `compute-sanitizer --tool memcheck | racecheck | synccheck | initcheck`
where that tool is open — the inputs are kept small so that its 10-1000x slow-down still ends in minutes.
Variants covered:

  pre-process : TMA and cooperative-copy staging, dense and sparse row staging, RGB8 / BGR8 / RGBA8 / I420,
                stretch and letterbox, EXACT / FAST / HALF arithmetic, fp32 and fp16 tensors, several CTAs per frame
                and several frames per CTA (ring wrap-around)
  decode      : RAW / SIGMOID / SOFTMAX, objectness read from the rows and PACKED (pipeline host path), the
                overflow path (> 256 survivors in one CTA slice), the candidate cap
  NMS         : register path (<= 512), shared-memory sort path (> 512), blocked path (crowd of one class,
                class-agnostic), the two-launch path above one wave, coordinate restore
  post (fused): the co-resident decode + NMS kernel of the pipeline, when present
  rows        : softmax tile (<= 128 columns) and long-row kernels, argmax
  attributes  : dominant colour, door state
  depth       : post-process, point cloud, inverse depth -> metric (cluster / DSMEM), per-detection distance
                (shared-memory sort and the global-workspace path)
  dispatcher  : upload -> pre-process -> download on the three event-ordered streams
"""
import numpy as np
import pytest
import torch

import synth
from oracle import oracle as O
from trackiellm_b200 import vision as V

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def pre(frames, ow, oh, **kw):
    out = V.preprocess_batch(torch.from_numpy(np.ascontiguousarray(frames)).cuda(), ow, oh, **kw)
    torch.cuda.synchronize()
    return out.cpu().numpy()


# ------------------------------------------------------------------------------------------------ pre-process
@pytest.mark.parametrize("w,h,ow,oh,n", [
    (320, 240, 160, 128, 3),     # TMA (960-byte rows), dense
    (301, 77, 128, 64, 2),       # cooperative copy (903-byte rows), dense
    (320, 240, 96, 64, 3),       # TMA, sparse (y_ratio 3.7)
    (301, 203, 64, 40, 2),       # cooperative copy, sparse
    (64, 48, 320, 320, 1),       # up-scaling: every source row feeds several output rows
])
def test_pre_rgb8_staging_variants(w, h, ow, oh, n):
    fr = np.stack([synth.rgb_frame(w, h, 10 + i) for i in range(n)])
    got = pre(fr, ow, oh)
    for i in range(n):
        assert np.array_equal(bits(got[i]), bits(O.preprocess_rgb8(fr[i], ow, oh)))
    fast = pre(fr, ow, oh, arith=V.ARITH_FAST)
    ref = np.stack([O.preprocess_rgb8(f, ow, oh) for f in fr])
    assert np.all(np.abs(fast - ref) <= 1e-5 + 1e-5 * np.abs(ref))


@pytest.mark.parametrize("arith", [V.ARITH_EXACT, V.ARITH_FAST, V.ARITH_HALF])
def test_pre_fp16_arith_variants(arith):
    fr = np.stack([synth.rgb_frame(320, 240, 20 + i) for i in range(2)])
    got = pre(fr, 160, 128, out_dtype=torch.float16, arith=arith)
    for i in range(2):
        ref = O.preprocess_rgb8(fr[i], 160, 128)
        if arith == V.ARITH_EXACT:
            assert np.array_equal(got[i].view(np.uint16), O.f32_to_f16(ref).view(np.uint16))
        else:
            assert np.all(np.abs(got[i].astype(np.float32) - ref) <= 1e-2 + 1e-2 * np.abs(ref))


def test_pre_rgba_bgr_letterbox():
    fr = synth.rgb_frame(320, 200, 6)
    rgba = np.concatenate([fr, np.full((200, 320, 1), 77, np.uint8)], axis=2)
    assert np.array_equal(bits(pre(rgba[None], 96, 96, fmt=V.PIXEL_RGBA8)[0]), bits(O.preprocess_rgb8(fr, 96, 96)))
    got = pre(rgba[None], 96, 96, fmt=V.PIXEL_RGBA8, out_dtype=torch.float16, arith=V.ARITH_HALF)[0].astype(np.float32)
    ref = O.preprocess_rgb8(fr, 96, 96)
    assert np.all(np.abs(got - ref) <= 1e-2 + 1e-2 * np.abs(ref))
    bgr = np.ascontiguousarray(fr[:, :, ::-1])
    assert np.array_equal(bits(pre(bgr.reshape(1, -1), 96, 96, width=320, height=200, fmt=V.PIXEL_BGR8)[0]), bits(ref))
    for (w, h) in ((320, 240), (240, 320)):
        f2 = synth.rgb_frame(w, h, 2)
        got = pre(f2[None], 160, 160, resize=V.RESIZE_LETTERBOX, pad_value=114.0)[0]
        assert np.array_equal(bits(got), bits(O.preprocess_rgb8(f2, 160, 160, resize=O.RESIZE_LETTERBOX, pad_value=114.0)))


@pytest.mark.parametrize("w,h,ow,oh", [(256, 128, 160, 96), (256, 256, 64, 64), (130, 66, 96, 64)])
def test_pre_i420_variants(w, h, ow, oh):
    buf = np.stack([synth.i420_frame(w, h, 8), synth.i420_frame(w, h, 9)])
    got = pre(buf, ow, oh, width=w, height=h, fmt=V.PIXEL_I420)
    for i in range(2):
        assert np.array_equal(bits(got[i]), bits(O.preprocess_rgb8(O.i420_to_rgb8(buf[i], w, h), ow, oh)))


def test_pre_many_frames_per_cta_ring_wraps():
    # 600 tiny frames: more frames than CTAs, every CTA walks several items and wraps its ring many times
    fr = np.stack([synth.rgb_frame(64, 32, 100 + i) for i in range(600)])
    got = pre(fr, 32, 24)
    for i in (0, 1, 299, 598, 599):
        assert np.array_equal(bits(got[i]), bits(O.preprocess_rgb8(fr[i], 32, 24)))


# ------------------------------------------------------------------------------------------------ decode / NMS
def post(ts, thr, iou, max_det, act=V.ACT_RAW, cap=4096, agnostic=False, restore=None):
    d = torch.from_numpy(np.ascontiguousarray(ts)).cuda()
    c, n = V.decode_detections(d, thr, act=act, max_candidates=cap)
    dets, nd = V.nms_classwise(c, n, iou, max_det, class_agnostic=agnostic, restore=restore)
    torch.cuda.synchronize()
    return V.candidates_to_numpy(c, n), n.cpu().numpy(), V.candidates_to_numpy(dets, nd)


@pytest.mark.parametrize("act", [V.ACT_RAW, V.ACT_SIGMOID, V.ACT_SOFTMAX])
def test_decode_nms_register_path(act):
    ts = np.stack([synth.logits_tensor(50 + i, num_proposals=3000, objects=12, dups=8, adversarial=(i == 0),
                                       as_logit=(act != V.ACT_RAW)) for i in range(2)])
    cands, counts, dets = post(ts, 0.5, 0.5, 300, act)
    for i in range(2):
        rc, total = O.decode(ts[i], 0.5, act)
        assert int(counts[i]) == total and cands[i].tobytes() == rc.tobytes()
        assert dets[i].tobytes() == O.nms_classwise(rc, 0.5, 300).tobytes()


def test_decode_overflow_and_cap_paths():
    t = synth.logits_tensor(4, num_proposals=6000)
    cands, counts, _ = post(t[None], -1.0, 0.5, 10, cap=1024)       # every row passes: > 256 per CTA slice, > cap
    rc, total = O.decode(t, -1.0, cap=1024)
    assert int(counts[0]) == total == 6000 and cands[0].tobytes() == rc.tobytes()


def test_nms_sorted_and_blocked_paths_and_restore():
    # > 512 candidates spread over 80 classes: shared-memory sort path, no class above 64
    t = synth.logits_tensor(6, num_proposals=8000, objects=70, dups=10)
    rc, total = O.decode(t, 0.5)
    assert total > 512
    _, _, dets = post(t[None], 0.5, 0.45, 400, restore=(1920, 1080, 640, 640))
    ref = O.restore_coords(O.nms_classwise(rc, 0.45, 400), 1920, 1080, 640, 640)
    assert dets[0].tobytes() == ref.tobytes()
    # a crowd of few classes (segments > 64) and class-agnostic mode: blocked path
    t = synth.logits_tensor(7, num_proposals=8000, objects=30, dups=30, num_classes=3)
    rc, _ = O.decode(t, 0.5)
    _, _, dets = post(t[None], 0.5, 0.5, 500)
    assert dets[0].tobytes() == O.nms_classwise(rc, 0.5, 500).tobytes()
    _, _, dets = post(t[None], 0.5, 0.5, 500, agnostic=True)
    ag = rc.copy(); ag["class_id"] = 0
    want = O.nms_classwise(ag, 0.5, 500)
    assert np.array_equal(dets[0]["index"], want["index"])


def test_nms_two_launch_path_above_one_wave():
    rng = np.random.default_rng(5)
    n, cap = 200, 1024
    sizes = rng.choice([0, 1, 40, 200, 512, 513, 700], n)
    host = np.zeros((n, cap), O.CAND_DTYPE)
    refs = []
    for f, k in enumerate(sizes):
        k = int(k)
        c = np.zeros(k, O.CAND_DTYPE)
        c["index"] = np.sort(rng.choice(25200, k, replace=False))
        cx, cy = rng.uniform(0, 640, k), rng.uniform(0, 640, k)
        w, h = rng.uniform(8, 200, k), rng.uniform(8, 200, k)
        c["x"], c["y"], c["w"], c["h"] = (cx - w / 2).astype(np.int32), (cy - h / 2).astype(np.int32), w.astype(np.int32), h.astype(np.int32)
        c["confidence"] = rng.choice(np.linspace(0.5, 1.0, 53, dtype=np.float32), k)
        c["class_id"] = rng.integers(0, 2 if k > 600 else 40, k)
        host[f, :k] = c
        refs.append(O.nms_classwise(c, 0.5, 100))
    dets, nd = V.nms_classwise(torch.from_numpy(host.view(np.int32).reshape(n, cap, 7)).cuda(),
                               torch.from_numpy(sizes.astype(np.int32)).cuda(), 0.5, 100)
    got = V.candidates_to_numpy(dets, nd)
    for f in range(n):
        assert got[f].tobytes() == refs[f].tobytes(), f


# ------------------------------------------------------------------------------------------------ pipeline
@pytest.mark.parametrize("stage", [0, 1, 2])
def test_pipeline_host_and_device_paths(stage):
    """process_device (the fused step) and process_host in the three host-logit modes (0 = PACKED decode: needs
    n * P >= 65536 rows to take the packing path)."""
    n, p = 3, 25200
    W, H = 320, 240
    frames = np.stack([synth.rgb_frame(W, H, 70 + i) for i in range(n)])
    logits = np.stack([synth.logits_tensor(70 + i, num_proposals=p, objects=14, dups=7, adversarial=(i == 0)) for i in range(n)])
    with V.Pipeline(max_batch=n, chunk_frames=2, width=W, height=H, tensor_w=96, tensor_h=64, out_dtype=torch.float32,
                    arith=V.ARITH_EXACT, num_proposals=p, num_classes=80, conf_thr=0.5, iou_thr=0.5, max_detections=300,
                    stage_host_logits=stage) as pipe:
        hf, hl = torch.from_numpy(frames).pin_memory(), torch.from_numpy(logits).pin_memory()
        _, hd, hc = pipe.alloc_outputs(n, pinned=True)
        ht = torch.empty((n, 3, 64, 96), dtype=torch.float32).pin_memory()
        pipe.process_host(hf, hl, hd, hc, tensor=ht)
        dt, dd, dc = pipe.alloc_outputs(n)
        pipe.process_device(hf.cuda(), hl.cuda(), dt, dd, dc)
        torch.cuda.synchronize()
        got_h = V.candidates_to_numpy(hd.cuda(), hc.cuda())
        got_d = V.candidates_to_numpy(dd, dc)
        for i in range(n):
            rc, _ = O.decode(logits[i], 0.5, cap=4096)
            ref = O.nms_classwise(rc, 0.5, 300)
            assert got_h[i].tobytes() == ref.tobytes() and got_d[i].tobytes() == ref.tobytes()
            t_ref = O.preprocess_rgb8(frames[i], 96, 64)
            assert np.array_equal(bits(ht[i].numpy()), bits(t_ref)) and np.array_equal(bits(dt[i].cpu().numpy()), bits(t_ref))


# ------------------------------------------------------------------------------------------------ rows / attributes / depth
@pytest.mark.parametrize("rows,cols", [(33, 80), (65, 129), (3, 5000), (1, 1)])
def test_rows_softmax_argmax(rows, cols):
    x = np.random.default_rng(rows * 1000 + cols).normal(0, 3, (rows, cols)).astype(np.float32)
    sm, am = V.argmax_rows(torch.from_numpy(x).cuda())
    torch.cuda.synchronize()
    assert np.array_equal(bits(sm.cpu().numpy()), bits(O.softmax_rows(x, det_exp=True)))
    assert np.array_equal(am.cpu().numpy().astype(np.uint32), O.argmax_rows(x))
    assert np.array_equal(bits(V.kernels_softmax(torch.from_numpy(x).cuda()).cpu().numpy()), bits(O.softmax_rows(x, det_exp=True)))


def test_attributes_small():
    fr = synth.rgb_frame(160, 120, 3)
    fr[20:60, 30:90] = (250, 10, 10)
    boxes = np.array([[30, 20, 60, 40], [0, 0, 160, 120], [-5, -5, 20, 20], [150, 110, 30, 30]], np.int32)
    ids, hist = V.dominant_color(torch.from_numpy(fr).cuda(), torch.from_numpy(boxes).cuda(), want_bins=True)
    st, ed = V.door_state(torch.from_numpy(fr).cuda(), torch.from_numpy(boxes).cuda(), want_edges=True)
    torch.cuda.synchronize()
    for i, b in enumerate(boxes):
        k, ref = O.dominant_color(fr, b)
        assert int(ids[i]) == k and np.array_equal(hist[i].cpu().numpy(), ref)
        if i < 2:      # the reference's door-state loop does not clip (tk_attribute_classifier.c:119-121): in-frame boxes only
            s, cnt = O.door_state(fr, b)
            assert int(st[i]) == s and int(ed[i]) == cnt


def test_depth_small():
    rng = np.random.default_rng(1)
    raw = rng.uniform(-2, 1200, (41, 53)).astype(np.float32)
    d = torch.from_numpy(raw).cuda()
    m = V.kernels_postprocess_depth_map(d, 0.37, 1.5).cpu().numpy()
    assert np.array_equal(bits(m), bits(raw * np.float32(0.37) + np.float32(1.5)))
    pts = V.kernels_depth_to_point_cloud(d, 300.0, 310.0, 26.0, 20.0).cpu().numpy()
    v, u = np.meshgrid(np.arange(41, dtype=np.float32), np.arange(53, dtype=np.float32), indexing="ij")
    ok = (raw > 0) & (raw <= 1000)
    ref = np.where(ok[..., None], np.stack([(u - np.float32(26.0)) * raw / np.float32(300.0),
                                            (v - np.float32(20.0)) * raw / np.float32(310.0), raw], axis=-1), 0).astype(np.float32)
    assert np.array_equal(bits(pts), bits(ref))
    for shape in ((2, 37, 53), (1, 256, 256)):
        inv = rng.uniform(0.0, 900.0, shape).astype(np.float32)
        got = V.inverse_depth_to_metric(torch.from_numpy(inv).cuda()).cpu().numpy()
        for i in range(shape[0]):
            assert np.array_equal(bits(got[i]), bits(O.inverse_depth_to_metric(inv[i])))


@pytest.mark.parametrize("dw,dh,big", [(100, 100, False), (256, 256, True)])
def test_distance_small_and_workspace(dw, dh, big):
    rng = np.random.default_rng(dw)
    n, k = 2, 6
    depth = (rng.random((n, dh, dw), dtype=np.float32) * 60 + 0.2).astype(np.float32)
    depth[rng.random((n, dh, dw)) < 0.1] = 0.0
    boxes = np.zeros((n, k, 4), np.int32)
    for j in range(k):
        boxes[:, j] = (j, 2 * j, 640 - 40 * j, 640 - 50 * j) if big else (5 * j, 3 * j, 200 + 30 * j, 150 + 20 * j)
    d = np.zeros((n, k), V.CAND_DTYPE)
    d["x"], d["y"], d["w"], d["h"] = boxes[..., 0], boxes[..., 1], boxes[..., 2], boxes[..., 3]
    out = V.detection_distance(torch.from_numpy(d.view(np.int32).reshape(n, k, 7)).cuda(), None,
                               torch.from_numpy(depth).cuda(), 640, 640)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for f in range(n):
        for j in range(k):
            assert bits(got[f, j]) == bits(np.float32(O.detection_distance(boxes[f, j], depth[f], 640, 640))), (f, j)


def test_dispatcher_streams():
    fr = synth.rgb_frame(320, 240, 0)
    out = np.empty((3, 128, 160), np.float32)
    with V.CudaDispatcher(0) as disp:
        bi, bo = disp.malloc(fr.nbytes), disp.malloc(out.nbytes)
        for _ in range(3):                       # repeated upload -> compute -> download on the event-ordered streams
            disp.upload_async(bi, fr)
            disp.preprocess_image(bi, 320, 240, bo, 160, 128)
            disp.download_async(out, bo)
        disp.synchronize()
        disp.free(bi); disp.free(bo)
    assert np.array_equal(bits(out), bits(O.preprocess_rgb8(fr, 160, 128)))
