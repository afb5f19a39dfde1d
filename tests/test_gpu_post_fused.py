"""GPU parity of the fused co-resident post-process kernel (tk_kernels_postprocess_fused, csrc/post.cu): decode +
class-wise NMS per frame in one kernel — bit-exact against the oracle AND against the two-kernel path
(tk_kernels_decode_detections + tk_kernels_nms_classwise), including the intermediate candidate list."""
import numpy as np
import pytest
import torch

import synth
from oracle import oracle as O
from trackiellm_b200 import vision as V

pytestmark = pytest.mark.gpu


LAYOUT = 0     # set per test by the `layout` fixture: 0 = a cluster of 8 CTAs per frame, 1 = persistent co-resident CTAs


@pytest.fixture(autouse=True, params=[0, 1], ids=["cluster", "persistent"])
def layout(request):
    global LAYOUT
    LAYOUT = request.param
    yield request.param


def fused(ts, thr, iou, max_det, act=V.ACT_RAW, cap=4096, agnostic=False, restore=None):
    d = torch.from_numpy(np.ascontiguousarray(ts)).cuda()
    if d.dim() == 2:
        d = d[None]
    c, n, dets, nd = V.postprocess_fused(d, thr, iou, max_det, act=act, max_candidates=cap, class_agnostic=agnostic,
                                         restore=restore, layout=LAYOUT)
    torch.cuda.synchronize()
    return V.candidates_to_numpy(c, n), n.cpu().numpy(), V.candidates_to_numpy(dets, nd)


def oracle_post(t, thr, iou, max_det, act=V.ACT_RAW, cap=4096, agnostic=False, restore=None):
    c, total = O.decode(t, thr, act, cap=cap)
    cc = c
    if agnostic:
        cc = c.copy(); cc["class_id"] = 0
    d = O.nms_classwise(cc, iou, max_det)
    if agnostic:                                   # records keep their own class id; only suppression ignores it
        by_index = {int(r["index"]): r for r in c}
        d = np.array([by_index[int(i)] for i in d["index"]], dtype=c.dtype) if len(d) else d
    if restore:
        d = O.restore_coords(d, *restore)
    return c, total, d


@pytest.mark.parametrize("act", [V.ACT_RAW, V.ACT_SIGMOID, V.ACT_SOFTMAX])
@pytest.mark.parametrize("seed,adv", [(0, True), (1, False)])
def test_fused_bitexact_full_head(act, seed, adv):
    t = synth.logits_tensor(seed, adversarial=adv, as_logit=(act != V.ACT_RAW))
    cands, counts, dets = fused(t, 0.5, 0.5, 500, act)
    rc, total, rd = oracle_post(t, 0.5, 0.5, 500, act)
    assert int(counts[0]) == total
    assert cands[0].tobytes() == rc.tobytes()
    assert dets[0].tobytes() == rd.tobytes()


@pytest.mark.parametrize("P,C", [(1, 80), (31, 80), (257, 3), (1000, 1), (25200, 0), (8 * 32 * 5 + 1, 80), (5000, 300), (40000, 20)])
def test_fused_ragged_shapes(P, C):
    """C = 300: rows too long to stage two per warp (staging depth 1 or the direct path); P = 40000: a bitmap that
    leaves little staging room."""
    t = synth.logits_tensor(3, num_proposals=P, num_classes=C, objects=5, dups=3)
    for act in (V.ACT_RAW, V.ACT_SOFTMAX):
        cands, counts, dets = fused(t, 0.25, 0.5, 100, act)
        rc, total, rd = oracle_post(t, 0.25, 0.5, 100, act)
        assert int(counts[0]) == total and cands[0].tobytes() == rc.tobytes() and dets[0].tobytes() == rd.tobytes()


def test_fused_crowds_take_the_global_scratch_path():
    # 4000 candidates of 4 classes: above the 512 the shared-memory layout holds, class segments above 64 (blocked path)
    t = synth.logits_tensor(6, objects=100, dups=40, num_classes=4)
    cands, counts, dets = fused(t, 0.5, 0.45, 1000)
    rc, total, rd = oracle_post(t, 0.5, 0.45, 1000)
    assert total == 4000 and int(counts[0]) == total
    assert cands[0].tobytes() == rc.tobytes() and dets[0].tobytes() == rd.tobytes()
    # 700 candidates over 80 classes: scratch path, segment-mask algorithm
    t = synth.logits_tensor(7, objects=70, dups=10)
    _, counts, dets = fused(t, 0.5, 0.5, 500, restore=(1920, 1080, 640, 640))
    rc, total, rd = oracle_post(t, 0.5, 0.5, 500, restore=(1920, 1080, 640, 640))
    assert total > 512 and dets[0].tobytes() == rd.tobytes()
    # class-agnostic
    _, _, dets = fused(t, 0.5, 0.5, 500, agnostic=True)
    _, _, rd = oracle_post(t, 0.5, 0.5, 500, agnostic=True)
    assert np.array_equal(dets[0]["index"], rd["index"])


def test_fused_all_pass_and_caps():
    t = synth.logits_tensor(4)
    for cap in (4096, 500, 64):
        cands, counts, dets = fused(t, -1.0, 0.5, 50, cap=cap)      # every row passes: count 25200 > cap
        rc, total, rd = oracle_post(t, -1.0, 0.5, 50, cap=cap)
        assert int(counts[0]) == total == 25200
        assert cands[0].tobytes() == rc.tobytes() and dets[0].tobytes() == rd.tobytes()
    ts = np.stack([synth.logits_tensor(8), synth.logits_tensor(9)])
    ts[1, :, 4] = 0.0                                                # an empty frame
    _, counts, dets = fused(ts, 0.5, 0.5, 100)
    assert int(counts[1]) == 0 and len(dets[1]) == 0 and len(dets[0]) > 0


def test_fused_equals_two_kernel_path_on_a_long_batch():
    """400 frames on 148 persistent CTAs (every CTA loops), small heads of mixed density — the fused kernel's four
    outputs against tk_kernels_decode_detections + tk_kernels_nms_classwise, byte for byte."""
    n, P = 400, 3000
    rng = np.random.default_rng(11)
    ts = np.stack([synth.logits_tensor(500 + i, num_proposals=P, objects=int(rng.integers(0, 60)), dups=int(rng.integers(1, 16)),
                                       adversarial=(i % 50 == 0)) for i in range(n)])
    d = torch.from_numpy(ts).cuda()
    c1, n1, d1, m1 = V.postprocess_fused(d, 0.5, 0.5, 300, max_candidates=1024, layout=LAYOUT)
    c2, n2 = V.decode_detections(d, 0.5, max_candidates=1024)
    d2, m2 = V.nms_classwise(c2, n2, 0.5, 300)
    torch.cuda.synchronize()
    assert torch.equal(n1, n2) and torch.equal(m1, m2)
    a, b = V.candidates_to_numpy(c1, n1), V.candidates_to_numpy(c2, n2)
    x, y = V.candidates_to_numpy(d1, m1), V.candidates_to_numpy(d2, m2)
    for f in range(n):
        assert a[f].tobytes() == b[f].tobytes() and x[f].tobytes() == y[f].tobytes(), f
    for f in (0, 50, 399):
        _, total, rd = oracle_post(ts[f], 0.5, 0.5, 300, cap=1024)
        assert x[f].tobytes() == rd.tobytes()


def test_fused_random_configs_bitexact():
    rng = np.random.default_rng(123)
    for case in range(24):
        P = int(rng.integers(1, 30000))
        C = int(rng.integers(1, 92))
        act = (V.ACT_RAW, V.ACT_SIGMOID, V.ACT_SOFTMAX)[case % 3]
        objects, dups = int(rng.integers(0, 90)), int(rng.integers(1, 30))
        thr = float(rng.choice([0.25, 0.5, 0.5, 0.7]))
        iou = float(rng.choice([0.3, 0.45, 0.5, 0.65]))
        cap = int(rng.choice([64, 500, 4096, 4096]))
        max_det = int(rng.choice([1, 20, 300, 500]))
        n = int(rng.integers(1, 4))
        ts = np.stack([synth.logits_tensor(7000 + 10 * case + i, num_proposals=P, num_classes=C, objects=objects, dups=dups,
                                           adversarial=(P >= 64 and i == 0), as_logit=(act != V.ACT_RAW)) for i in range(n)])
        cands, counts, dets = fused(ts, thr, iou, max_det, act, cap=cap)
        for i in range(n):
            rc, total, rd = oracle_post(ts[i], thr, iou, max_det, act, cap=cap)
            assert int(counts[i]) == total, (case, i)
            assert cands[i].tobytes() == rc.tobytes() and dets[i].tobytes() == rd.tobytes(), (case, i, P, C, act, cap)


@pytest.mark.parametrize("n", [3, 40])
def test_pipeline_fused_equals_separate(n):
    """process_device with the cluster-fused post-process (fuse_postprocess = 3) and with the persistent form forked
    onto the side stream (= 2) against the three kernels in a row (= 1, the default): tensors, detections and counts
    identical; repeated calls on one stream stay ordered."""
    W, H, P = 320, 240, 4000
    frames = np.stack([synth.rgb_frame(W, H, 70 + i) for i in range(n)])
    logits = np.stack([synth.logits_tensor(70 + i, num_proposals=P, objects=20, dups=7, adversarial=(i == 0)) for i in range(n)])
    outs = []
    for mode in (1, 2, 3) if LAYOUT == 0 else (1, 2):
        with V.Pipeline(max_batch=n, width=W, height=H, tensor_w=96, tensor_h=64, out_dtype=torch.float32, arith=V.ARITH_EXACT,
                        num_proposals=P, num_classes=80, conf_thr=0.5, iou_thr=0.5, max_detections=300,
                        fuse_postprocess=mode) as pipe:
            df, dl = torch.from_numpy(frames).cuda(), torch.from_numpy(logits).cuda()
            dt, dd, dc = pipe.alloc_outputs(n)
            for rep in range(3):
                if rep == 1:                       # other logits in between: a stale result would show
                    pipe.process_device(df, torch.zeros_like(dl), dt, dd, dc)
                else:
                    pipe.process_device(df, dl, dt, dd, dc)
            torch.cuda.synchronize()
            outs.append((dt.cpu().numpy().copy(), V.candidates_to_numpy(dd, dc), dc.cpu().numpy().copy()))
    for other in outs[1:]:
        assert np.array_equal(outs[0][0].view(np.uint32), other[0].view(np.uint32))
        assert np.array_equal(outs[0][2], other[2])
        for f in range(n):
            assert outs[0][1][f].tobytes() == other[1][f].tobytes()
    for f in range(n):
        rc, _ = O.decode(logits[f], 0.5, cap=4096)
        assert outs[1][1][f].tobytes() == O.nms_classwise(rc, 0.5, 300).tobytes()


def test_no_candidate_is_dropped_silently():
    """with a threshold low enough that more than 4096 rows pass, the capped kernels used to keep the first
    4096 by index and say nothing.  Now (a) the separate launchers take any cap (NMS above 4096 candidates sorts in a
    global work area), (b) the reference-typed entry point tk_b200_postprocess_detections, which has no cap in its
    signature, answers what the UNTRUNCATED oracle answers (the reference keeps every candidate,
    src/vision/tk_object_detector.c:339-355), (c) the pipeline reports the overflow."""
    if LAYOUT == 1:
        pytest.skip("layout-independent")
    t = synth.logits_tensor(21, objects=200, dups=30)
    thr = 0.2                                   # rows with U(0, 0.3) objectness pass too
    rc, total = O.decode(t, thr)                # no cap
    assert total > 8000
    want = O.nms_classwise(rc, 0.5, 500)
    d = torch.from_numpy(t[None]).cuda()
    c, n = V.decode_detections(d, thr, max_candidates=25200)
    dets, nd = V.nms_classwise(c, n, 0.5, 500)
    torch.cuda.synchronize()
    assert int(n[0]) == total
    assert V.candidates_to_numpy(dets, nd)[0].tobytes() == want.tobytes()
    _, _, got = fused(t, thr, 0.5, 500, cap=25200)
    assert got[0].tobytes() == want.tobytes()
    recs = V.postprocess_detections(t, conf_thr=thr, iou_thr=0.5, max_detections=500)
    assert [(r[0], int(np.float32(r[2]).view(np.uint32)), r[3]) for r in recs] == \
        [(int(k["class_id"]), int(np.float32(k["confidence"]).view(np.uint32)), (int(k["x"]), int(k["y"]), int(k["w"]), int(k["h"]))) for k in want]
    with V.Pipeline(max_batch=1, width=64, height=48, tensor_w=32, tensor_h=32, out_dtype=torch.float32, conf_thr=thr,
                    max_detections=500) as pipe:
        _, hd, hc = pipe.alloc_outputs(1, pinned=True)
        pipe.process_host(None, torch.from_numpy(t[None]).pin_memory(), hd, hc)
        frames_over, max_count = pipe.candidate_overflow()
        assert frames_over == 1 and max_count == total
