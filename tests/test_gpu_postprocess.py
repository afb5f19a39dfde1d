"""GPU parity: decode + ordered compaction, class-wise NMS, tk_cuda_tensor_nms — bit-exact vs the oracle."""
import numpy as np
import pytest
import torch

import synth
from oracle import oracle as O
from trackiellm_b200 import vision as V

pytestmark = pytest.mark.gpu


def same(a, b):
    if a.shape != b.shape:
        print(f"shape mismatch: {a.shape} vs {b.shape}")
        return False
    if a.tobytes() == b.tobytes():
        return True
    ab = a.view(np.uint32).reshape(len(a), -1)
    bb = b.view(np.uint32).reshape(len(b), -1)
    bad = np.nonzero((ab != bb).any(axis=1))[0]
    for i in bad[:8]:
        print(f"record {i}: gpu {a[i]} {ab[i]}  oracle {b[i]} {bb[i]}")
    return False


def gpu_decode(t, thr, act=V.ACT_RAW, cap=4096):
    d = torch.from_numpy(np.ascontiguousarray(t)).cuda()
    if d.dim() == 2:
        d = d[None]
    c, n = V.decode_detections(d, thr, act=act, max_candidates=cap)
    torch.cuda.synchronize()
    return c, n


@pytest.mark.parametrize("act", [V.ACT_RAW, V.ACT_SIGMOID, V.ACT_SOFTMAX])
@pytest.mark.parametrize("seed,adv", [(0, True), (1, False), (2, True)])
def test_decode_bitexact(act, seed, adv):
    t = synth.logits_tensor(seed, adversarial=adv, as_logit=(act != V.ACT_RAW))
    c, n = gpu_decode(t, 0.5, act)
    ref, total = O.decode(t, 0.5, act)
    got = V.candidates_to_numpy(c, n)[0]
    assert int(n[0]) == total
    assert same(got, ref.view(V.CAND_DTYPE))


@pytest.mark.parametrize("P,C", [(1, 80), (31, 80), (257, 3), (1000, 1), (25200, 0), (8 * 32 * 5 + 1, 80), (5000, 300)])
def test_decode_ragged_shapes(P, C):
    t = synth.logits_tensor(3, num_proposals=P, num_classes=C, objects=5, dups=3)
    for act in (V.ACT_RAW, V.ACT_SOFTMAX):
        c, n = gpu_decode(t, 0.25, act)
        ref, total = O.decode(t, 0.25, act, cap=4096)
        assert int(n[0]) == total and same(V.candidates_to_numpy(c, n)[0], ref.view(V.CAND_DTYPE))


def test_decode_all_pass_and_cap():
    t = synth.logits_tensor(4)
    c, n = gpu_decode(t, -1.0, cap=4096)            # every row passes: count 25200 > cap
    ref, total = O.decode(t, -1.0, cap=4096)
    assert int(n[0]) == total == 25200
    assert same(V.candidates_to_numpy(c, n)[0], ref.view(V.CAND_DTYPE))
    c, n = gpu_decode(t, 2.0)                       # nothing passes
    assert int(n[0]) == 0


def test_decode_batch_matches_singles():
    ts = np.stack([synth.logits_tensor(s, adversarial=(s == 0)) for s in range(6)])
    c, n = gpu_decode(ts, 0.5)
    lists = V.candidates_to_numpy(c, n)
    for i in range(6):
        ref, total = O.decode(ts[i], 0.5)
        assert int(n[i]) == total and same(lists[i], ref.view(V.CAND_DTYPE))


def full_post(ts, thr, iou, max_det, act=V.ACT_RAW, cap=4096, agnostic=False):
    c, n = gpu_decode(ts, thr, act, cap)
    d, m = V.nms_classwise(c, n, iou, max_det, class_agnostic=agnostic)
    torch.cuda.synchronize()
    return V.candidates_to_numpy(d, m)


@pytest.mark.parametrize("act", [V.ACT_RAW, V.ACT_SIGMOID, V.ACT_SOFTMAX])
def test_decode_nms_bitexact_batch(act):
    ts = np.stack([synth.logits_tensor(s, adversarial=(s % 2 == 0), as_logit=(act != V.ACT_RAW)) for s in range(8)])
    got = full_post(ts, 0.5, 0.5, 500, act)
    for i in range(8):
        c, _ = O.decode(ts[i], 0.5, act, cap=4096)
        ref = O.nms_classwise(c, 0.5, 500)
        assert same(got[i], ref.view(V.CAND_DTYPE)), i


@pytest.mark.parametrize("iou,max_det", [(0.0, 500), (0.3, 7), (0.7, 500), (1.0, 500)])
def test_nms_thresholds_and_caps(iou, max_det):
    t = synth.logits_tensor(5, adversarial=True, objects=80, dups=20)
    got = full_post(t, 0.5, iou, max_det)[0]
    c, _ = O.decode(t, 0.5)
    assert same(got, O.nms_classwise(c, iou, max_det).view(V.CAND_DTYPE))


def test_nms_large_dense_crowd():
    # 4000 candidates, few classes, heavy overlap: exercises many 64-blocks and the sweep
    t = synth.logits_tensor(6, objects=100, dups=40, num_classes=4)
    got = full_post(t, 0.5, 0.45, 1000)[0]
    c, total = O.decode(t, 0.5, cap=4096)
    assert total == 4000
    assert same(got, O.nms_classwise(c, 0.45, 1000).view(V.CAND_DTYPE))


def test_nms_idempotent_on_gpu():
    t = synth.logits_tensor(7)
    c, n = gpu_decode(t, 0.5)
    d, m = V.nms_classwise(c, n, 0.5, 500)
    d2, m2 = V.nms_classwise(d, m, 0.5, 500)
    torch.cuda.synchronize()
    assert same(V.candidates_to_numpy(d, m)[0], V.candidates_to_numpy(d2, m2)[0])


def test_nms_more_frames_than_sms_mixed_sizes():
    """Batches above the SM count take the two-launch path (frames of <= 512 candidates with 25 KB of shared
    memory, looping CTAs; the crowded frames in a second launch): candidate counts below, at and above the split,
    empty frames, a count above the cap — every frame against the oracle."""
    rng = np.random.default_rng(77)
    n, cap = 700, 2048
    sizes = rng.choice([0, 1, 63, 64, 65, 300, 511, 512, 513, 900, 2048, 2500], n)
    host = np.zeros((n, cap), O.CAND_DTYPE)
    refs = []
    for f, k in enumerate(sizes):
        kk = min(int(k), cap)
        c = np.zeros(kk, O.CAND_DTYPE)
        c["index"] = np.sort(rng.choice(25200, kk, replace=False))
        cx, cy = rng.uniform(0, 640, kk), rng.uniform(0, 640, kk)
        if kk > 600:                                       # a crowd of one class: the blocked path
            cx, cy = rng.normal(320, 40, kk), rng.normal(320, 40, kk)
        w, h = rng.uniform(8, 200, kk), rng.uniform(8, 200, kk)
        c["x"], c["y"], c["w"], c["h"] = (cx - w / 2).astype(np.int32), (cy - h / 2).astype(np.int32), w.astype(np.int32), h.astype(np.int32)
        c["confidence"] = rng.choice(np.linspace(0.5, 1.0, 97, dtype=np.float32), kk)   # many exact ties
        c["class_id"] = rng.integers(0, 3 if kk > 600 else 80, kk)
        host[f, :kk] = c
        refs.append(O.nms_classwise(c, 0.5, 300))
    cands = torch.from_numpy(host.view(np.int32).reshape(n, cap, V.CAND_WORDS)).cuda()
    counts = torch.from_numpy(sizes.astype(np.int32)).cuda()
    dets, ndet = V.nms_classwise(cands, counts, 0.5, 300)
    got = V.candidates_to_numpy(dets, ndet)
    for f in range(n):
        assert same(got[f], refs[f].view(V.CAND_DTYPE)), (f, int(sizes[f]))


def test_nms_empty_frames():
    ts = np.stack([synth.logits_tensor(8), synth.logits_tensor(9)])
    ts[1, :, 4] = 0.0
    got = full_post(ts, 0.5, 0.5, 100)
    assert len(got[1]) == 0 and len(got[0]) > 0


def test_cuda_tensor_nms_reference_signature():
    # tk_cuda_tensor_nms (src/gpu/extensions/cuda/tk_cuda_tensor_ops.h:513-522): class-agnostic, score >= thr
    rng = np.random.default_rng(0)
    k = 1500
    boxes = np.concatenate([rng.uniform(0, 500, (k, 2)), rng.uniform(5, 120, (k, 2))], axis=1).astype(np.float32)
    scores = rng.uniform(0, 1, k).astype(np.float32)
    scores[10] = 0.25                                   # == threshold: kept by >=
    cands = np.zeros(k, O.CAND_DTYPE)
    cands["index"] = np.arange(k)
    cands["x"], cands["y"], cands["w"], cands["h"] = boxes.astype(np.int32).T
    cands["confidence"] = scores
    ref = O.nms_classwise(cands[scores >= 0.25], 0.4, 200)
    with V.CudaDispatcher(0) as disp:
        bb, bs, bi = disp.malloc(boxes.nbytes), disp.malloc(scores.nbytes), disp.malloc(200 * 4)
        disp.upload_async(bb, boxes)
        disp.upload_async(bs, scores)
        n = disp.tensor_nms(bb, bs, k, bi, 200, 0.4, 0.25)
        idx = np.full(200, -7, np.int32)
        disp.download_async(idx, bi)
        disp.synchronize()
        for b in (bb, bs, bi):
            disp.free(b)
    assert n == len(ref)
    assert np.array_equal(idx[:n], ref["index"]) and np.all(idx[n:] == -1)


def test_random_post_configs_bitexact():
    """40 seeded random head shapes / thresholds / caps / activations / crowd densities — ragged proposal counts
    (not multiples of the cluster slice), 1..91 classes, candidate caps that truncate, class-agnostic mode via
    identical class ids — every kept record bit-for-bit against the oracle."""
    rng = np.random.default_rng(77)
    for case in range(40):
        P = int(rng.integers(1, 30000))
        C = int(rng.integers(1, 92))
        act = (V.ACT_RAW, V.ACT_SIGMOID, V.ACT_SOFTMAX)[case % 3]
        objects = int(rng.integers(0, 90))
        dups = int(rng.integers(1, 30))
        thr = float(rng.choice([0.25, 0.5, 0.5, 0.7]))
        iou = float(rng.choice([0.3, 0.45, 0.5, 0.65]))
        cap = int(rng.choice([64, 500, 4096, 4096]))
        max_det = int(rng.choice([1, 20, 300, 500]))
        agnostic = bool(case % 5 == 4)
        n = int(rng.integers(1, 4))
        ts = np.stack([synth.logits_tensor(9000 + 10 * case + i, num_proposals=P, num_classes=C, objects=objects, dups=dups,
                                           adversarial=(P >= 64 and i == 0), as_logit=(act != V.ACT_RAW)) for i in range(n)])
        got = full_post(ts, thr, iou, max_det, act, cap=cap, agnostic=agnostic)
        for i in range(n):
            c, total = O.decode(ts[i], thr, act, cap=cap)
            if agnostic:
                c = c.copy()
                cls = c["class_id"].copy()
                c["class_id"] = 0
                ref = O.nms_classwise(c, iou, max_det)
                # the oracle ran on class 0; put the real class ids back by proposal index
                back = {int(k): int(v) for k, v in zip(c["index"], cls)}
                ref["class_id"] = [back[int(k)] for k in ref["index"]]
            else:
                ref = O.nms_classwise(c, iou, max_det)
            assert same(got[i], ref.view(V.CAND_DTYPE)), (case, P, C, act, objects, dups, thr, iou, cap, max_det, agnostic, i)
