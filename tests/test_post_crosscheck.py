"""CPU: the C oracle's decode + IoU + class-wise NMS (oracle/tk_oracle.c) against a second, independent statement of the
same path in NumPy (tests/np_post.py, written from SURVEY Appendix A.2 / A.3 and the reference lines cited there).

The reference cannot pin these functions (its postprocess_detections is a stub, src/vision/tk_object_detector.c:351,357,371),
so the oracle was the only statement of them in round 1 — written by the same hand as the kernels.  This differential
fuzz removes the common mode: two statements, different authorship of every helper (float->int, NaN, total order, IoU),
same answers bit for bit in raw mode on the 40 random configurations the GPU parity test uses; in sigmoid / softmax mode,
where the oracle and the kernels share a software expf, scores agree within 1e-5 relative of a libm-grade exp and the
kept lists are identical whenever that tolerance cannot reorder them."""
import numpy as np
import pytest

import np_post as N
import synth
from oracle import oracle as O


def random_configs(count=40):
    rng = np.random.default_rng(77)            # the stream of tests/test_gpu_postprocess.py::test_random_post_configs_bitexact
    for case in range(count):
        P = int(rng.integers(1, 30000))
        C = int(rng.integers(1, 92))
        act = case % 3
        objects = int(rng.integers(0, 90))
        dups = int(rng.integers(1, 30))
        thr = float(rng.choice([0.25, 0.5, 0.5, 0.7]))
        iou = float(rng.choice([0.3, 0.45, 0.5, 0.65]))
        cap = int(rng.choice([64, 500, 4096, 4096]))
        max_det = int(rng.choice([1, 20, 300, 500]))
        agnostic = bool(case % 5 == 4)
        n = int(rng.integers(1, 4))
        yield case, P, C, act, objects, dups, thr, iou, cap, max_det, agnostic, n


def test_raw_mode_bit_exact_on_the_random_configs():
    checked = 0
    for case, P, C, act, objects, dups, thr, iou, cap, max_det, agnostic, n in random_configs():
        for i in range(n):
            t = synth.logits_tensor(9000 + 10 * case + i, num_proposals=P, num_classes=C, objects=objects, dups=dups,
                                    adversarial=(P >= 64 and i == 0), as_logit=False)
            c_or, total_or = O.decode(t, thr, O.ACT_RAW, cap=cap)
            c_np, total_np = N.decode(t, thr, N.RAW, cap=cap)
            assert total_or == total_np, (case, i)
            assert c_or.tobytes() == c_np.tobytes(), (case, i, P, C)
            k_in = c_or.copy()
            if agnostic:
                k_in["class_id"] = 0
            k_or = O.nms_classwise(k_in, iou, max_det)
            k_np = N.nms_classwise(c_np, iou, max_det, class_agnostic=agnostic)
            assert np.array_equal(k_or["index"], k_np["index"]), (case, i)
            if not agnostic:
                assert k_or.tobytes() == k_np.tobytes(), (case, i)
            checked += 1
    assert checked >= 40


def test_full_head_adversarial_rows_bit_exact():
    for seed in (0, 2):
        t = synth.logits_tensor(seed, adversarial=True)
        c_or, total = O.decode(t, 0.5)
        c_np, total_np = N.decode(t, 0.5)
        assert total == total_np and c_or.tobytes() == c_np.tobytes()
        for thr in (0.0, 0.3, 0.5, 1.0):
            assert O.nms_classwise(c_or, thr, 500).tobytes() == N.nms_classwise(c_np, thr, 500).tobytes()


def test_nms_statement_on_oracle_candidates_all_modes():
    """The NMS / IoU / order statement alone, on the oracle's own candidate lists in every activation mode (so that
    exp rounding cannot enter): crowds, ties, NaN scores, zero-area boxes."""
    for case, P, C, act, objects, dups, thr, iou, cap, max_det, agnostic, n in random_configs(24):
        t = synth.logits_tensor(9000 + 10 * case, num_proposals=P, num_classes=C, objects=objects, dups=dups,
                                adversarial=(P >= 64), as_logit=(act != 0))
        c, _ = O.decode(t, thr, act, cap=cap)
        assert O.nms_classwise(c, iou, max_det).tobytes() == N.nms_classwise(c, iou, max_det).tobytes(), case
    rng = np.random.default_rng(3)
    for k in (1, 65, 700, 3000):                        # crowds of few classes, many exactly equal scores
        c = np.zeros(k, N.REC)
        c["index"] = np.sort(rng.choice(25200, k, replace=False))
        cx, cy = rng.normal(320, 60, k), rng.normal(320, 60, k)
        w, h = rng.uniform(0, 200, k), rng.uniform(0, 200, k)
        c["x"], c["y"], c["w"], c["h"] = (cx - w / 2).astype(np.int32), (cy - h / 2).astype(np.int32), w.astype(np.int32), h.astype(np.int32)
        c["confidence"] = rng.choice(np.linspace(0.5, 1.0, 31, dtype=np.float32), k)
        c["class_id"] = rng.integers(0, 3, k)
        assert O.nms_classwise(c, 0.5, 500).tobytes() == N.nms_classwise(c, 0.5, 500).tobytes(), k


def test_iou_statement_edge_cases():
    def rec(x, y, w, h):
        r = np.zeros(1, N.REC); r["x"], r["y"], r["w"], r["h"] = x, y, w, h
        return r
    cases = [((0, 0, 10, 10), (0, 0, 10, 10)), ((0, 0, 10, 10), (10, 0, 10, 10)), ((0, 0, 10, 10), (11, 0, 10, 10)),
             ((100, 100, 100, 100), (100, 125, 100, 50)), ((5, 5, 0, 0), (5, 5, 0, 0)), ((-50, -7, 10, 4), (-45, -6, 10, 4)),
             ((2147483647, 0, 5, 5), (2147483640, 0, 10, 5)), ((0, 0, -5, 10), (0, 0, 10, 10)), ((0, 0, 65536, 65536), (0, 0, 65536, 65536))]
    for a, b in cases:
        ra, rb = rec(*a), rec(*b)
        got = N.iou_int(ra[0], rb)[0]
        want = np.float32(O.iou(ra[0], rb[0]))
        assert np.float32(got).view(np.uint32) == want.view(np.uint32), (a, b, got, want)


@pytest.mark.parametrize("act", [N.SIGMOID, N.SOFTMAX])
def test_activation_modes_within_libm_tolerance(act):
    """Scores through a libm-grade exp (NumPy, fp64 exp rounded to fp32) against the oracle's shared software expf (< 3 ulp of exp):
    <= 1e-5 relative; candidate rows, boxes and classes identical; kept indices identical wherever the two score
    lists sort the same way (a 1e-5 difference can only swap exact near-ties)."""
    same_order, frames = 0, 0
    for seed in range(6):
        t = synth.logits_tensor(300 + seed, adversarial=False, as_logit=True)
        c_or, total = O.decode(t, 0.5, act)
        c_np, total_np = N.decode(t, 0.5, act)
        # a row whose sigmoid(objectness) sits within 1e-6 of the threshold may flip; the generator keeps them away
        assert total == total_np
        for f in ("index", "x", "y", "w", "h", "class_id"):
            assert np.array_equal(c_or[f], c_np[f]), f
        rel = np.abs(c_or["confidence"] - c_np["confidence"]) / np.maximum(np.abs(c_or["confidence"]), 1e-30)
        assert rel.max() <= 1e-5, rel.max()
        frames += 1
        o1 = np.lexsort((c_or["index"], -N.total_order_key(c_or["confidence"])))
        o2 = np.lexsort((c_np["index"], -N.total_order_key(c_np["confidence"])))
        if np.array_equal(o1, o2):
            same_order += 1
            assert np.array_equal(O.nms_classwise(c_or, 0.5, 500)["index"], N.nms_classwise(c_np, 0.5, 500)["index"])
    assert same_order >= frames // 2


def test_shared_expf_against_libm():
    """The one helper oracle and product share: |tk_dm_expf(x) - exp(x)| < 3 ulp (measured 2.62) over the range decode
    uses — 3.6e-7 relative, far inside the 1e-5 score tolerance — and exact at 0."""
    xs = np.concatenate([np.linspace(-30, 30, 20001), np.array([-103.9, -87.3, -1e-8, 0.0, 1e-8, 88.7])]).astype(np.float32)
    got = np.array([O.det_expf(float(x)) for x in xs], np.float32)
    want64 = np.exp(xs.astype(np.float64))
    want = want64.astype(np.float32)
    ulp = np.maximum(np.spacing(np.abs(want)), np.float32(1e-45))
    assert np.all(np.abs(got.astype(np.float64) - want64) <= 3.0 * ulp.astype(np.float64) + 1e-300)
    assert O.det_expf(0.0) == 1.0
