"""Second, independent statement of the detector post-process in NumPy — TEST INFRASTRUCTURE ONLY.

Written from SURVEY.md Appendix A.2 / A.3 and the reference lines they cite, NOT from oracle/tk_oracle.c, so that the C
oracle (which the kernels are checked against) has a cross-check that does not share its author's helpers:

  decode   src/vision/tk_object_detector.c:343-351  row pointer, strict `row[4] > thr`, C-truncated box, then the fixed
           choices of SURVEY §8(c): class = first maximum of the class logits, confidence = act(obj) * act(cls)
  IoU      src/vision/src/object_analysis.rs:71-91   integer rects, 0 when disjoint or union <= 0
  NMS      src/gpu/metal/tk_metal_kernels.metal:864-875 made sequential and class-wise: order (confidence desc,
           proposal index asc), a box is dropped when a KEPT earlier box of its class has IoU > thr (strict)

Activations use NumPy's float64 exp rounded to fp32 (not the shared software expf): sigmoid / softmax scores agree with
the oracle to ~1e-6 relative, not bit for bit — see tests/test_post_crosscheck.py for the tolerance contract.
"""
from __future__ import annotations

import numpy as np

REC = np.dtype([("index", "<i4"), ("x", "<i4"), ("y", "<i4"), ("w", "<i4"), ("h", "<i4"),
                ("confidence", "<f4"), ("class_id", "<u4")])
RAW, SIGMOID, SOFTMAX = 0, 1, 2


def c_int_cast(v: np.ndarray) -> np.ndarray:
    """(int)v of tk_object_detector.c:347-350: truncation toward zero; NaN -> 0 and saturation where C leaves it undefined."""
    v = np.asarray(v, np.float32).astype(np.float64)
    t = np.trunc(np.nan_to_num(v, nan=0.0, posinf=2.0**31, neginf=-2.0**31))
    return np.clip(t, -2.0**31, 2.0**31 - 1).astype(np.int64).astype(np.int32)


def sigmoid32(v: np.ndarray) -> np.ndarray:
    v = np.asarray(v, np.float32)
    with np.errstate(over="ignore", invalid="ignore"):
        e = np.exp(-v.astype(np.float64)).astype(np.float32)                 # exp rounded to fp32, as a libm expf would give
        return (np.float32(1.0) / (np.float32(1.0) + e)).astype(np.float32)


def decode(logits: np.ndarray, thr: float, act: int = RAW, cap: int | None = None):
    """-> (records of the rows that pass, in ascending row order, at most `cap`; number of rows that pass)."""
    t = np.asarray(logits, np.float32)
    P, L = t.shape
    C = L - 5
    obj = t[:, 4] if act == RAW else sigmoid32(t[:, 4])
    with np.errstate(invalid="ignore"):
        rows = np.nonzero(obj > np.float32(thr))[0]                           # strict; NaN never passes
    total = len(rows)
    if cap is not None:
        rows = rows[:cap]
    out = np.zeros(len(rows), REC)
    out["index"] = rows
    r = t[rows]
    half_w = (r[:, 2] / np.float32(2)).astype(np.float32)
    half_h = (r[:, 3] / np.float32(2)).astype(np.float32)
    with np.errstate(invalid="ignore", over="ignore"):
        out["x"] = c_int_cast((r[:, 0] - half_w).astype(np.float32))
        out["y"] = c_int_cast((r[:, 1] - half_h).astype(np.float32))
    out["w"] = c_int_cast(r[:, 2])
    out["h"] = c_int_cast(r[:, 3])
    if C == 0:
        cls_score = np.ones(len(rows), np.float32)
        best = np.zeros(len(rows), np.int64)
    else:
        cl = r[:, 5:]
        # first maximum of a sequential scan `if (v[j] > v[best]) best = j`: a NaN at j >= 1 never wins, a NaN at 0 is
        # never displaced; -0 == +0
        best = np.zeros(len(rows), np.int64)
        cur = cl[:, 0].copy()
        with np.errstate(invalid="ignore"):
            for j in range(1, C):
                win = cl[:, j] > cur
                best[win] = j
                cur[win] = cl[win, j]
        bv = cl[np.arange(len(rows)), best]
        if act == RAW:
            cls_score = bv
        elif act == SIGMOID:
            cls_score = sigmoid32(bv)
        else:   # softmax over the class logits, max-subtracted, summed in index order (tests/tk_gpu_softmax_test.cpp:31-49)
            with np.errstate(invalid="ignore", over="ignore"):
                e = np.exp((cl - bv[:, None]).astype(np.float32).astype(np.float64)).astype(np.float32)
                s = np.zeros(len(rows), np.float32)
                for j in range(C):
                    s = (s + e[:, j]).astype(np.float32)
                cls_score = (np.float32(1.0) * np.exp(np.zeros(len(rows))).astype(np.float32) / s).astype(np.float32)
                cls_score = np.where(np.isnan(bv), np.float32(np.nan), cls_score).astype(np.float32)
    with np.errstate(invalid="ignore"):
        conf = (obj[rows].astype(np.float32) * cls_score.astype(np.float32)).astype(np.float32)
    cb = conf.view(np.uint32).copy()
    cb[np.isnan(conf)] = 0x7FC00000                                           # one NaN pattern
    out["confidence"] = cb.view(np.float32)
    out["class_id"] = best
    return out, total


def iou_int(a, bs: np.ndarray) -> np.ndarray:
    """IoU of record `a` against records `bs` (object_analysis.rs:71-91; i32 arithmetic wraps as release-mode Rust)."""
    i32 = np.int32
    with np.errstate(over="ignore"):
        ax, ay, aw, ah = (np.full(len(bs), a[k], i32) for k in ("x", "y", "w", "h"))
        bx, by, bw, bh = (bs[k].astype(i32) for k in ("x", "y", "w", "h"))
        x_left, y_top = np.maximum(ax, bx), np.maximum(ay, by)
        x_right, y_bottom = np.minimum(ax + aw, bx + bw), np.minimum(ay + ah, by + bh)
        disjoint = (x_right < x_left) | (y_bottom < y_top)
        inter = (x_right - x_left).astype(np.float32) * (y_bottom - y_top).astype(np.float32)
        area_a, area_b = (aw * ah).astype(np.float32), (bw * bh).astype(np.float32)
        union = ((area_a + area_b).astype(np.float32) - inter).astype(np.float32)
        with np.errstate(divide="ignore", invalid="ignore"):
            v = np.where(union > 0, inter / union, np.float32(0)).astype(np.float32)
    return np.where(disjoint, np.float32(0), v)


def total_order_key(f: np.ndarray) -> np.ndarray:
    """IEEE-754 totalOrder as int64 keys: -NaN < -inf < ... < -0 < +0 < ... < +inf < +NaN."""
    i = np.asarray(f, np.float32).view(np.int32).astype(np.int64)
    return np.where(i < 0, -(i & 0x7FFFFFFF) - 1, i)


def nms_classwise(cands: np.ndarray, iou_thr: float, max_det: int, class_agnostic: bool = False) -> np.ndarray:
    if len(cands) == 0 or max_det == 0:
        return cands[:0].copy()
    order = np.lexsort((cands["index"], -total_order_key(cands["confidence"])))   # last key is primary
    s = cands[order]
    kept: list[int] = []
    thr = np.float32(iou_thr)
    for k in range(len(s)):
        if kept:
            ks = s[kept]
            same = np.ones(len(ks), bool) if class_agnostic else (ks["class_id"] == s[k]["class_id"])
            if same.any() and np.any(iou_int(s[k], ks[same]) > thr):
                continue
        kept.append(k)
    return s[kept[:max_det]].copy()
