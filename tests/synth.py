"""Seeded synthetic inputs shared by tests, smoke() and bench.py.

All generators use numpy default_rng(seed) (PCG64) so the oracle and the GPU see
identical bytes.  Nothing here touches /root/reference.
"""
from __future__ import annotations

import numpy as np

NUM_PROPOSALS = 25200      # YOLOv5n 640x640 head
NUM_CLASSES = 80           # COCO, src/vision/tk_vision_pipeline.c:368-378


def rgb_frame(w: int, h: int, seed: int, kind: str = "random") -> np.ndarray:
    """(h, w, 3) u8.  kinds: random | zeros (tests/tk_vision_full_test.c:30-32) | ramp | white."""
    if kind == "zeros":
        return np.zeros((h, w, 3), np.uint8)
    if kind == "white":
        return np.full((h, w, 3), 255, np.uint8)
    if kind == "ramp":
        y, x, c = np.meshgrid(np.arange(h), np.arange(w), np.arange(3), indexing="ij")
        return ((x + 2 * y + 85 * c) & 255).astype(np.uint8)
    return np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)


def i420_frame(w: int, h: int, seed: int) -> np.ndarray:
    """Flat I420 buffer: Y w*h, U (w/2)*(h/2), V (w/2)*(h/2) (src/ffi/c_api/tk_ffi_api.c:917-931)."""
    n = w * h + 2 * (w // 2) * (h // 2)
    return np.random.default_rng(seed).integers(0, 256, n, dtype=np.uint8)


def logits_tensor(seed: int, num_proposals: int = NUM_PROPOSALS, num_classes: int = NUM_CLASSES,
                  objects: int = 40, dups: int = 12, adversarial: bool = False, as_logit: bool = False) -> np.ndarray:
    """(P, 5+C) f32 detector-head output: `objects` true boxes, each emitting `dups` jittered
    near-duplicates with high objectness; every other row has objectness < 0.3."""
    rng = np.random.default_rng(seed)
    p, c = num_proposals, num_classes
    t = np.empty((p, 5 + c), np.float32)
    t[:, 0:2] = rng.uniform(0, 640, (p, 2))
    t[:, 2:4] = rng.uniform(4, 200, (p, 2))
    t[:, 4] = rng.uniform(0, 0.3, p)
    t[:, 5:] = rng.uniform(0, 0.2, (p, c))
    n_obj = min(objects, max(p // max(dups, 1), 0))
    rows = rng.permutation(p)[: n_obj * dups].reshape(n_obj, dups) if n_obj else np.empty((0, dups), int)
    for o in range(n_obj):
        cx, cy = rng.uniform(40, 600, 2)
        w, h = np.exp(rng.uniform(np.log(16), np.log(320), 2))
        k = int(rng.integers(0, c)) if c else 0
        for r in rows[o]:
            t[r, 0:4] = np.array([cx, cy, w, h]) + rng.normal(0, 3, 4)
            t[r, 4] = rng.uniform(0.55, 0.99)
            if c:
                t[r, 5 + k] = rng.uniform(0.6, 0.99)
    if adversarial and p >= 64:
        a = p // 2
        # exact duplicate rows (tie-break by index), same class
        t[a] = t[a + 1] = t[a + 2]
        t[a:a + 3, 4] = 0.9
        # objectness exactly at the threshold (strict >): must be rejected at thr 0.5
        t[a + 3, 4] = 0.5
        # IoU exactly 0.5 pair: boxes 100x100 at x=100 and 100x50-overlap -> inter 5000, union 10000 -> wait for strict >
        t[a + 4, 0:5] = [150, 150, 100, 100, 0.95]
        t[a + 5, 0:5] = [150, 175, 100, 50, 0.94]        # contained: inter 5000 / union 10000 = 0.5 exactly
        t[a + 4, 5:] = 0.01; t[a + 5, 5:] = 0.01
        if c:
            t[a + 4, 5] = 0.9; t[a + 5, 5] = 0.9
        # zero-area boxes (union == 0 rule) and negative corner (truncation toward zero)
        t[a + 6, 0:5] = [300, 300, 0.4, 0.4, 0.93]
        t[a + 7, 0:5] = [300, 300, 0.4, 0.4, 0.92]
        t[a + 8, 0:5] = [2.5, 3.5, 20.7, 30.9, 0.91]
        t[a + 9, 0:5] = [-50.5, -7.25, 10.5, 4.5, 0.9]
        # NaN / inf hygiene
        t[a + 10, 4] = np.nan
        t[a + 11, 4] = 0.97; t[a + 11, 5] = np.nan          # NaN at class 0 sticks (sequential first-max)
        t[a + 12, 4] = 0.96
        if c > 3:
            t[a + 12, 5 + 3] = np.nan                        # NaN elsewhere is never selected
        t[a + 13, 0:5] = [1e12, -1e12, 3e9, 5, 0.955]        # saturating float->int
        # equal class scores: first maximum wins
        t[a + 14, 4] = 0.9
        if c > 7:
            t[a + 14, 5:] = 0.1; t[a + 14, 5 + 2] = 0.8; t[a + 14, 5 + 7] = 0.8
    if as_logit:
        with np.errstate(divide="ignore", invalid="ignore"):
            q = np.clip(t[:, 4:].astype(np.float64), 1e-6, 1 - 1e-6)
            lg = np.log(q / (1 - q)).astype(np.float32)
            lg[np.isnan(t[:, 4:])] = np.nan
            t[:, 4:] = lg
    return t
