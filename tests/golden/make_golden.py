"""Generate tests/golden/golden.npz from the reference's OWN compiled code (oracle/_ref).

Run in the build container (needs /root/reference to have been compiled by
`make -C oracle`); the GPU box only reads the committed .npz.  Inputs are seeded
so the script is reproducible:  python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import synth  # noqa: E402
from oracle import oracle as O  # noqa: E402

assert O.have_ref() and O.have_ref_softmax(), "build oracle/_ref first (make -C oracle)"
g = {}

# (a2) tk_preprocessor_resize_and_normalize_to_chw — small cases stored whole
for name, (w, h, ow, oh, seed, kind) in {
    "pre_small_a": (64, 48, 40, 40, 11, "random"),
    "pre_small_b": (33, 17, 20, 12, 12, "random"),
    "pre_small_ramp": (96, 54, 32, 32, 0, "ramp"),
}.items():
    fr = synth.rgb_frame(w, h, seed, kind)
    g[name + "_in"] = fr
    g[name + "_out"] = O.ref_preprocess(fr, ow, oh)

# ... and BASELINE config 1 / 2 sizes stored as SHA-256 of the fp32 bytes
for name, (w, h, seed, kind) in {
    "pre_cfg1_zeros": (640, 480, 0, "zeros"),       # tests/tk_vision_full_test.c:27-34
    "pre_cfg1_random": (640, 480, 0, "random"),
    "pre_cfg1_ramp": (640, 480, 0, "ramp"),
    "pre_cfg1_white": (640, 480, 0, "white"),
    "pre_1080p_random": (1920, 1080, 7, "random"),
}.items():
    out = O.ref_preprocess(synth.rgb_frame(w, h, seed, kind), 640, 640)
    g[name + "_sha256"] = np.frombuffer(hashlib.sha256(out.tobytes()).digest(), np.uint8)

# (a10) softmax_cpu on the reference test's own input (tests/tk_gpu_softmax_test.cpp:68-70)
x = (np.arange(4 * 256) % 256).astype(np.float32).reshape(4, 256)
g["softmax_kat_in"] = x
g["softmax_kat_out"] = O.ref_softmax(x)
x80 = np.random.default_rng(5).normal(0, 3, (64, 80)).astype(np.float32)
g["softmax_80_in"] = x80
g["softmax_80_out"] = O.ref_softmax(x80)

# (a11) tk_classify_dominant_color / tk_classify_door_state
COLORS = list(O.COLOR_NAMES)
fr = synth.rgb_frame(160, 120, 21)
boxes = np.array([[0, 0, 160, 120], [10, 10, 80, 80], [-20, -20, 60, 60], [100, 60, 100, 100], [50, 50, 1, 1]], np.int32)
g["color_frame"] = fr
g["color_boxes"] = boxes
g["color_ids"] = np.array([COLORS.index(O.ref_dominant_color(fr, b)) for b in boxes], np.int32)
solid = {"red": (255, 0, 0), "yellow": (255, 255, 0), "green": (0, 255, 0), "cyan": (0, 255, 255), "blue": (0, 0, 255),
         "magenta": (255, 0, 255), "black": (5, 5, 5), "white": (250, 250, 250), "gray": (128, 128, 128),
         "orange": (255, 120, 0), "pink": (255, 100, 180), "brown": (120, 70, 20)}
g["solid_rgb"] = np.array(list(solid.values()), np.uint8)
ids = []
for rgb in solid.values():
    f = np.empty((20, 20, 3), np.uint8); f[:] = rgb
    ids.append(COLORS.index(O.ref_dominant_color(f, (0, 0, 20, 20))))
g["solid_ids"] = np.array(ids, np.int32)
# door: tests/tk_attribute_classifier_test.c:56-93 (grey frame, two 5-px black bars) + an open one
door = np.full((100, 100, 3), 128, np.uint8)
door[:, 30:35] = 0; door[:, 60:65] = 0
g["door_frame"] = door
g["door_box"] = np.array([10, 10, 80, 80], np.int32)
g["door_closed"] = np.array([O.ref_door_state(door, (10, 10, 80, 80)) == "closed"])
stripes = np.full((100, 100, 3), 128, np.uint8)
stripes[::4] = 255; stripes[1::4] = 255; stripes[2::4] = 0; stripes[3::4] = 0
g["door2_frame"] = stripes
g["door2_closed"] = np.array([O.ref_door_state(stripes, (10, 10, 80, 80)) == "closed"])

np.savez_compressed(os.path.join(HERE, "golden.npz"), **g)
print("wrote golden.npz with", len(g), "arrays;", os.path.getsize(os.path.join(HERE, "golden.npz")), "bytes")
print("door:", g["door_closed"], g["door2_closed"], "solid:", [COLORS[i] for i in ids], "boxes:", g["color_ids"])
