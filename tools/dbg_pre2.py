"""Debug: locate mismatching output rows of the pre-process kernel for a given geometry."""
import sys
import numpy as np, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import synth
from oracle import oracle as O
from trackiellm_b200 import vision as V
w, h, ow, oh, n = [int(x) for x in sys.argv[1:6]]
fr = np.stack([synth.rgb_frame(w, h, i) for i in range(n)])
d = torch.from_numpy(fr).cuda()
ref = np.stack([O.preprocess_rgb8(fr[i], ow, oh) for i in range(n)])
for rep in range(6):
    out = V.preprocess_batch(d, ow, oh)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    bad = np.argwhere((got.view(np.uint32) != ref.view(np.uint32)).any(axis=(1, 3)))   # (frame, row)
    print("rep", rep, "bad rows:", len(bad), [tuple(b) for b in bad[:12]])
