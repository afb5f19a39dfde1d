"""Experiment: stage timings of the device-resident step under different L2 fetch granularities."""
import sys
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import bench
from trackiellm_b200 import _capi as capi, vision as V

B = 64
fr, lg = bench.make_inputs(bench.WORKLOADS["1080p-rgb"], 8, 0)
import numpy as np
fr = np.concatenate([fr] * 8); lg = np.concatenate([lg] * 8)
df, dl = torch.from_numpy(fr).cuda(), torch.from_numpy(lg).cuda()
pipe = V.Pipeline(max_batch=B)
t, d, c = pipe.alloc_outputs(B)
lib = capi.lib()
for g in (0, 32, 64, 128, 32):
    print("granularity ->", lib.tk_b200_l2_fetch_granularity(g))
    for _ in range(5):
        pipe.process_device(df, dl, t, d, c)
    torch.cuda.synchronize()
    pipe.stage_timing(100)
    for _ in range(100):
        pipe.process_device(df, dl, t, d, c)
    torch.cuda.synchronize()
    ms, n = pipe.stage_times()
    print("   us/launch: pre %.1f decode %.1f nms %.1f" % tuple(1e3 * m / n for m in ms))
