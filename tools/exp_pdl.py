"""Experiment: step time of the bench workload with NO stage events between the kernels (151.0 us per 64-frame step
on B200 against 163 us with the four event records per step), and a checksum of tensor + detections + counts.
It was also the A/B harness for decode launched as a programmatic dependent of the pre-process kernel (a library
build that read TK_B200_PDL; 157.8 us, same checksum, dropped - DESIGN 9.1); the variable is ignored now.
TK_B200_PDL_PRE=0 turns off the pre-process kernel's early launch behind the previous step's NMS (150.9 vs 147.6 us)."""
import os, sys, zlib, torch
sys.path.insert(0, ".")
import bench as Bn
from trackiellm_b200 import vision as V
B = 64
fr, lg = Bn.make_inputs(B, 0)
d_frames = torch.from_numpy(fr).cuda(); d_logits = torch.from_numpy(lg).cuda()
pipe = V.Pipeline(device_id=0, max_batch=B, chunk_frames=8, width=Bn.SRC_W, height=Bn.SRC_H, fmt=V.PIXEL_RGB8,
                  tensor_w=Bn.DST, tensor_h=Bn.DST, out_dtype=torch.float16, num_proposals=Bn.P, num_classes=Bn.NC,
                  act=V.ACT_RAW, conf_thr=Bn.CONF_THR, iou_thr=Bn.IOU_THR, max_detections=Bn.MAX_DET)
d_tensor, d_dets, d_counts = pipe.alloc_outputs(B)
for _ in range(10):
    pipe.process_device(d_frames, d_logits, d_tensor, d_dets, d_counts)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for _ in range(5):
    torch.cuda.synchronize(); e0.record()
    for _ in range(200):
        pipe.process_device(d_frames, d_logits, d_tensor, d_dets, d_counts)
    e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) * 5)
crc = zlib.crc32(d_dets.cpu().numpy().tobytes() + d_counts.cpu().numpy().tobytes() + d_tensor.cpu().numpy().tobytes())
print(f"PDL_PRE={os.environ.get('TK_B200_PDL_PRE', '1')}: {best:.1f} us/step crc={crc:08x}", flush=True)
