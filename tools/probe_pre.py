"""GPU probe: preprocess kernel parity vs oracle + quick timing (dev tool)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from trackiellm_b200 import vision as V
from oracle import oracle as O

def check(w, h, ow, oh, dtype, arith, n=2, seed=0):
    rng = np.random.default_rng(seed)
    fr = rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8)
    d = torch.from_numpy(fr).cuda()
    out = V.preprocess_batch(d, ow, oh, out_dtype=dtype, arith=arith)
    torch.cuda.synchronize()
    got = out.float().cpu().numpy()
    worst = 0.0; exact = True
    for i in range(n):
        ref = O.preprocess_rgb8(fr[i], ow, oh)
        if dtype == torch.float16:
            ref16 = O.f32_to_f16(ref)
            exact &= np.array_equal(ref16.view(np.uint16), out[i].cpu().numpy().view(np.uint16))
            ref = ref16.astype(np.float32)
        else:
            exact &= np.array_equal(ref.view(np.uint32), got[i].view(np.uint32))
        worst = max(worst, float(np.max(np.abs(got[i] - ref) / (1e-5 + 1e-5*np.abs(ref)) if dtype==torch.float32 else np.abs(got[i]-ref))))
    print(f"{w}x{h}->{ow}x{oh} {dtype} arith={arith} n={n}: bit-exact={exact} worst={worst:.3g}", flush=True)

for (w,h) in [(640,480),(1920,1080),(333,257),(64,48)]:
    for dt in (torch.float32, torch.float16):
        for ar in (V.ARITH_EXACT, V.ARITH_FAST):
            check(w,h,640,640,dt,ar)
check(1920,1080,256,256,torch.float32,V.ARITH_EXACT)
check(1920,1080,641,333,torch.float32,V.ARITH_EXACT)
check(1920,1080,640,640,torch.float32,V.ARITH_EXACT,n=5,seed=3)

# timing
for dt, ar in [(torch.float16, V.ARITH_FAST), (torch.float16, V.ARITH_EXACT), (torch.float32, V.ARITH_EXACT), (torch.float32, V.ARITH_FAST)]:
    n = 64
    d = torch.randint(0, 256, (n,1080,1920,3), dtype=torch.uint8, device="cuda")
    out = torch.empty((n,3,640,640), dtype=dt, device="cuda")
    for _ in range(3): V.preprocess_batch(d,640,640,out_dtype=dt,arith=ar,out=out)
    torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    iters=20
    for _ in range(iters): V.preprocess_batch(d,640,640,out_dtype=dt,arith=ar,out=out)
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/iters
    byts = n*(1920*1080*3 + 640*640*3*out.element_size())
    print(f"timing {dt} arith={ar}: {ms:.3f} ms/batch64  {n/ms*1e3:.0f} frames/s  {byts/ms/1e6:.0f} GB/s  frac={byts/ms/1e6/6553.6:.3f}", flush=True)
