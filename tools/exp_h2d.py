"""Experiment (run under torchrun, one rank per GPU): aggregate host-to-device bandwidth when every rank copies
400 MB from its own pinned buffer at the same time.  (A write-combined allocation was tried as well: no difference.)"""
import ctypes as C, os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, ".")
from trackiellm_b200 import _capi as capi

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("gloo", rank=rank, world_size=world)
lib = capi.lib()
N = 400 * 1024 * 1024
dev = torch.empty(N, dtype=torch.uint8, device="cuda")
for name, fn in (("pinned", lib.tk_b200_host_alloc),):
    p = C.c_void_p()
    assert fn(C.byref(p), C.c_size_t(N)) == 0
    C.memset(p, 1, N)
    host = torch.from_numpy(np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(N,)))
    for _ in range(2):
        dev.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(10):
        dev.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gbs = torch.tensor([10 * N / dt / 1e9], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(gbs)
    if rank == 0:
        print(f"{world} GPUs, {name}: aggregate H2D {float(gbs[0]):.1f} GB/s ({float(gbs[0]) / world:.1f} per GPU)", flush=True)
    del host
    lib.tk_b200_host_free(p)
