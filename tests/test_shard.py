"""Multi-GPU host logic on CPU: frame sharding and the final host gather over gloo, world_size 2 and 3.
The per-rank "device output" is produced by the oracle here (no GPU in this container); the GPU
version of the same check is test_gpu_sharded_equals_single below."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import synth
from oracle import oracle as O
from trackiellm_b200 import shard as S

MAX_DET = 50
P = 600


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _oracle_block(seeds):
    """Padded [n, MAX_DET] detections + counts for the frames with these seeds (what a rank's GPU writes)."""
    dets = np.zeros((len(seeds), MAX_DET), dtype=S.CAND_DTYPE)
    counts = np.zeros(len(seeds), dtype=np.int32)
    for k, s in enumerate(seeds):
        t = synth.logits_tensor(s, num_proposals=P, objects=6, dups=5)
        c, _ = O.decode(t, 0.5, cap=4096)
        d = O.nms_classwise(c, 0.5, MAX_DET).view(S.CAND_DTYPE)
        dets[k, :len(d)] = d
        counts[k] = len(d)
    return dets, counts


def test_shard_range_tiles_exactly():
    for n in (0, 1, 7, 64, 256, 1000):
        for world in (1, 2, 3, 4, 8):
            blocks = [S.shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        S.shard_range(4, 2, 2)
    assert [S.stream_owner(s, 8) for s in range(10)] == [0, 1, 2, 3, 4, 5, 6, 7, 0, 1]


def test_pack_and_single_process_gather():
    dets, counts = _oracle_block([1, 2, 3])
    counts[1] = MAX_DET + 7                       # a count above the cap is clamped, as the kernels clamp it
    flat, offs = S.gather_detections(dets, counts)
    assert offs[0] == 0 and offs[-1] == len(flat) and len(offs) == 4
    assert offs[2] - offs[1] == MAX_DET
    # int32 [n, max_det, 7] words (the device layout) are accepted as well
    words = dets.view(np.int32).reshape(3, MAX_DET, 7)
    flat2, offs2 = S.gather_detections(words, counts)
    assert flat2.tobytes() == flat.tobytes() and np.array_equal(offs, offs2)


def _worker(rank, world, port, n_frames, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = S.shard_range(n_frames, rank, world)
        dets, counts = _oracle_block(list(range(100 + lo, 100 + hi)))
        res = S.gather_detections(dets, counts)
        if rank == 0:
            flat, offs = res
            np.savez(out_path, flat=flat.view(np.uint8), offs=offs)
        else:
            assert res is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_frames", [(2, 7), (3, 5), (2, 1)])
def test_gloo_gather_equals_single_process(tmp_path, world, n_frames):
    out = str(tmp_path / "gathered.npz")
    mp.spawn(_worker, args=(world, _free_port(), n_frames, out), nprocs=world, join=True)
    got = np.load(out)
    dets, counts = _oracle_block(list(range(100, 100 + n_frames)))
    flat, offs = S.pack_detections(dets, counts)
    assert got["flat"].tobytes() == flat.tobytes()
    assert np.array_equal(got["offs"], np.concatenate([[0], np.cumsum(offs)]))


@pytest.mark.gpu
def test_gpu_sharded_equals_single():
    """Two emulated ranks on one GPU (sequentially): each processes its block through its own pipeline
    handle; the union equals the single-pipeline result bit for bit."""
    from trackiellm_b200 import vision as V
    n, W, H = 9, 320, 240
    frames = np.stack([synth.rgb_frame(W, H, 50 + i) for i in range(n)])
    logits = np.stack([synth.logits_tensor(50 + i, num_proposals=P, objects=6, dups=5) for i in range(n)])
    hf, hl = torch.from_numpy(frames).pin_memory(), torch.from_numpy(logits).pin_memory()
    kw = dict(max_batch=n, width=W, height=H, tensor_w=64, tensor_h=64, num_proposals=P, num_classes=80,
              max_detections=MAX_DET)
    with V.Pipeline(**kw) as pipe:
        _, d, c = pipe.alloc_outputs(n, pinned=True)
        pipe.process_host(hf, hl, d, c)
        want_flat, want_counts = S.pack_detections(d.numpy(), c.numpy())
    parts = []
    for rank in range(2):
        sp = S.ShardedPipeline(rank, 2, 0, **kw)
        parts.append(sp.process_host_block(hf, hl))        # no process group: returns this block only
        sp.close()
    flat = np.concatenate([p[0] for p in parts])
    assert flat.tobytes() == want_flat.tobytes()
    assert sum(len(p[1]) - 1 for p in parts) == n
