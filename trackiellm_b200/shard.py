"""Frame sharding across the GPUs of one box: one process per GPU, no data-path collective.

Frames (pre-process) and logit tensors (post-process) are independent units — the reference has no
cross-frame state on this path (its only stateful piece, the Rust tracker
src/vision/src/object_analysis.rs:49-51, is downstream).  Rank g owns the contiguous block
[g*B/G, (g+1)*B/G) of a batch, or camera stream s -> rank s % G (BASELINE config 4), runs its own
tk_b200_pipeline on its own device, and the only exchange is a final HOST gather of the per-frame
detection lists (torch.distributed over gloo; NCCL is not used — nothing crosses NVLink).
"""
from __future__ import annotations

import numpy as np

CAND_DTYPE = np.dtype([("index", "<i4"), ("x", "<i4"), ("y", "<i4"), ("w", "<i4"), ("h", "<i4"),
                       ("confidence", "<f4"), ("class_id", "<u4")])


def shard_range(num_units: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block of `num_units` owned by `rank`: [rank*B/G, (rank+1)*B/G) in integer arithmetic.
    Blocks tile [0, B) exactly and differ in size by at most one."""
    if world <= 0 or not (0 <= rank < world) or num_units < 0:
        raise ValueError("bad shard arguments")
    return (num_units * rank) // world, (num_units * (rank + 1)) // world


def stream_owner(stream_id: int, world: int) -> int:
    """Camera stream -> rank, preserving per-stream frame order on one device."""
    return stream_id % world


def pack_detections(dets: np.ndarray, counts: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """[n, max_det] padded records + [n] counts -> (flat records of the kept detections, counts)."""
    dets = np.ascontiguousarray(dets)
    if dets.dtype != CAND_DTYPE:
        dets = dets.reshape(dets.shape[0], -1, 7).view(CAND_DTYPE).reshape(dets.shape[0], -1)
    counts = np.asarray(counts, dtype=np.int64)
    cap = dets.shape[1] if dets.ndim == 2 else 0
    flat = [dets[i, :min(int(c), cap)] for i, c in enumerate(counts)]
    return (np.concatenate(flat) if flat else np.zeros(0, CAND_DTYPE)), np.minimum(counts, cap)


def gather_detections(local_dets: np.ndarray, local_counts: np.ndarray, group=None, dst: int = 0):
    """Final host gather.  Every rank passes the padded detections of ITS block; `dst` receives
    (records of all frames in global frame order, offsets[B+1]) — frame f's detections are
    records[offsets[f]:offsets[f+1]] — and the other ranks receive None.  Works without
    torch.distributed initialised (single process)."""
    import torch.distributed as dist
    flat, counts = pack_detections(local_dets, local_counts)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return flat, np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    payload = (flat.tobytes(), counts.tolist())
    gathered = [None] * world if rank == dst else None
    dist.gather_object(payload, gathered, dst=dst, group=group)
    if rank != dst:
        return None
    recs = [np.frombuffer(b, dtype=CAND_DTYPE) for b, _ in gathered]     # blocks are contiguous in rank order
    cnts = np.concatenate([np.asarray(c, dtype=np.int64) for _, c in gathered]) if gathered else np.zeros(0, np.int64)
    return np.concatenate(recs), np.concatenate([[0], np.cumsum(cnts)]).astype(np.int64)


class ShardedPipeline:
    """This rank's slice of a batched job: wraps vision.Pipeline on the rank's device."""

    def __init__(self, rank: int, world: int, local_device: int, **pipeline_kwargs):
        from . import vision as V
        self.rank, self.world = rank, world
        self.pipe = V.Pipeline(device_id=local_device, **pipeline_kwargs)

    def close(self) -> None:
        self.pipe.close()

    def process_host_block(self, frames_all, logits_all, group=None, dst: int = 0):
        """frames_all / logits_all: the whole job's pinned host batch (every rank sees the same arrays, e.g. a
        shared-memory ring filled by the capture threads); this rank processes its block and the detections
        are gathered on `dst`."""
        import torch
        n = frames_all.shape[0] if frames_all is not None else logits_all.shape[0]
        lo, hi = shard_range(n, self.rank, self.world)
        _, dets, counts = self.pipe.alloc_outputs(max(hi - lo, 1), pinned=True)
        if hi > lo:
            self.pipe.process_host(frames_all[lo:hi] if frames_all is not None else None,
                                   logits_all[lo:hi] if logits_all is not None else None, dets, counts)
        d = dets.numpy()[:hi - lo] if isinstance(dets, torch.Tensor) else dets[:hi - lo]
        c = counts.numpy()[:hi - lo] if isinstance(counts, torch.Tensor) else counts[:hi - lo]
        return gather_detections(d, c, group=group, dst=dst)
