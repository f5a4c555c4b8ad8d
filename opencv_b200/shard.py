"""Frame sharding across GPUs: frames are independent, so frame i goes to rank i mod G,
every rank runs its own detector handle on its own GPU, and the only exchange is a HOST gather of the
variable-length box lists, in frame order.  There is no collective on the data path.

The reference's closest analogue is samples/gpu/multi.cpp:27-68 (one host thread per device).
"""
import numpy as np


def shard_indices(n_frames, rank, world):
    """Indices of the frames rank `rank` processes (round robin, as SURVEY 8e)."""
    return list(range(rank, n_frames, world))


def detect_sharded(detect_batch, frames, rank, world):
    """Runs `detect_batch(list_of_frames) -> list of (N_i, 4) int32 arrays` on this rank's shard.
    `frames` is the full sequence (any indexable); only the shard is touched.  -> {frame index: rects}"""
    idx = shard_indices(len(frames), rank, world)
    out = detect_batch([frames[i] for i in idx]) if idx else []
    assert len(out) == len(idx)
    return {i: np.asarray(r, np.int32).reshape(-1, 4) for i, r in zip(idx, out)}


def gather_detections(local, n_frames, rank=0, world=1, dst=0, group=None):
    """Host gather of the per-frame box lists onto rank `dst` (torch.distributed.gather_object: pickled over the
    process group's CPU path, a few KB per frame; `group` = a gloo group when the default group is NCCL).
    Returns the list in frame order on `dst`, None elsewhere."""
    if world == 1:
        return [local[i] for i in range(n_frames)]
    import torch.distributed as dist
    parts = [None] * world if rank == dst else None
    dist.gather_object(local, parts, dst=dst, group=group)
    if rank != dst:
        return None
    merged = {}
    for p in parts:
        merged.update(p)
    assert sorted(merged) == list(range(n_frames)), "a frame is missing or duplicated"
    return [merged[i] for i in range(n_frames)]
