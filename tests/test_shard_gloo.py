"""Multi-process host logic of the frame-sharded path on CPU: world_size 2, gloo.  The detector is replaced by a
deterministic stand-in (the CUDA path needs a GPU); what is tested is the partition and the host gather."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from opencv_b200.shard import detect_sharded, gather_detections, shard_indices


def fake_rects(frame):
    n = int(frame.sum()) % 5
    return np.array([[int(frame[0]) + k, k, 64, 128] for k in range(n)], np.int32).reshape(-1, 4)


def _worker(rank, world, port, n_frames, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    frames = [np.array([i, 3 * i + 1], np.int64) for i in range(n_frames)]
    local = detect_sharded(lambda fs: [fake_rects(f) for f in fs], frames, rank, world)
    assert sorted(local) == shard_indices(n_frames, rank, world)
    allr = gather_detections(local, n_frames, rank, world)
    if rank == 0:
        q.put([r.tolist() for r in allr])
    else:
        assert allr is None
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def test_shard_indices_cover_once():
    for n in [0, 1, 7, 512]:
        for w in [1, 2, 4, 8]:
            got = sorted(i for r in range(w) for i in shard_indices(n, r, w))
            assert got == list(range(n))


def test_gather_world2_gloo():
    n_frames, world = 7, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    exp = [fake_rects(np.array([i, 3 * i + 1], np.int64)).tolist() for i in range(n_frames)]
    assert got == exp


def test_gather_world1():
    frames = [np.array([i, i], np.int64) for i in range(5)]
    local = detect_sharded(lambda fs: [fake_rects(f) for f in fs], frames, 0, 1)
    out = gather_detections(local, 5)
    assert [r.tolist() for r in out] == [fake_rects(f).tolist() for f in frames]
