"""Copy-only ceiling of the end-to-end path at N ranks: every rank uploads its step's frames (32 x 8.3 MB pinned buffers,
one cudaMemcpyAsync per frame, as the library does) at the same time, no kernels.  Run under torchrun; rank 0 prints the
aggregate.  The e2e frames/s of bench.py at N GPUs cannot exceed aggregate_GB/s / 8.29 MB per 1080p BGRA frame."""
import json
import os
import sys

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nb = 8294400
xs = [torch.empty(nb, dtype=torch.uint8).pin_memory() for _ in range(32)]
for x in xs:
    x.fill_(rank + 1)                       # first touch
ys = [torch.empty(nb, dtype=torch.uint8, device="cuda") for _ in range(32)]
for a, b in zip(xs, ys):
    b.copy_(a, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
reps = 20
e0.record()
for _ in range(reps):
    for a, b in zip(xs, ys):
        b.copy_(a, non_blocking=True)
e1.record(); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
gbs_rank = reps * 32 * nb / (e0.elapsed_time(e1) * 1e-3) / 1e9
agg = world * reps * 32 * nb / (float(ms.item()) * 1e-3) / 1e9
if rank == 0:
    print(json.dumps({"ranks": world, "h2d_GBps_rank0": gbs_rank, "h2d_GBps_aggregate_max_over_ranks": agg,
                      "fps_ceiling_1080p_bgra": agg * 1e9 / nb, "host_cpus": len(os.sched_getaffinity(0))}))
if world > 1:
    dist.destroy_process_group()
