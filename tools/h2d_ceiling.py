"""Host->device copy ceiling of the box, alone and with every rank copying at once.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/h2d_ceiling.py

Every rank owns one pinned 295 MB buffer (the bench's heatmap batch) and copies it to its GPU with cudaMemcpyAsync,
20 times: first rank by rank (the others idle), then all ranks together after a barrier.  Rank 0 prints one JSON line:
GB/s per rank alone, GB/s per rank concurrently, and the aggregate -- the frames/s the host can feed at most is
aggregate / 9.216 MB per frame.  No batched-memcpy API is used.
"""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
if hasattr(os, "sched_setaffinity") and world > 1 and os.environ.get("PIN", "1") == "1":
    cpus = sorted(os.sched_getaffinity(0))
    per = max(1, len(cpus) // world)
    os.sched_setaffinity(0, cpus[local * per:(local + 1) * per] or cpus)
nbytes = 32 * 5 * 15 * 128 * 240 * 4
host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
host.fill_(1)
devbuf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
REPS = 20


def copy_gbs():
    torch.cuda.synchronize(dev)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(REPS):
        devbuf.copy_(host, non_blocking=True)
    e.record()
    torch.cuda.synchronize(dev)
    return nbytes * REPS / (s.elapsed_time(e) * 1e-3) / 1e9


def barrier():
    if world > 1:
        dist.barrier()


copy_gbs()
alone = 0.0
for r in range(world):
    barrier()
    if r == rank:
        alone = copy_gbs()
barrier()
together = copy_gbs()
vals = torch.tensor([alone, together], device=dev, dtype=torch.float64)
allv = [torch.zeros_like(vals) for _ in range(world)]
if world > 1:
    dist.all_gather(allv, vals)
else:
    allv = [vals]
if rank == 0:
    a = [float(v[0]) for v in allv]
    t = [float(v[1]) for v in allv]
    agg = sum(t)
    print(json.dumps({"check": "pinned host -> device copy ceiling, 295 MB buffers, cudaMemcpyAsync x 20", "n_gpus": world,
                      "cpus": len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else None,
                      "gbs_alone_per_rank": [round(x, 2) for x in a], "gbs_concurrent_per_rank": [round(x, 2) for x in t],
                      "gbs_concurrent_aggregate": round(agg, 1), "concurrent_over_alone": round(agg / max(sum(a), 1e-9), 3),
                      "frames_per_s_ceiling": round(agg * 1e9 / (5 * 15 * 128 * 240 * 4), 0)}))
if world > 1:
    dist.destroy_process_group()
