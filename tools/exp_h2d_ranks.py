#!/usr/bin/env python
"""Kernel-free host->device copy test under torchrun: every rank copies the cfg2 step's source pictures (945 MB, pinned) to its
own GPU, all ranks at once, with and without binding each rank to its own share of the host cores.  Shows the host's ceiling
for bench.py's e2e figure at N GPUs (per-rank GB/s, and the pictures/s that bandwidth allows).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/exp_h2d_ranks.py"""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
bind = os.environ.get("GVC_BIND", "1") == "1"
if bind and world > 1:
    avail = sorted(os.sched_getaffinity(0))
    per = max(1, len(avail) // world)
    os.sched_setaffinity(0, set(avail[local * per:(local + 1) * per]))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
FB, N = 1920 * 1080 * 3 // 2, 304
host = torch.empty(N * FB, dtype=torch.uint8).pin_memory()
host.fill_(rank + 1)
dev = torch.empty_like(host, device="cuda")
s = torch.cuda.Stream()
def step(reps):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    with torch.cuda.stream(s):
        for _ in range(reps):
            dev.copy_(host, non_blocking=True)
    s.synchronize()
    return time.perf_counter() - t0
step(2)
dt = step(10)
gbs = 10 * host.numel() / dt / 1e9
t = torch.tensor([gbs], device="cuda")
if world > 1:
    allv = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allv, t)
    vals = [float(v) for v in allv]
else:
    vals = [gbs]
if rank == 0:
    print(json.dumps({"ranks": world, "bound_to_own_cores": bind, "h2d_gbs_per_rank": [round(v, 1) for v in vals], "h2d_gbs_total": round(sum(vals), 1),
                      "pictures_per_s_this_allows": round(min(vals) * 1e9 / FB * world), "cpu_affinity_rank0": len(os.sched_getaffinity(0))}))
if world > 1:
    dist.destroy_process_group()
