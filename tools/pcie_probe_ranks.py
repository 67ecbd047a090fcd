#!/usr/bin/env python
"""PCIe probe on ALL ranks at once: pure pinned H2D + D2H, no kernels of ours — what the box gives N GPUs that stream
through host memory concurrently.  This is the ceiling of bench.py's end-to-end leg at N GPUs (every rank moves
3.9 GB up and 4.8 GB down per SF10 scan).  Launch: python tools/pcie_probe_ranks.py | torchrun --nproc-per-node N ...
One JSON line on rank 0: per-direction GB/s per rank (min / median / max over ranks) and the aggregate."""
import json
import os
import sys

import torch

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local_rank = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local_rank)
dev = f"cuda:{local_rank}"
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device(dev))

GB = 1 << 30
n = GB
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d_a = torch.empty(n, dtype=torch.uint8, device=dev)
d_b = torch.empty(n, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()


def run(do_h2d, do_d2h, reps=4):
    def once():
        if do_h2d:
            with torch.cuda.stream(s1):
                d_a.copy_(h_in, non_blocking=True)
        if do_d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_b, non_blocking=True)
    once()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_stream(torch.cuda.current_stream())
    s2.wait_stream(torch.cuda.current_stream())
    for _ in range(reps):
        once()
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / reps
    gbs = (n * (int(do_h2d) + int(do_d2h))) / ms / 1e6
    t = torch.tensor([gbs], dtype=torch.float64, device=dev)
    if dist:
        allv = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allv, t)
        vals = sorted(float(x.item()) for x in allv)
    else:
        vals = [gbs]
    return {"per_rank_min": round(vals[0], 1), "per_rank_median": round(vals[len(vals) // 2], 1),
            "per_rank_max": round(vals[-1], 1), "aggregate": round(sum(vals), 1)}


out = {"n_gpus": world, "bytes_per_copy": n, "unit": "GB/s",
       "h2d_only": run(True, False), "d2h_only": run(False, True), "both": run(True, True)}
if rank == 0:
    print(json.dumps(out))
if dist:
    dist.destroy_process_group()
