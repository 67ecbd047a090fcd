#!/usr/bin/env python
"""BASELINE config 3: lineitem SF10 read_arrow with pushed-down filters (l_shipdate range, l_quantity < 24) + the
Q6-style aggregate, record batches of ONE stream sharded contiguously across the GPUs (strong scaling); only the
per-GPU {rows, selected, sum} are combined (NCCL all_reduce of 12 int64 lanes).  Device-resident; CUDA events;
max over ranks.  One JSON line on rank 0.  Launch: python tools/bench_filter.py  |  torchrun ... tools/bench_filter.py"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=10.0)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    import torch
    import duckdb_arrow_b200 as b2a
    from duckdb_arrow_b200 import parallel as P
    from tools import lineitem
    from bench import build_stream, measured_peak

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    rows = lineitem.ROWS_PER_SF.get(args.sf if args.sf not in (1.0, 10.0, 100.0) else int(args.sf),
                                    int(6_001_215 * args.sf))
    host, size = build_stream(rows, 42, pin=False)  # the same stream on every rank; each uploads only its share
    reader = b2a.ArrowIPCReader.open_buffers([(host.data_ptr(), size)])
    scan = P.sharded_scan(reader, rank, world, device=local_rank)
    my_rows = sum(reader.batch_info(b).rows for b in range(scan.batch_lo, scan.batch_hi))
    dev = f"cuda:{local_rank}"
    d_image = torch.empty(scan.image_bytes + 256, dtype=torch.uint8, device=dev)
    d_image[:scan.image_bytes].copy_(host[scan.image_lo:scan.image_lo + scan.image_bytes])
    scan.set_device_image(d_image)
    d_out = torch.empty(scan.out_bytes, dtype=torch.uint8, device=dev)
    preds = [("l_shipdate", ">=", 8766), ("l_shipdate", "<", 9131), ("l_quantity", "<", 2400)]
    stream = torch.cuda.current_stream()

    def step():
        scan.decode(d_out)
        sel, count, total = scan.filter(preds, d_out)
        s = scan.sum_product("l_extendedprice", "l_discount", sel, count, d_out)  # reads the result back (sync)
        return P.combine_counts({"rows": my_rows, "selected": int(total.item()), "sum": s}, device=dev)

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        res = step()
    barrier()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(args.steps):
        res = step()
    z.record(stream)
    barrier()
    ms = a.elapsed_time(z) / args.steps
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    if rank == 0:
        peak, kind = measured_peak()
        print(json.dumps({
            "metric": "read_arrow + pushed-down filters + Q6 aggregate, rows/s (lineitem, device-resident)",
            "unit": "rows/s", "value": res["rows"] / (ms / 1e3), "n_gpus": world, "ms_per_step": ms,
            "scaling": "strong", "steps": args.steps, "rows": res["rows"], "selected": res["selected"],
            "selectivity": res["selected"] / max(1, res["rows"]), "sum_extendedprice_x_discount": str(res["sum"]),
            "collective": "all_reduce(SUM) of 12 int64 lanes (rows, selected, 128-bit sum as limbs)",
            "config": {"workload": f"lineitem SF{args.sf:g}, filters l_shipdate in [1994-01-01,1995-01-01) and "
                                   "l_quantity < 24, batches sharded contiguously across GPUs"}}))
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
