#!/usr/bin/env python
"""bench.py — read_arrow throughput on synthetic TPC-H lineitem (BASELINE.json metric).

A "step" = one full read_arrow scan (all 16 columns, every record batch) of the rank's lineitem IPC stream.
  value : rows/s, whole job, stream image already resident in HBM (device-resident decode, CUDA events)
  e2e   : same metric through the C ABI with HOST buffers: pinned host stream -> H2D -> decode -> D2H of the
          DuckDB vectors into pinned host memory, all inside the timed region
  roofline : algorithmic bytes of one decode launch / its CUDA-event duration vs the measured HBM copy peak
  cpu_baseline : the oracle port of the reference's scan loop, timed on a bounded sample on the host cores

python bench.py [--gpus N] [--steps K] [--warmup W] [--sf 10] [--impl reference]
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            p = [x.strip() for x in l.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
            except ValueError:
                continue
            for n, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def build_stream(rows: int, seed: int, pin: bool):
    """lineitem IPC stream written by pyarrow straight into (pinned) host memory."""
    import pyarrow as pa
    import torch
    from tools import lineitem
    cap = int(rows * 150) + (1 << 20)
    host = torch.empty(cap, dtype=torch.uint8, pin_memory=pin)
    sink = pa.FixedSizeBufferWriter(pa.py_buffer(host.numpy()))
    t = time.time()
    n_batches = lineitem.write_stream(sink, rows, seed=seed, unique=16)
    size = sink.tell()
    log(f"[bench] generated {rows} rows / {n_batches} batches / {size / 1e9:.2f} GB in {time.time() - t:.1f}s")
    return host, size


def cpu_reference_rate(host_np, size, max_batches, seconds, threads):
    """Reference's read_arrow on the host cores (oracle port): per record-batch message DecodeBody through the 4 KiB
    bounce buffer (a1) + nanoarrow FULL validation (a2) + the 2 048-row scan loop (a4-a9), over a bounded sample of
    the same stream.  threads > 1 models a glob of `threads` files (the reference scans ONE thread per file,
    src/file_scanner/arrow_file_scan.cpp:35-41).  Returns rows/s and what was sampled."""
    from oracle import oracle_py as O
    O.lib()
    msgs = O.stream_messages(host_np[:size], max_batches)
    rows = sum(len(m[2]) for m in msgs)
    if not msgs:
        return 0.0, "empty"
    passes, t0 = 0, time.time()
    if threads <= 1:
        while True:
            O.read_arrow_checksum(msgs)
            passes += 1
            if time.time() - t0 >= seconds:
                break
    else:
        from concurrent.futures import ThreadPoolExecutor
        parts = [msgs[i::threads] for i in range(threads)]
        parts = [p for p in parts if p]
        with ThreadPoolExecutor(len(parts)) as ex:
            while True:
                list(ex.map(O.read_arrow_checksum, parts))
                passes += 1
                if time.time() - t0 >= seconds:
                    break
    dt = time.time() - t0
    return rows * passes / dt, f"{len(msgs)} batches ({rows} rows) x {passes} passes in {dt:.1f}s"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--sf", type=float, default=10.0)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--slices", type=int, default=0, help="pipeline slices of the host path (0 = library default)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--columns", default="", help="comma-separated projection (experiments only; the metric is all 16)")
    ap.add_argument("--profile", action="store_true", help="short run for ncu: no e2e leg, no CPU baseline, no clock soak")
    args = ap.parse_args()
    steps, warmup = max(1, args.steps), max(3, args.warmup) if args.impl != "reference" else max(0, args.warmup)

    from tools import lineitem
    rows = lineitem.ROWS_PER_SF.get(args.sf if args.sf not in (1.0, 10.0, 100.0) else int(args.sf),
                                    int(6_001_215 * args.sf))
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    workload = f"lineitem SF{args.sf:g} read_arrow full scan, 16 columns, {lineitem.BATCH_ROWS}-row batches, per GPU"
    config = {"workload": workload, "rows_per_gpu": rows, "batch_rows": lineitem.BATCH_ROWS,
              "l2": "inputs larger than L2 (no flush needed)", "sharding": "one IPC stream per GPU (files of a glob)"}

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return
        import torch  # pinned memory not needed
        host, size = build_stream(min(rows, 64 * lineitem.BATCH_ROWS), 42, pin=False)
        cores = os.cpu_count() or 1
        per = max(2.0, min(20.0, 120.0 / (steps + warmup + 1)))
        for _ in range(warmup):
            cpu_reference_rate(host.numpy(), size, 64, 1.0, cores)
        vals, sample = [], ""
        t0 = time.time()
        for _ in range(steps):
            v, sample = cpu_reference_rate(host.numpy(), size, 64, per, cores)
            vals.append(v)
        dt = (time.time() - t0) / steps
        v = float(np.mean(vals))
        one, one_sample = cpu_reference_rate(host.numpy(), size, 64, per, 1)
        print(json.dumps({
            "impl": "reference", "metric": "read_arrow rows/s (lineitem, all columns)", "value": v, "unit": "rows/s",
            "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": config,
            "cpu_baseline": {"value": v, "unit": "rows/s", "cores": cores, "kind": "port",
                             "sample": "oracle port of the reference read_arrow (DecodeBody 4 KiB bounce copy + FULL "
                                       "validation + scan loop), in-memory stream split over all host threads as if it "
                                       "were a glob of files, " + sample,
                             "single_file_value": one, "single_file_cores": 1,
                             "single_file_note": "what the reference does on ONE file: one scan thread per file "
                                                 "(arrow_file_scan.cpp:35-41); " + one_sample},
            "e2e": {"value": v, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    # ------------------------------------------------------------------ B200 arm
    import torch
    import duckdb_arrow_b200 as b2a
    from duckdb_arrow_b200._lib import check, lib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the scan path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    from duckdb_arrow_b200 import parallel
    numa = parallel.bind_to_device_numa_node(local_rank) if os.environ.get("B2A_NUMA_BIND", "1") != "0" else {}
    log(f"[bench] rank {rank}: numa {numa}, {len(os.sched_getaffinity(0))} cpus")
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))

    host, size = build_stream(rows, 42 + rank, pin=True)
    reader = b2a.ArrowIPCReader.open_buffers([(host.data_ptr(), size)])
    if args.columns:
        reader.set_projection(args.columns.split(","))
        config["columns"] = args.columns
    scan = b2a.ArrowScan(reader, device=local_rank)
    assert reader.num_rows == rows
    log(f"[bench] rank {rank}: image {scan.image_bytes / 1e9:.2f} GB, out {scan.out_bytes / 1e9:.2f} GB, "
        f"algorithmic {scan.algorithmic_bytes / 1e9:.2f} GB ({scan.algorithmic_bytes / rows:.1f} B/row), "
        f"{scan.n_tasks} tasks, {scan.n_tiles} tiles")
    dev = f"cuda:{local_rank}"
    d_image = torch.empty(scan.image_bytes + 256, dtype=torch.uint8, device=dev)
    d_out = torch.empty(scan.out_bytes, dtype=torch.uint8, device=dev)
    d_image[:scan.image_bytes].copy_(host[scan.image_lo:scan.image_lo + scan.image_bytes], non_blocking=True)
    scan.set_device_image(d_image)
    stream = torch.cuda.current_stream()

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident decode
    sampler = ClockSampler(local_rank)
    sampler.start()  # sampled from the warm-up to the end of the e2e leg (the device-timed region alone is ~ms)
    for _ in range(warmup):
        scan.decode(d_out)
    scan.finish()
    barrier()
    launches0 = scan.gpu_launches
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    t_all0, t_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_all0.record(stream)
    for a, z in evs:
        a.record(stream)
        scan.decode(d_out)
        z.record(stream)
    t_all1.record(stream)
    barrier()
    scan.finish()
    launches = scan.gpu_launches - launches0
    total_ms = t_all0.elapsed_time(t_all1)
    kernel_ms = float(np.mean([a.elapsed_time(z) for a, z in evs]))
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    r = torch.tensor([float(rows)], dtype=torch.float64, device=dev)
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(r, op=dist.ReduceOp.SUM)  # the only cross-GPU step of the path: combine per-GPU row counts
    total_ms, total_rows = float(t.item()), float(r.item())
    ms_per_step = total_ms / steps
    value = total_rows / (ms_per_step / 1e3)

    # ---- e2e: host stream -> H2D -> decode -> D2H host vectors, through b2a_read_arrow_host.
    # Headline = host-alias scan (DirectConversion columns alias the pinned host stream like the reference's zero-copy
    # vectors; only buffers that need conversion cross PCIe).  "full_copy" = every column up, decoded and down.
    e2e = None
    if not args.no_e2e and not args.profile:
        def run_e2e(host_alias):
            sc = b2a.ArrowScan(reader, device=local_rank, host_alias=host_alias)
            h_out = torch.empty(sc.out_bytes, dtype=torch.uint8, pin_memory=True)
            e_steps = max(2, min(steps, 5))

            def one():
                check(lib.b2a_read_arrow_host(sc._h, C.c_void_p(h_out.data_ptr()), args.slices))

            one()
            one()
            barrier()
            t0 = time.perf_counter()
            for _ in range(e_steps):
                one()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            te = torch.tensor([dt], dtype=torch.float64, device=dev)
            if dist:
                dist.all_reduce(te, op=dist.ReduceOp.MAX)
            rate = total_rows / (float(te.item()) / e_steps)
            res = {"value": rate, "unit": "rows/s", "h2d_bytes_per_step": int(sc.upload_bytes),
                   "d2h_bytes_per_step": int(sc.out_bytes), "steps": e_steps, "ms_per_step": float(te.item()) / e_steps * 1e3,
                   "launches_per_step": (sc.gpu_launches // (e_steps + 2))}
            sc.close()
            del h_out
            return res

        e2e = run_e2e(True)
        e2e["what"] = ("b2a_read_arrow_host on a B2A_SCAN_HOST_ALIAS scan: pinned host IPC stream -> H2D of the buffers "
                       "that need conversion -> decode -> D2H of the converted vectors (pinned); fixed-width columns "
                       "alias the host stream (reference DirectConversion)")
        full = run_e2e(False)
        e2e["full_copy"] = {k: full[k] for k in ("value", "h2d_bytes_per_step", "d2h_bytes_per_step", "ms_per_step")}

    # keep the GPU busy for a moment so that the 100 ms nvidia-smi sampler sees clocks under load
    t_busy = time.time()
    while time.time() - t_busy < 1.0 and not args.profile:
        for _ in range(20):
            scan.decode(d_out)
        torch.cuda.synchronize()
    clocks = sampler.stop()
    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return

    peak, peak_kind = measured_peak()
    achieved = scan.algorithmic_bytes / (kernel_ms / 1e3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get("b2a_decode_kernel_dram_bytes_per_launch")
        except Exception:
            traffic = None
    cores = 1
    if args.profile:
        cpu_v, cpu_sample = 0.0, "skipped (--profile)"
    else:
        cpu_v, cpu_sample = cpu_reference_rate(host.numpy(), size, 64, args.cpu_seconds, cores)
    out = {
        "metric": "read_arrow rows/s (lineitem, all columns)", "value": value, "unit": "rows/s",
        "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config,
        "gb_per_s": value * scan.algorithmic_bytes / rows / 1e9,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_kind": peak_kind, "kernel": "b2a_decode_kernel",
                     "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": int(scan.algorithmic_bytes),
                     "frac_of_nominal_8TBs": achieved / 8000.0},
        "cpu_baseline": {"value": cpu_v, "unit": "rows/s", "cores": cores, "kind": "port", "sample": cpu_sample,
                         "host_cores_available": os.cpu_count()},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    print(json.dumps(out))
    if dist:
        dist.destroy_process_group()


def _stdout_for_the_json_line_only():
    """The contract is ONE JSON line on stdout.  Native libraries write there too (NCCL prints its version line to fd 1
    when NCCL_DEBUG is set in the environment): send fd 1 to stderr for the whole run and keep the real stdout for
    print()."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)


if __name__ == "__main__":
    _stdout_for_the_json_line_only()
    main()
    sys.stdout.flush()
