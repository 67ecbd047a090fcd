#!/usr/bin/env python
"""read_arrow of a compressed lineitem stream (IpcWriteOptions(compression='zstd'), what the reference reads through
ZSTD_decompress per buffer; --codec lz4 = LZ4_FRAME bodies, which the reference cannot read): device-resident
decompress + decode, CUDA events.  One JSON line.
python tools/bench_zstd.py [--sf 1] [--steps 5] [--codec zstd|lz4]"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=1.0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--level", type=int, default=1)
    ap.add_argument("--codec", default="zstd", choices=["zstd", "lz4"])
    ap.add_argument("--columns", default="")
    args = ap.parse_args()
    import pyarrow as pa
    import pyarrow.ipc as ipc
    import torch
    import duckdb_arrow_b200 as b2a
    from tools import lineitem
    from bench import measured_peak
    rows = int(6_001_215 * args.sf)
    t0 = time.time()
    sink = pa.BufferOutputStream()
    codec = pa.Codec("zstd", compression_level=args.level) if args.codec == "zstd" else pa.Codec("lz4")
    opt = ipc.IpcWriteOptions(compression=codec)
    with ipc.new_stream(sink, lineitem.schema(), options=opt) as w:
        for b in lineitem.iter_batches(rows, seed=42, unique=16):
            w.write_batch(b)
    buf = sink.getvalue()
    print(f"[bench_compressed] {rows} rows -> {buf.size / 1e9:.3f} GB compressed in {time.time() - t0:.1f}s", file=sys.stderr)
    reader = b2a.ArrowIPCReader.open_buffers([buf])
    if args.columns:
        reader.set_projection(args.columns.split(","))
    scan = b2a.ArrowScan(reader, device=0)
    dev = "cuda:0"
    d_image = torch.empty(scan.device_image_bytes, dtype=torch.uint8, device=dev)
    host = torch.frombuffer(buf, dtype=torch.uint8)
    d_image[:scan.image_bytes].copy_(host[scan.image_lo:scan.image_lo + scan.image_bytes])
    scan.set_device_image(d_image)
    d_out = torch.empty(scan.out_bytes, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()
    for _ in range(args.warmup):
        scan.decode(d_out)
    scan.finish()
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(args.steps):
        scan.decode(d_out)
    z.record(stream)
    torch.cuda.synchronize()
    scan.finish()
    ms = a.elapsed_time(z) / args.steps
    # CPU reference point: pyarrow (libzstd, one thread per column by default) decompressing the same stream
    t0 = time.time()
    n = sum(len(b) for b in ipc.open_stream(buf))
    cpu_s = time.time() - t0
    peak, kind = measured_peak()
    print(json.dumps({
        "metric": f"read_arrow rows/s, {args.codec.upper()}-compressed lineitem (device-resident decompress + decode)",
        "unit": "rows/s", "value": rows / (ms / 1e3), "ms_per_step": ms, "rows": rows,
        "compressed_bytes": int(scan.image_bytes), "decompressed_bytes": int(scan.scratch_bytes),
        "ratio": scan.scratch_bytes / max(1, scan.image_bytes), "jobs": int(scan.n_decompress_jobs),
        "decompressed_gb_per_s": scan.scratch_bytes / (ms / 1e3) / 1e9,
        "pyarrow_cpu_decompress_rows_per_s": n / cpu_s, "pyarrow_cpu_threads": pa.cpu_count(),
        "config": {"workload": f"lineitem SF{args.sf:g}, " + (f"zstd level {args.level}" if args.codec == "zstd" else "lz4 frame") + ", 122880-row batches"}}))


if __name__ == "__main__":
    main()
