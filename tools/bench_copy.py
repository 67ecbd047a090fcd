#!/usr/bin/env python
"""Writer-path throughput (BASELINE config 5): COPY lineitem TO (FORMAT ARROWS), device-resident.

Input = the DuckDB vectors of lineitem in HBM (built by decoding a synthetic lineitem IPC stream with the scan
kernels and concatenating the per-batch slabs into one vector per column); output = the complete IPC stream in HBM.
Prints one JSON line: rows/s and GB/s of pass 1 (measure) + pass 2 (encode), and of the encode kernel alone.
Not the headline metric (that is bench.py); numbers land in profiles/."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=10.0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--row-group-size", type=int, default=122_880)
    ap.add_argument("--verify-batches", type=int, default=3)
    ap.add_argument("--columns", default="", help="comma-separated subset (experiments)")
    ap.add_argument("--large-offsets", action="store_true", help="SET arrow_large_buffer_size = true")
    ap.add_argument("--string-view", action="store_true", help="SET produce_arrow_string_view = true")
    args = ap.parse_args()
    import pyarrow as pa
    import pyarrow.ipc as ipc
    import torch
    import duckdb_arrow_b200 as b2a
    from tools import lineitem
    from bench import build_stream, measured_peak

    rows = lineitem.ROWS_PER_SF.get(args.sf if args.sf not in (1.0, 10.0, 100.0) else int(args.sf),
                                    int(6_001_215 * args.sf))
    host, size = build_stream(rows, 42, pin=True)
    reader = b2a.ArrowIPCReader.open_buffers([(host.data_ptr(), size)])
    scan = b2a.ArrowScan(reader, device=0)
    d_image = torch.empty(scan.image_bytes + 256, dtype=torch.uint8, device="cuda:0")
    d_image[:scan.image_bytes].copy_(host[scan.image_lo:scan.image_lo + scan.image_bytes])
    scan.set_device_image(d_image)
    d_out = scan.decode()
    scan.finish()
    cols = []
    want = set(args.columns.split(",")) if args.columns else None
    for f in reader.columns:
        if want and f.name not in want:
            continue
        parts = []
        for b in range(reader.num_batches):
            co = scan.column(b, f.index)
            parts.append(d_out[co.data_off: co.data_off + co.rows * f.out_width])
        data = torch.cat(parts)
        cols.append(b2a.DeviceVector(f.type, rows, data, None, name=f.name, decimal=(f.decimal_width, f.decimal_scale)))
    del d_out, scan
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream()

    def new_writer():
        return b2a.ArrowWriter(cols, rows, args.row_group_size, arrow_large_buffer_size=args.large_offsets,
                               produce_arrow_string_view=args.string_view)

    w = new_writer()
    w.measure()
    d_stream = torch.empty(w.stream_bytes, dtype=torch.uint8, device="cuda:0")
    for _ in range(args.warmup):
        w.encode(d_stream)
    w.finish()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for a, z in ev:
        a.record(stream)
        w.encode(d_stream)
        z.record(stream)
    torch.cuda.synchronize()
    w.finish()
    encode_ms = float(np.mean([a.elapsed_time(z) for a, z in ev]))
    # pass 1 (create + measure): host planning + 2 small kernels + one sync, timed by wall clock
    m = []
    for _ in range(max(2, args.steps // 2)):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        w2 = new_writer()
        w2.measure()
        torch.cuda.synchronize()
        m.append((time.perf_counter() - t0) * 1e3)
        w2.close()
    measure_ms = float(np.median(m))
    # verify a few batches against the source through pyarrow
    h = torch.empty(w.stream_bytes, dtype=torch.uint8, pin_memory=True)
    h.copy_(d_stream)
    torch.cuda.synchronize()
    out = h.numpy()
    w.headers(out)
    src = ipc.open_stream(pa.py_buffer(host.numpy()[:size]))
    got = ipc.open_stream(pa.py_buffer(out))
    ok = True
    for i, (a, b) in enumerate(zip(src, got)):
        if i >= args.verify_batches:
            break
        ok = ok and a.num_rows == b.num_rows and all(
            a.column(n).to_pylist() == b.column(n).to_pylist() for n in b.schema.names)
    peak, kind = measured_peak()
    alg = w.algorithmic_bytes
    print(json.dumps({
        "metric": "COPY lineitem TO (FORMAT ARROWS) rows/s, device-resident", "unit": "rows/s",
        "value": rows / ((encode_ms + measure_ms) / 1e3), "encode_only_value": rows / (encode_ms / 1e3),
        "encode_ms": encode_ms, "measure_ms": measure_ms, "rows": rows, "n_batches": w.n_batches,
        "stream_bytes": w.stream_bytes, "algorithmic_bytes": alg,
        "roofline": {"bound": "hbm", "kernel": "b2a_encode_kernel", "achieved": alg / (encode_ms / 1e3) / 1e9,
                     "peak": peak, "peak_kind": kind, "unit": "GB/s", "frac": alg / (encode_ms / 1e3) / 1e9 / peak},
        "verified_batches_equal_source": bool(ok), "gpu_launches_per_copy": int(w2.gpu_launches if False else 3),
        "config": {"workload": f"lineitem SF{args.sf:g} COPY TO arrows, row_group_size {args.row_group_size}",
                   "arrow_large_buffer_size": args.large_offsets, "produce_arrow_string_view": args.string_view}}))


if __name__ == "__main__":
    main()
