"""TPC-H lineitem as the reference would write it: DuckDB dbgen -> the reference's ArrowAppender -> IPC stream.

  CALL dbgen(sf)                       duckdb/extension/tpch/dbgen/dbgen.cpp (the reference's DuckDB, oracle/_ref)
  COPY lineitem TO ... (FORMAT arrows) = per row_group_size (122 880, src/writer/write_arrow_stream.cpp:33) rows:
      ArrowConverter::ToArrowArray (the reference's appender, run here through oracle/ref_py.export)
      + framing of every buffer incl. the all-valid bitmaps (tools/ipc_frame.py — nanoarrow's encoder is not on disk)

This is the workload BASELINE.json's north_star names ("synthetic TPC-H lineitem (DuckDB dbgen, written to IPC by the
reference's own COPY)").  Bench / test infrastructure: uses oracle/_ref, never imported by the product.

The stream is cached under /dev/shm (B2A_DATA_DIR) so that the reference arm, the B200 arm and every --gpus N run of one
box generate it once; concurrent ranks wait for the one that got the lock.
"""
from __future__ import annotations

import fcntl
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH_ROWS = 122_880
ROWS = {0.01: 60_175, 0.1: 600_572, 1: 6_001_215, 10: 59_986_052}


def data_dir() -> str:
    d = os.environ.get("B2A_DATA_DIR", "/dev/shm/b2a_data")
    os.makedirs(d, exist_ok=True)
    return d


def available() -> bool:
    from oracle import ref_py
    return ref_py.available()


def generate(path: str, sf: float, batch_rows: int = BATCH_ROWS, threads: int | None = None, log=None) -> dict:
    from oracle import ref_py
    from tools import ipc_frame
    t0 = time.time()
    ref_py.set_threads(threads or max(1, min(os.cpu_count() or 1, 32)))
    ref_py.query(f"CALL dbgen(sf={sf:g})")
    t1 = time.time()
    rows = n_batches = size = 0
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        schema_written = False
        for b in ref_py.export("SELECT * FROM lineitem", batch_rows):
            if not schema_written:
                size += f.write(ipc_frame.schema_message(b.schema))
                schema_written = True
            size += f.write(ipc_frame.record_batch_message(b))
            rows += b.num_rows
            n_batches += 1
        size += f.write(ipc_frame.EOS)
    for t in ("lineitem", "orders", "partsupp", "part", "customer", "supplier", "nation", "region"):
        ref_py.query(f"DROP TABLE {t}")
    ref_py.set_threads(1)
    os.replace(tmp, path)
    info = {"rows": rows, "batches": n_batches, "bytes": size, "dbgen_s": t1 - t0, "export_s": time.time() - t1}
    if log:
        log(f"[dbgen] sf={sf:g}: {rows} rows / {n_batches} batches / {size / 1e9:.2f} GB "
            f"(dbgen {info['dbgen_s']:.1f}s + appender/framing {info['export_s']:.1f}s) -> {path}")
    return info


def lineitem_path(sf: float, batch_rows: int = BATCH_ROWS, log=None) -> str:
    """Path of the cached dbgen lineitem stream of scale factor `sf` (generated on first use, under a file lock)."""
    path = os.path.join(data_dir(), f"lineitem_sf{sf:g}_rg{batch_rows}.arrows")
    if os.path.exists(path):
        return path
    with open(path + ".lock", "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        try:
            if not os.path.exists(path):
                generate(path, sf, batch_rows, log=log)
        finally:
            fcntl.flock(lk, fcntl.LOCK_UN)
    return path


def message_index(path: str):
    """[(offset, metadata_bytes, body_bytes, rows)] of the record-batch messages + the schema, by walking the message
    prefixes (what the reference's reader does sequentially, ipc_file_stream_reader.cpp:96-132)."""
    import pyarrow as pa
    import pyarrow.ipc as ipc
    out = []
    with pa.memory_map(path) as mm:
        reader = ipc.MessageReader.open_stream(mm)
        pos, schema = 0, None
        for msg in reader:
            meta = msg.metadata.size
            start = pos
            pos += 8 + meta + msg.body.size
            if msg.type == "schema":
                schema = ipc.read_schema(msg)
            elif msg.type == "record batch":
                out.append((start, 8 + meta, msg.body.size))
    return schema, out


if __name__ == "__main__":
    sf = float(sys.argv[1]) if len(sys.argv) > 1 else 1
    p = lineitem_path(sf, log=lambda *a: print(*a, file=sys.stderr))
    print(p, os.path.getsize(p))
