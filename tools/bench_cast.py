#!/usr/bin/env python
"""Device casts of the multi-file mapper, one column at a time: a one-row file fixes the column's type, a second file holds N
rows of another type; `b2a_multifile_finalize_run` (= one b2a_cast_kernel launch) is timed with CUDA events, inputs resident.
Prints one JSON line per case (appended to gpurun_out/bench_cast.jsonl as it goes).  Explains numbers, not the headline."""
import ctypes as C
import decimal
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import pyarrow as pa
    import pyarrow.compute as pc
    import pyarrow.ipc as ipc
    import torch
    import duckdb_arrow_b200 as b2a
    from duckdb_arrow_b200._lib import check, lib
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
    steps = 5
    rng = np.random.default_rng(1)
    tmp = tempfile.mkdtemp(dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    out_path = os.path.join(ROOT, "gpurun_out", "bench_cast.jsonl")

    def write(path, arr, batch=122_880):
        t = pa.table({"v": arr})
        with ipc.new_stream(path, t.schema) as w:
            w.write_table(t, max_chunksize=batch)

    doubles = pa.array(rng.random(n) * 1e6)
    cents = pa.array(rng.integers(-10**12, 10**12, n), pa.int64())
    cases = [
        ("VARCHAR -> DOUBLE", pa.array([1.0]), pc.cast(doubles, pa.string())),
        ("DOUBLE -> VARCHAR", pa.array(["x"]), doubles),
        ("BIGINT -> DECIMAL(18,2)", pa.array([decimal.Decimal("1.00")], pa.decimal128(18, 2)), pa.array(rng.integers(-10**15, 10**15, n), pa.int64())),
        ("VARCHAR -> BIGINT", pa.array([1], pa.int64()), pc.cast(cents, pa.string())),
    ]
    for k, (name, first, second) in enumerate(cases):
        a, b = os.path.join(tmp, f"a{k}.arrows"), os.path.join(tmp, f"b{k}.arrows")
        write(a, first)
        write(b, second)
        mfs = b2a.read_arrow([a, b])
        f = mfs.files[1]
        image = f.reader.image()
        scan = b2a.ArrowScan(f.reader, mfs.device)
        scan.upload()
        base = image.ctypes.data + scan.image_lo
        scan.decode(string_ptr_base=base)
        fin = C.c_void_p()
        check(lib.b2a_multifile_finalize_create(mfs._h, f.index, scan._h, C.byref(fin)))
        fb, nc = C.c_uint64(), C.c_uint64()
        check(lib.b2a_multifile_finalize_sizes(fin, C.byref(fb), C.byref(nc)))
        d_final = torch.empty(fb.value, dtype=torch.uint8, device=f"cuda:{mfs.device}")
        h_final = torch.empty(16, dtype=torch.uint8, pin_memory=True)

        def run():
            check(lib.b2a_multifile_finalize_run(fin, C.c_void_p(scan.d_out.data_ptr()), C.c_void_p(d_final.data_ptr()), base,
                                                 h_final.data_ptr(), scan._stream_ptr()))
        for _ in range(2):
            run()
        check(lib.b2a_multifile_finalize_finish(fin, scan._stream_ptr()))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        line = {"cast": name, "rows": n, "casts": int(nc.value), "ms": ms, "rows_per_s": n / ms * 1e3,
                "source_bytes": int(os.path.getsize(b))}
        print(json.dumps(line), flush=True)
        with open(out_path, "a") as fh:
            fh.write(json.dumps(line) + "\n")
        lib.b2a_multifile_finalize_destroy(fin)
        scan.finish()
        scan.close()


if __name__ == "__main__":
    main()
