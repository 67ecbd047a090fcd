#!/usr/bin/env python
"""BASELINE config 4: multi-file glob read_arrow of lineitem files (SF100 / 64 rows each) with union_by_name and
dictionary-encoded strings; files of the glob are sharded over the GPUs (8 files per GPU: 64 files at 8 GPUs).

Every second file has a different column order and a narrower `l_linenumber` (INTEGER instead of BIGINT), so the
union_by_name binding produces a non-trivial mapping and one device cast per such file; l_returnflag / l_linestatus /
l_shipinstruct / l_shipmode are dictionary encoded (int8 / int32 indices) and come out as selection vectors over a
dictionary decoded once per file.  Device-resident (file images uploaded once, then timed with CUDA events, max over
ranks) plus one wall-clock pass that includes bind + H2D from the page cache.  One JSON line on rank 0.
Launch: python tools/bench_multifile.py | torchrun --nproc-per-node N ... tools/bench_multifile.py"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def write_files(tmp: str, rows: int, n_files: int, seed: int):
    import pyarrow as pa
    from tools import lineitem
    base = lineitem.schema(dictionary_strings=True)
    order_b = [base.names[i] for i in (15, 3, 0, 1, 2, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14)]
    fields_b = [pa.field("l_linenumber", pa.int32()) if n == "l_linenumber" else base.field(n) for n in order_b]
    schema_b = pa.schema(fields_b)
    pa_, pb_ = os.path.join(tmp, "variant_a.arrows"), os.path.join(tmp, "variant_b.arrows")
    with pa.OSFile(pa_, "wb") as fa, pa.OSFile(pb_, "wb") as fb:
        with pa.ipc.new_stream(fa, base) as wa, pa.ipc.new_stream(fb, schema_b) as wb:
            for b in lineitem.iter_batches(rows, seed=seed, dictionary_strings=True, unique=8):
                wa.write_batch(b)
                cols = [b.column(base.get_field_index(n)) for n in order_b]
                cols[1] = cols[1].cast(pa.int32())
                wb.write_batch(pa.RecordBatch.from_arrays(cols, schema=schema_b))
    # the glob: n_files names; identical contents are fine for throughput (each file is uploaded and decoded on its own)
    gdir = os.path.join(tmp, "glob")
    os.makedirs(gdir)
    for i in range(n_files):
        os.symlink(pb_ if i % 2 else pa_, os.path.join(gdir, f"lineitem_{i:03d}.arrows"))
    return os.path.join(gdir, "*.arrows")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--files-per-gpu", type=int, default=8)
    ap.add_argument("--rows-per-file", type=int, default=600_037_902 // 64)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--streams", type=int, default=4)
    args = ap.parse_args()
    import torch
    import duckdb_arrow_b200 as b2a
    from duckdb_arrow_b200 import parallel as P
    from duckdb_arrow_b200._lib import check, lib
    from bench import measured_peak

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device(dev))
    tmp = tempfile.mkdtemp(prefix=f"b2a_mf_{rank}_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        # every rank writes its own share of the glob (the files a rank reads are the ones it would be handed by
        # parallel.file_shard over the whole glob; identical generator, rank-specific seed)
        pattern = write_files(tmp, args.rows_per_file, args.files_per_gpu, 100 + rank)
        t0 = time.perf_counter()
        mfs = b2a.read_arrow(pattern, union_by_name=True)
        bind_s = time.perf_counter() - t0
        stream = torch.cuda.current_stream()
        sp = C.c_void_p(stream.cuda_stream)
        files = []
        t0 = time.perf_counter()
        for f in mfs.files:
            scan = b2a.ArrowScan(f.reader, device=local_rank)
            scan.upload()
            d_out = torch.empty(scan.out_bytes, dtype=torch.uint8, device=dev)
            fin = C.c_void_p()
            check(lib.b2a_multifile_finalize_create(mfs._h, f.index, scan._h, C.byref(fin)))
            fb, nc = C.c_uint64(), C.c_uint64()
            check(lib.b2a_multifile_finalize_sizes(fin, C.byref(fb), C.byref(nc)))
            d_final = torch.empty(max(fb.value, 256), dtype=torch.uint8, device=dev)
            files.append((f, scan, d_out, fin, d_final, nc.value, fb.value))

        # files are independent: their decode (+cast) launches go round-robin over a few streams so that the tail of one
        # file's persistent grid overlaps the ramp-up of the next
        side = [torch.cuda.Stream(device=dev) for _ in range(args.streams)]

        def step():
            for st in side:
                st.wait_stream(stream)
            for i, (f, scan, d_out, fin, d_final, n_casts, _) in enumerate(files):
                st = side[i % len(side)]
                with torch.cuda.stream(st):
                    scan.decode(d_out)
                    if n_casts:
                        check(lib.b2a_multifile_finalize_run(fin, C.c_void_p(d_out.data_ptr()),
                                                             C.c_void_p(d_final.data_ptr()), 0, 0,
                                                             C.c_void_p(st.cuda_stream)))
            for st in side:
                stream.wait_stream(st)

        def finish():
            for f, scan, d_out, fin, d_final, n_casts, _ in files:
                scan.finish()
                check(lib.b2a_multifile_finalize_finish(fin, sp))

        step()
        finish()
        torch.cuda.synchronize()
        first_pass_s = time.perf_counter() - t0  # open scans + H2D from the page cache + decode + casts

        def barrier():
            torch.cuda.synchronize()
            if dist:
                dist.barrier()
            torch.cuda.synchronize()

        for _ in range(max(3, args.warmup)):
            step()
        finish()
        barrier()
        launches0 = sum(x[1].gpu_launches + int(lib.b2a_multifile_finalize_gpu_launches(x[3])) for x in files)
        a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(args.steps):
            step()
        z.record(stream)
        barrier()
        finish()
        launches = sum(x[1].gpu_launches + int(lib.b2a_multifile_finalize_gpu_launches(x[3])) for x in files) - launches0
        ms = a.elapsed_time(z) / args.steps
        t = torch.tensor([ms, first_pass_s], dtype=torch.float64, device=dev)
        if dist:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, first_pass_s = [float(x) for x in t.tolist()]
        my_rows = sum(x[0].reader.num_rows for x in files)
        # algorithmic bytes: IPC buffers read + vectors written by the scans, + cast input/output
        alg = sum(x[1].algorithmic_bytes for x in files)
        cast_bytes = sum(x[0].reader.num_rows * 12 for x in files if x[5])  # INTEGER (4 B) in, BIGINT (8 B) out
        tot = P.combine_counts({"rows": my_rows, "alg": alg + cast_bytes, "files": len(files)}, device=dev)
        if rank == 0:
            peak, kind = measured_peak()
            gbs = tot["alg"] / (ms / 1e3) / 1e9
            print(json.dumps({
                "metric": "multi-file read_arrow rows/s (lineitem, dictionary-encoded strings, union_by_name, "
                          "device-resident)",
                "unit": "rows/s", "value": tot["rows"] / (ms / 1e3), "n_gpus": world, "scaling": "weak",
                "ms_per_step": ms, "steps": args.steps, "files": tot["files"], "rows": tot["rows"],
                "columns": mfs.names, "types": mfs.types,
                "casts_per_step": sum(1 for x in files if x[5]), "gpu_launches_per_step": launches // args.steps,
                "algorithmic_gb_per_s": gbs,
                "roofline": {"bound": "hbm", "achieved": gbs / world, "peak": peak, "peak_kind": kind, "unit": "GB/s",
                             "frac": gbs / world / peak, "per": "GPU"},
                "first_pass_wall_s": first_pass_s, "bind_s": bind_s,
                "first_pass_rows_per_s": tot["rows"] / first_pass_s,
                "collective": "all_reduce(SUM) of {rows, bytes, files} (12 int64 lanes)",
                "config": {"workload": f"{tot['files']} lineitem files x {args.rows_per_file} rows (SF100/64 each), "
                                       "4 dictionary-encoded string columns, union_by_name over two schema variants "
                                       "(column order, INTEGER vs BIGINT l_linenumber), files sharded over GPUs"}}))
        for x in files:
            lib.b2a_multifile_finalize_destroy(x[3])
            x[1].close()
        mfs.close()
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
