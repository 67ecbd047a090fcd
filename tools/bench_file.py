"""read_arrow('lineitem.arrows') end to end FROM A FILE (page cache -> host vectors): the case SURVEY §8f N-3 is about.
The file is mmap'ed by the reader (a pageable image); b2a_read_arrow_host then packs the upload ranges into pinned
staging buffers on a few host threads, sends one large copy per slice and scatters on the device
(scanner/host_stage.hpp).  B2A_H2D_STAGE=0 keeps one cudaMemcpyAsync per range from the pageable mapping instead.
One JSON line."""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=2.0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--slices", type=int, default=0)
    ap.add_argument("--dir", default="/dev/shm" if os.path.isdir("/dev/shm") else "/tmp")
    args = ap.parse_args()
    import torch
    import duckdb_arrow_b200 as b2a
    from duckdb_arrow_b200._lib import lib, check
    from tools import lineitem

    rows = int(6_001_215 * args.sf)
    path = os.path.join(args.dir, f"b2a_bench_lineitem_sf{args.sf}.arrows")
    with open(path, "wb") as f:
        lineitem.write_stream(f, rows)
    try:
        reader = b2a.ArrowIPCReader.open_file(path)
        sc = b2a.ArrowScan(reader, device=0, host_alias=True)
        h_out = torch.empty(sc.out_bytes, dtype=torch.uint8, pin_memory=True)

        def one():
            check(lib.b2a_read_arrow_host(sc._h, C.c_void_p(h_out.data_ptr()), args.slices))

        one()
        one()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            one()
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) / args.steps * 1e3
        print(json.dumps({"metric": "read_arrow(path) rows/s, file in page cache -> host vectors (host-alias scan)",
                          "value": reader.num_rows / ms * 1e3, "unit": "rows/s", "ms_per_step": ms, "rows": reader.num_rows,
                          "h2d_bytes_per_step": int(sc.upload_bytes), "d2h_bytes_per_step": int(sc.out_bytes),
                          "stage": os.environ.get("B2A_H2D_STAGE", "auto"),
                          "threads": os.environ.get("B2A_STAGE_THREADS", "default"),
                          "file_bytes": os.path.getsize(path), "dir": args.dir}))
        sc.close()
    finally:
        os.unlink(path)


if __name__ == "__main__":
    main()
