#!/usr/bin/env python
"""SQL through the drop-in: the reference's DuckDB (oracle/_ref/libduckdb_ref.so) LOADs integration/duckdb/
libnanoarrow_b2a.so and runs statements over read_arrow('<dbgen SF file>'); next to it the same statements over the
reference's own arrow_scan of the same stream (its CPU converters).  Wall clock, one JSON line.  Uses oracle/ (the
compiled reference) as the HOST APPLICATION of the extension and as the baseline — a bench tool, not the product.
python tools/bench_sql.py [--sf 10] [--threads 16] [--repeat 3]"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

Q6 = ("SELECT sum(l_extendedprice * l_discount) FROM {} WHERE l_shipdate >= DATE '1994-01-01' AND "
      "l_shipdate < DATE '1995-01-01' AND l_discount BETWEEN 0.05 AND 0.07 AND l_quantity < 24")
STATEMENTS = {
    "q6": Q6,
    "count_star": "SELECT count(*) FROM {}",
    "sum_quantity": "SELECT sum(l_quantity) FROM {}",
    "count_all_columns": ("SELECT " + ", ".join(f"count({c})" for c in (
        "l_orderkey l_partkey l_suppkey l_linenumber l_quantity l_extendedprice l_discount l_tax l_returnflag l_linestatus "
        "l_shipdate l_commitdate l_receiptdate l_shipinstruct l_shipmode l_comment").split()) + " FROM {}"),
    "all_columns": ("SELECT count(l_orderkey), count(l_partkey), count(l_suppkey), count(l_linenumber), sum(l_quantity), "
                    "sum(l_extendedprice), sum(l_discount), sum(l_tax), max(l_returnflag), max(l_linestatus), "
                    "max(l_shipdate), max(l_commitdate), max(l_receiptdate), max(l_shipinstruct), max(l_shipmode), "
                    "max(l_comment) FROM {}"),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", type=float, default=10.0)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--repeat", type=int, default=3)
    ap.add_argument("--skip-reference", action="store_true")
    ap.add_argument("--only", default="", help="comma-separated statement names")
    args = ap.parse_args()
    import pyarrow.ipc as ipc
    from oracle import ref_py as R
    from bench import stream_path
    path, kind = stream_path(args.sf)
    R.load_extension(os.path.join(ROOT, "integration", "duckdb", "libnanoarrow_b2a.so"))
    R.set_threads(args.threads)
    out = {"file": path, "data": kind, "threads": args.threads, "statements": {}}
    for name, sql in STATEMENTS.items():
        if args.only and name not in args.only.split(","):
            continue
        times, ans = [], None
        for _ in range(args.repeat + 1):
            t0 = time.perf_counter()
            ans = R.query(sql.format(f"read_arrow('{path}')"))
            times.append(time.perf_counter() - t0)
        rec = {"b200_extension_s": min(times[1:]), "first_call_s": times[0], "answer": ans[0]}
        if not args.skip_reference:
            # the reference's converters under the same DuckDB: arrow_scan over a pyarrow reader of the same file
            t0 = time.perf_counter()
            R.register("b2a_sql_ref", ipc.open_stream(path))
            ref_ans = R.query(sql.format("b2a_sql_ref"))
            rec["reference_arrow_scan_s"] = time.perf_counter() - t0
            rec["same_answer"] = ref_ans[0] == ans[0]
            rec["speedup"] = rec["reference_arrow_scan_s"] / rec["b200_extension_s"]
        out["statements"][name] = rec
        print(f"[bench_sql] {name}: {rec}", file=sys.stderr)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
