"""Writes tests/golden/dbgen/*: a small lineitem stream produced by the REFERENCE (oracle/_ref: DuckDB dbgen -> the
reference's ArrowAppender; framed by tools/ipc_frame.py with the appender's validity bitmaps), the reference's own
answers for it, and the reference appender's buffers for a few writer inputs.  Run in the container that holds
/root/reference; the outputs are committed so that the GPU box (which has no /root/reference) tests against vectors the
reference itself produced.

  python tools/make_golden_dbgen.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_py  # noqa: E402
from tools import dbgen_lineitem, ipc_frame  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "dbgen")
SF, BATCH = 0.002, 2500
Q6 = ("SELECT count(*), sum(l_extendedprice * l_discount) FROM lineitem WHERE l_shipdate >= DATE '1994-01-01' AND "
      "l_shipdate < DATE '1995-01-01' AND l_discount BETWEEN 0.05 AND 0.07 AND l_quantity < 24")


def main():
    os.makedirs(OUT, exist_ok=True)
    ref_py.query(f"CALL dbgen(sf={SF})")
    batches = list(ref_py.export("SELECT * FROM lineitem", BATCH))
    path = os.path.join(OUT, f"lineitem_sf{SF}_rg{BATCH}.arrows")
    with open(path, "wb") as f:
        rows, size = ipc_frame.write_stream(f, batches[0].schema, batches)
    answers = {
        "source": f"reference DuckDB {ref_py.source_id()} (oracle/_ref): CALL dbgen(sf={SF}); export of lineitem in "
                  f"batches of {BATCH} rows through ArrowConverter::ToArrowArray",
        "rows": rows, "bytes": size, "batches": len(batches),
        "q6_count_sum": ref_py.query(Q6)[0],
        "count_quantity_lt_24": ref_py.query("SELECT count(*) FROM lineitem WHERE l_quantity < 24")[0][0],
        "count_shipdate_1994": ref_py.query("SELECT count(*) FROM lineitem WHERE l_shipdate >= DATE '1994-01-01' AND "
                                            "l_shipdate < DATE '1995-01-01'")[0][0],
        "sum_quantity": ref_py.query("SELECT sum(l_quantity) FROM lineitem")[0][0],
        "max_comment": ref_py.query("SELECT max(l_comment) FROM lineitem")[0][0],
        "first_row": ref_py.query("SELECT * FROM lineitem LIMIT 1")[0],
    }
    with open(os.path.join(OUT, "answers.json"), "w") as f:
        json.dump(answers, f, indent=1)
    print(path, rows, size)


if __name__ == "__main__":
    main()
