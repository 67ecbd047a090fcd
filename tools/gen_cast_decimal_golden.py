"""tests/golden/cast_decimal.json: the REFERENCE'S results (the DuckDB compiled from /root/reference/duckdb into oracle/_ref) for
the casts into and out of DECIMAL that the multi-file mapper runs on the device — integer -> DECIMAL / HUGEINT, DECIMAL ->
DECIMAL (scale up, scale down with rounding, both with their range checks), DECIMAL -> integer, FLOAT / DOUBLE -> DECIMAL
(raw = the bit patterns of the source values), DECIMAL -> DOUBLE / FLOAT (results = bit patterns) — for the
GPU test in tests/test_multifile.py; /root/reference does not exist on the GPU box, the fixture travels.  Rows on which the
reference's cast fails carry null.   Run in the build container:  python tools/gen_cast_decimal_golden.py"""
import json
import os
import sys

import numpy as np
import pyarrow as pa

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_py as R  # noqa: E402
from test_cast_parse import INT_RANGE, _export_column, _unscaled, arrow_array_of, decimal_values, real_values  # noqa: E402

INTS = {"TINYINT": pa.int8(), "SMALLINT": pa.int16(), "INTEGER": pa.int32(), "BIGINT": pa.int64(), "UTINYINT": pa.uint8(),
        "USMALLINT": pa.uint16(), "UINTEGER": pa.uint32(), "UBIGINT": pa.uint64()}


def main():
    rng = np.random.default_rng(99)
    cases = []
    k = 0

    def run(src, dst, raw, arr):
        nonlocal k
        k += 1
        R.register(f"t_gd_{k}", pa.table({"i": pa.array(range(len(raw)), pa.int64()), "c": arr}))
        col = _export_column(f"SELECT i, TRY_CAST(c AS {dst}) AS r FROM t_gd_{k}")
        if dst in ("DOUBLE", "FLOAT"):
            bits = col.to_numpy(zero_copy_only=False).view(np.uint64 if dst == "DOUBLE" else np.uint32)
            result = [str(int(b)) for b in bits]
        elif dst in INT_RANGE:
            result = [None if v is None else str(v) for v in col.to_pylist()]
        else:
            result = [None if v is None else str(v) for v in _unscaled(col)]
        cases.append({"from": src, "to": dst, "raw": [str(v) for v in raw], "result": result})

    for name, t in INTS.items():
        info = np.iinfo(t.to_pandas_dtype())
        vals = [int(info.min), int(info.max), 0, 1, 99, 100, 999, 1000, 9999, 10000] + \
               [int(x) for x in rng.integers(info.min, info.max, 40, dtype=t.to_pandas_dtype(), endpoint=True)] + \
               [int(x) for x in rng.integers(max(info.min, -100000), min(info.max, 100000), 40, endpoint=True)]
        vals = [v for v in vals if info.min <= v <= info.max]
        for dst in ["DECIMAL(4,1)", "DECIMAL(9,2)", "DECIMAL(18,3)", "DECIMAL(22,2)", "DECIMAL(38,19)"]:
            run(name, dst, vals, pa.array(vals, t))
    pairs = [((4, 2), (9, 4)), ((4, 2), (4, 3)), ((4, 2), (3, 2)), ((4, 2), (4, 0)), ((9, 4), (18, 6)), ((9, 4), (5, 1)),
             ((15, 2), (18, 4)), ((15, 2), (38, 10)), ((15, 2), (15, 0)), ((15, 2), (12, 1)), ((18, 6), (9, 2)), ((18, 0), (38, 20)),
             ((38, 10), (38, 0)), ((38, 10), (38, 12)), ((38, 10), (18, 2)), ((25, 5), (20, 0)), ((38, 37), (38, 1)), ((20, 10), (4, 1))]
    for (sw, ss), (dw, ds) in pairs:
        vals = decimal_values(rng, sw, ss, 60)
        run(f"DECIMAL({sw},{ss})", f"DECIMAL({dw},{ds})", vals, arrow_array_of(f"DECIMAL({sw},{ss})", vals))
    for w, s in [(4, 2), (9, 3), (15, 2), (18, 4), (18, 18), (20, 5), (38, 10), (38, 38)]:
        vals = decimal_values(rng, w, s, 80)
        for dst in ("DOUBLE", "FLOAT"):
            run(f"DECIMAL({w},{s})", dst, vals, arrow_array_of(f"DECIMAL({w},{s})", vals))
    for w, s in [(4, 2), (9, 0), (15, 2), (18, 4), (20, 5), (38, 10)]:
        vals = decimal_values(rng, w, s, 50)
        for dst in INT_RANGE:
            run(f"DECIMAL({w},{s})", dst, vals, arrow_array_of(f"DECIMAL({w},{s})", vals))
    for w, s in [(4, 2), (9, 2), (9, 9), (15, 2), (18, 6), (25, 5), (38, 10)]:
        for as_float in (False, True):
            arr = real_values(rng, w, s, 40, as_float)
            bits = arr.view(np.uint32 if as_float else np.uint64)
            # (the raw column holds the BIT PATTERNS of the source values)
            run("FLOAT" if as_float else "DOUBLE", f"DECIMAL({w},{s})", [int(b) for b in bits], pa.array(arr))
    doc = {"source": "TRY_CAST(c AS <to>) of oracle/_ref (reference DuckDB), tools/gen_cast_decimal_golden.py", "cases": cases}
    path = os.path.join(ROOT, "tests", "golden", "cast_decimal.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=0)
    print(path, len(cases), "cases", sum(len(c["raw"]) for c in cases), "values",
          sum(r is None for c in cases for r in c["result"]), "failing casts")


if __name__ == "__main__":
    if not R.available():
        sys.exit("oracle/_ref is not built")
    main()
