"""tests/golden/cast_to_varchar.json: raw values (days, microseconds, unscaled decimals) of BOOLEAN / DATE / TIME / TIMESTAMP /
DECIMAL columns and the text the REFERENCE'S `CAST(c AS VARCHAR)` gives for each (the DuckDB compiled from
/root/reference/duckdb into oracle/_ref, fed through its own arrow scan), for the GPU test of the multi-file mapper's device
cast (tests/test_multifile.py) — /root/reference does not exist on the GPU box, the fixture travels.
Run in the build container:  python tools/gen_cast_varchar_golden.py"""
import json
import os
import sys

import pyarrow as pa

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_py as R  # noqa: E402
from test_cast_parse import cast_to_varchar_cases  # noqa: E402


def main():
    types, low = {}, []
    for k, (type_name, (arr, raw)) in enumerate(cast_to_varchar_cases(seed=77, n=240).items()):
        if type_name == "TIMESTAMP":  # the reference throws for the first partial day of the int64 range
            low = [v for v in raw if -(2 ** 63 - 1) < v < -9223372022400000000]
            keep = [i for i, v in enumerate(raw) if v not in set(low)]
            arr, raw = arr.take(pa.array(keep)), [raw[i] for i in keep]
        R.register(f"t_gold_{k}", pa.table({"i": pa.array(range(len(raw)), pa.int64()), "c": arr}))
        rows = R.query(f"SELECT i, typeof(c), CAST(c AS VARCHAR) FROM t_gold_{k} ORDER BY i")
        assert len(rows) == len(raw) and rows[0][1] == type_name
        types[type_name] = {"raw": [str(v) for v in raw], "text": [r[2] for r in rows]}
    doc = {"source": "CAST(c AS VARCHAR) of oracle/_ref (reference DuckDB), tools/gen_cast_varchar_golden.py",
           "types": types, "timestamp_out_of_range": [str(v) for v in low]}
    path = os.path.join(ROOT, "tests", "golden", "cast_to_varchar.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=0)
    print(path, {t: len(c["raw"]) for t, c in types.items()}, len(low), "out of range")


if __name__ == "__main__":
    if not R.available():
        sys.exit("oracle/_ref is not built")
    main()
