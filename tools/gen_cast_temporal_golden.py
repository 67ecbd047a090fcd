"""tests/golden/cast_temporal.json: strings and the REFERENCE'S `TRY_CAST(s AS DATE)` / `TRY_CAST(s AS TIMESTAMP)` for each (days /
microseconds since the epoch, null where the cast fails; the DuckDB compiled from /root/reference/duckdb into oracle/_ref), for
the GPU test of the multi-file mapper's device cast (tests/test_multifile.py) — /root/reference does not exist on the GPU box,
the fixture travels.   Run in the build container:  python tools/gen_cast_temporal_golden.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_py as R  # noqa: E402
from test_cast_parse import _reference_temporal, temporal_strings  # noqa: E402


def main():
    rng = np.random.default_rng(123)
    dates, stamps = temporal_strings(rng, 700)
    doc = {"source": "TRY_CAST(s AS DATE / TIMESTAMP) of oracle/_ref (reference DuckDB), tools/gen_cast_temporal_golden.py"}
    for sql_type, strings in (("DATE", dates + stamps[:300]), ("TIMESTAMP", stamps + dates[:300])):
        strings = list(dict.fromkeys(strings))
        doc[sql_type] = {"strings": strings, "result": _reference_temporal(strings, sql_type, "gold_" + sql_type.lower())}
    path = os.path.join(ROOT, "tests", "golden", "cast_temporal.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=0)
    print(path, {t: (len(doc[t]["strings"]), sum(r is None for r in doc[t]["result"])) for t in ("DATE", "TIMESTAMP")})


if __name__ == "__main__":
    if not R.available():
        sys.exit("oracle/_ref is not built")
    main()
