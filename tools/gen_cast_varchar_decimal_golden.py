"""tests/golden/cast_varchar_decimal.json: strings and the REFERENCE'S `TRY_CAST(s AS DECIMAL(w,s))` for several (w,s) (unscaled
integers, null where the cast fails; the DuckDB compiled from /root/reference/duckdb into oracle/_ref), for the GPU test of the
multi-file mapper's device cast (tests/test_multifile.py) — /root/reference does not exist on the GPU box, the fixture travels.
Run in the build container:  python tools/gen_cast_varchar_decimal_golden.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_py as R  # noqa: E402
from test_cast_parse import _reference_decimal, decimal_strings  # noqa: E402


def main():
    rng = np.random.default_rng(321)
    strings = decimal_strings(rng, 500)
    targets = {}
    for k, (w, s) in enumerate([(4, 2), (9, 0), (9, 9), (15, 2), (18, 6), (22, 4), (38, 10)]):
        targets[f"DECIMAL({w},{s})"] = [None if v is None else str(v) for v in _reference_decimal(strings, w, s, f"gold{k}")]
    doc = {"source": "TRY_CAST(s AS DECIMAL(w,s)) of oracle/_ref (reference DuckDB), tools/gen_cast_varchar_decimal_golden.py",
           "strings": strings, "targets": targets}
    path = os.path.join(ROOT, "tests", "golden", "cast_varchar_decimal.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=0)
    print(path, len(strings), "strings", {t: sum(v is not None for v in r) for t, r in targets.items()})


if __name__ == "__main__":
    if not R.available():
        sys.exit("oracle/_ref is not built")
    main()
