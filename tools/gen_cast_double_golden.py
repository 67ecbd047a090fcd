"""tests/golden/cast_double.json: strings and the bits of the REFERENCE'S `CAST(s AS DOUBLE)` for each (the DuckDB compiled from
/root/reference/duckdb into oracle/_ref; TryDoubleCast over its vendored fast_float), for the GPU test of the multi-file
mapper's device cast (tests/test_multifile.py) — /root/reference does not exist on the GPU box, the fixture travels.
Includes strings the reference resolves with its big-decimal fallback (midpoints of adjacent doubles, up to ~770 digits) and
'_' separators; the one shape where the reference reads past the end of its input (see kernels/cast_parse_double.cuh) is
listed apart as "undefined_in_reference" — the device kernel reports those.
Run in the build container:  python tools/gen_cast_double_golden.py"""
import ctypes
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_py as R  # noqa: E402
from test_cast_parse import HAND_DOUBLE, _reference_double_bits  # noqa: E402


def main():
    so = os.path.join(ROOT, "tests", "native", "libcast_parse_host.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I", os.path.join(ROOT, "duckdb_arrow_b200", "src"),
                           os.path.join(ROOT, "tests", "native", "cast_parse_host.cpp"), "-o", so])
    lib = ctypes.CDLL(so)
    lib.b2a_test_varchar_to_double.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_ulonglong)]
    lib.b2a_test_varchar_to_double.restype = ctypes.c_uint
    rng = np.random.default_rng(2024)
    strings = list(HAND_DOUBLE) + ["  1.5", "2.5  ", " \t-3e2 \n"]
    for _ in range(1200):
        nd = int(rng.integers(1, 24))
        digits = "".join(str(d) for d in rng.integers(0, 10, nd))
        cut = int(rng.integers(0, nd + 1))
        body = digits[:cut] + ("." + digits[cut:] if cut < nd else "")
        if body.startswith("."):
            body = "0" + body
        r = rng.random()
        exp = "" if r < 0.3 else "e" + str(int(rng.integers(-30, 30))) if r < 0.6 else "e" + str(int(rng.integers(-345, 330)))
        strings.append(("-" if rng.random() < 0.3 else "") + body + exp)
    for _ in range(600):
        bits = int(rng.integers(0, 2 ** 63 - 1)) & 0x7FEFFFFFFFFFFFFF
        strings.append(repr(float(np.frombuffer(np.uint64(bits).tobytes(), dtype=np.float64)[0])))
    for _ in range(60):  # more than 19 digits on a rounding boundary: the fallback's
        m = int(rng.integers(2 ** 52, 2 ** 53)) * 2 + 1
        strings.append(str((m << int(rng.integers(12, 20))) >> 1))
    from fractions import Fraction
    for _ in range(40):  # exact midpoints of adjacent doubles and their neighbours
        bits = int(rng.integers(1, 0x7FE0000000000000))
        lo, hi = (float(np.frombuffer(np.uint64(b).tobytes(), dtype=np.float64)[0]) for b in (bits, bits + 1))
        mid = (Fraction(lo) + Fraction(hi)) / 2
        e10 = 0
        while mid.denominator != 1:
            mid *= 10
            e10 -= 1
        text = str(mid.numerator)
        strings += [text + "e" + str(e10), text + "1e" + str(e10 - 1)]
    strings += ["1" + "0" * 800, "0." + "0" * 800 + "1", "9" * 800, "1_2_3_4_5_6_7_8_9_0_1e5", "1_2_3_4_5_6_7_8_9_0_1 ",
                "1_2_3_4_5_6_7_8_9_0_1", "1_234567890123456789012345", "1234567890123456789012_345.5", "1.2345678901234567890_12345"]
    strings = list(dict.fromkeys(strings))
    want = _reference_double_bits(strings)
    values, invalid, fallback = [], [], []
    for s, w in zip(strings, want):
        b = s.encode()
        out = ctypes.c_ulonglong(0)
        rc = lib.b2a_test_varchar_to_double(b, len(b), ctypes.byref(out))
        if w is None:
            assert rc == 2, (s, rc)
            invalid.append(s)
        elif rc == 3:
            assert "_" in s, s
            fallback.append([s, "%016x" % w])
        else:
            assert rc == 0 and out.value == w, (s, rc, hex(out.value), hex(w))
            values.append([s, "%016x" % w])
    doc = {"source": "SELECT TRY_CAST(s AS DOUBLE) of oracle/_ref (reference DuckDB), tools/gen_cast_double_golden.py",
           "values": values, "invalid": invalid, "undefined_in_reference": fallback}
    path = os.path.join(ROOT, "tests", "golden", "cast_double.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=0)
    print(path, len(values), "values", len(invalid), "invalid", len(fallback), "undefined in the reference")


if __name__ == "__main__":
    if not R.available():
        sys.exit("oracle/_ref is not built")
    main()
