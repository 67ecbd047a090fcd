"""VARCHAR -> BIGINT of the multi-file mapper's device cast (src/kernels/cast_parse.cuh, called by b2a_cast_kernel) pinned to
the REFERENCE'S OWN cast: the header is `__host__ __device__`, compiled here with g++ (tests/native/cast_parse_host.cpp), and
every string goes through `SELECT TRY_CAST(s AS BIGINT)` of the DuckDB compiled from /root/reference/duckdb (oracle/_ref):
TrySimpleIntegerCast<int64_t> (duckdb/src/include/duckdb/common/operator/integer_cast_operator.hpp:20-458) — plain integers,
'_' separators, 0x / 0b literals, fractions (rounding on the first fraction digit), exponents, the int64 limits."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import ref_py as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built (needs /root/reference)")


@pytest.fixture(scope="module")
def parse():
    src = os.path.join(ROOT, "tests", "native", "cast_parse_host.cpp")
    so = os.path.join(ROOT, "tests", "native", "libcast_parse_host.so")
    hdr = os.path.join(ROOT, "duckdb_arrow_b200", "src", "kernels", "cast_parse.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I",
                               os.path.join(ROOT, "duckdb_arrow_b200", "src"), src, "-o", so])
    lib = ctypes.CDLL(so)
    lib.b2a_test_varchar_to_bigint.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_longlong)]
    lib.b2a_test_varchar_to_bigint.restype = ctypes.c_uint

    def f(s: str):
        b = s.encode()
        out = ctypes.c_longlong(0)
        rc = lib.b2a_test_varchar_to_bigint(b, len(b), ctypes.byref(out))
        return None if rc else int(out.value)
    return f


HAND = """0 1 -1 +1 -0 +0 007 42 -42 9223372036854775807 9223372036854775808 -9223372036854775808 -9223372036854775809
1. .1 . -. +. 1.4 1.5 -1.5 2.5 -2.5 0.5 -0.5 0.49999 .5 -.5 1.05 1.95 1.10 1.100 1.1000000000000000000001 99.95
9223372036854775807.4 9223372036854775807.5 -9223372036854775808.4 -9223372036854775808.5 9223372036854775806.5
1e3 1E3 1e+3 1e-3 15e-1 25e-1 -25e-1 1.5e1 1.25e1 1.25e2 1.25e3 12345e-2 12355e-2 5e-1 4e-1 1e0 1e 1e+ e5 .e5 1.e2 .5e1
1e18 9e18 1e19 -1e19 9.3e18 92233720368547758e2 1e-400 1e400 1e32767 1e32768 1e-32768 1.5e-1 0e10 0.0e10 1.23456789e5
1_000 1__000 _1 1_ 1_000_000 1_0.5 1.5_5 1.5_ 1_e3 1e1_0 1e_1 1e1_
0x10 0X1f 0xff_ff 0x 0x_1 0x1_ 0xg 0x7fffffffffffffff 0x8000000000000000 -0x10 +0x10 0x1.5 0x1e3
0b101 0B11 0b 0b2 0b1_0 0b_1 -0b1 0b111111111111111111111111111111111111111111111111111111111111111
0b1111111111111111111111111111111111111111111111111111111111111111
abc 1a a1 1-1 --1 +-1 1+ 12abc nan inf - + 1,000 1.2.3 1..2 ١٢٣""".split()


def _reference(strings):
    out = []
    for i in range(0, len(strings), 200):
        part = strings[i:i + 200]
        vals = ", ".join("(%d, '%s')" % (k, s.replace("'", "''")) for k, s in enumerate(part))
        rows = R.query(f"SELECT k, TRY_CAST(s AS BIGINT) IS NULL, TRY_CAST(s AS BIGINT) FROM (VALUES {vals}) t(k, s) ORDER BY k")
        assert len(rows) == len(part)
        out += [None if r[1] == "true" else int(r[2]) for r in rows]
    return out


def test_varchar_to_bigint_equals_the_reference_cast(parse):
    rng = np.random.default_rng(7)
    strings = list(HAND)
    strings += [" " + s for s in HAND[:40]] + [s + " " for s in HAND[:60]] + [s + "  \t" for s in HAND[:20]] + ["1 2", " ", ""]
    alphabet = "0123456789" * 3 + "..eE+-_ xXbB" + "af"
    for _ in range(3000):
        n = int(rng.integers(1, 12))
        strings.append("".join(alphabet[i] for i in rng.integers(0, len(alphabet), n)))
    for _ in range(1500):  # well-formed decimals / exponents around the rounding and range edges
        mant = str(int(rng.integers(0, 10 ** int(rng.integers(1, 19)))))
        frac = "" if rng.random() < 0.3 else "." + "".join(str(d) for d in rng.integers(0, 10, int(rng.integers(0, 6))))
        exp = "" if rng.random() < 0.4 else "e" + str(int(rng.integers(-22, 22)))
        strings.append(("-" if rng.random() < 0.4 else "") + mant + frac + exp)
    strings = [s for s in strings if "\\" not in s]
    want = _reference(strings)
    got = [parse(s) for s in strings]
    bad = [(s, w, g) for s, w, g in zip(strings, want, got) if w != g]
    assert not bad, f"{len(bad)} of {len(strings)} differ, first: {bad[:10]}"
    assert sum(w is not None for w in want) > 1000 and sum(w is None for w in want) > 500
