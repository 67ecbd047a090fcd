"""Multi-file read_arrow (src/file_scanner): binding on the host (CPU tests) and the scanned rows on the GPU, pinned
to the expectations of the reference's test/sql/multifile_reading.test (files under tests/golden/data/multifile are
the reference's own fixtures)."""
import math
import os

import numpy as np
import pyarrow as pa
import pytest

from logical import column_to_py

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "data")
MF = os.path.join(GOLDEN, "multifile")

PA_TYPE = {"VARCHAR": pa.string(), "DOUBLE": pa.float64(), "FLOAT": pa.float32(), "BIGINT": pa.int64(),
           "INTEGER": pa.int32(), "SMALLINT": pa.int16(), "TINYINT": pa.int8(), "UBIGINT": pa.uint64(),
           "UINTEGER": pa.uint32(), "USMALLINT": pa.uint16(), "UTINYINT": pa.uint8(), "TIMESTAMP": pa.int64(),
           "DATE": pa.int32()}


def rows_of(mfs):
    """materialise the scan as Python tuples, file by file, batch by batch (the reference's output order for one
    thread)"""
    out = []
    for f in mfs.files:
        f.scan()
        for b in range(f.reader.num_batches):
            cols = []
            for gc in mfs.columns:
                v = f.column(b, gc.index)
                if hasattr(v, "value"):  # ConstantColumn
                    val = v.value
                    if val is not None and gc.type_id == "BIGINT":
                        val = int(val)
                    cols.append([val] * v.rows)
                else:
                    vals = column_to_py(v, PA_TYPE[gc.type_id])
                    cols.append([x.decode() if isinstance(x, bytes) else x for x in vals])
            out += list(zip(*cols))
    return out


def mf(*names):
    return [os.path.join(MF, n) for n in names]


# ----------------------------------------------------------------------------------------------- host binding (CPU)
def test_bind_first_file_schema(b2a):
    s = b2a.read_arrow(mf("glob/f1.arrow", "different_order.arrows"))
    assert s.names == ["fruit", "variety", "weight"]
    assert s.types == ["VARCHAR", "VARCHAR", "DOUBLE"]
    m = [s.files[1].map(i) for i in range(3)]
    assert [x.kind for x in m] == [0, 0, 0]  # all LOCAL, matched by name
    # different_order.arrows is (fruit, weight, variety)
    assert [x.local_flat for x in m] == [0, 2, 1]


def test_bind_glob_sorted_and_counts(b2a):
    s = b2a.read_arrow(os.path.join(MF, "glob", "*.arrow"))
    assert [os.path.basename(f.path) for f in s.files] == ["f1.arrow", "f2.arrow", "f3.arrow"]
    assert s.num_rows == 6
    s2 = b2a.read_arrow([os.path.join(GOLDEN, "test.arrows")] * 2)
    assert s2.num_rows == 30974  # multifile_reading.test:11-14


def test_bind_schema_mismatch_errors(b2a):
    # multifile_reading.test:49-57
    for files in ([os.path.join(GOLDEN, "test.arrows"), mf("glob/f1.arrow")[0]],
                  mf("fruit_extra.arrows", "glob/f1.arrow")):
        with pytest.raises(b2a.B2AError, match="If you are trying to read files with different schemas, try setting "
                                               "union_by_name=True"):
            b2a.read_arrow(files)


def test_bind_union_by_name(b2a):
    s = b2a.read_arrow(mf("fruit_extra.arrows", "glob/f1.arrow"), union_by_name=True)
    assert s.names == ["fruit", "variety", "weight", "tasteness"]
    assert s.files[1].map(3).kind == 2  # constant NULL
    # multifile_reading.test:104-112: typeof(weight)
    three = mf("different_type_int.arrows", "glob/f1.arrow", "different_type.arrows")
    assert b2a.read_arrow(three, union_by_name=True).types[0] == "VARCHAR"
    assert b2a.read_arrow(three).types[0] == "BIGINT"
    assert b2a.read_arrow(mf("different_type.arrows", "glob/f1.arrow")).types[2] == "VARCHAR"
    assert b2a.read_arrow(mf("glob/f1.arrow", "different_type.arrows")).types[2] == "DOUBLE"


def test_bind_filename_and_hive(b2a):
    s = b2a.read_arrow(os.path.join(MF, "glob", "*.arrow"), filename=True)
    assert s.names[-1] == "filename" and s.types[-1] == "VARCHAR"
    assert s.files[2].map(3).constant.decode().endswith("glob/f3.arrow")
    h = b2a.read_arrow(os.path.join(MF, "hive", "*", "*.arrow"), hive_partitioning=True)
    assert h.names == ["fruit", "variety", "weight", "part"] and h.types[-1] == "VARCHAR"
    assert [f.map(3).constant.decode() for f in h.files] == ["a", "a", "b", "b"]
    with pytest.raises(b2a.B2AError, match="Option filename adds column"):
        b2a.read_arrow(os.path.join(MF, "glob", "*.arrow"), filename="fruit")
    with pytest.raises(b2a.B2AError, match="No files found that match the pattern"):
        b2a.read_arrow(os.path.join(MF, "glob", "*.nothing"))


def test_bind_projection_narrows_readers(b2a):
    s = b2a.read_arrow(os.path.join(MF, "glob", "*.arrow"), columns=["weight", "variety"])
    assert s.names == ["weight", "variety"]
    assert s.files[0].reader.projection is None  # python mirror untouched; the C++ reader holds the projection
    with pytest.raises(b2a.B2AError, match="not found in FROM clause"):
        b2a.read_arrow(os.path.join(MF, "glob", "*.arrow"), columns=["nope"])


def test_hive_type_autodetect(b2a, tmp_path):
    import pyarrow.ipc as ipc
    t = pa.table({"x": pa.array([1, 2], pa.int32())})
    for year in (2023, 2024):
        d = tmp_path / f"year={year}" / "day=2024-01-02"
        d.mkdir(parents=True)
        with ipc.new_stream(str(d / "p.arrows"), t.schema) as w:
            w.write_table(t)
    s = b2a.read_arrow(str(tmp_path / "*" / "*" / "*.arrows"), hive_partitioning=True)
    assert dict(zip(s.names, s.types)) == {"x": "INTEGER", "day": "DATE", "year": "BIGINT"}
    s = b2a.read_arrow(str(tmp_path / "*" / "*" / "*.arrows"), hive_partitioning=True, hive_types_autocast=False)
    assert dict(zip(s.names, s.types)) == {"x": "INTEGER", "day": "VARCHAR", "year": "VARCHAR"}


# ----------------------------------------------------------------------------------------------- scanned rows (GPU)
F1 = [("apple", "gala", 134.2), ("orange", "navel", 142.1)]
F2 = [("apple", "honeycrisp", 158.6), ("orange", "valencia", 96.7)]
F3 = [("apple", "fuji", None), ("orange", "cara cara", None)]


@pytest.mark.gpu
def test_glob_rows(b2a, cuda_device):
    # multifile_reading.test:17-25
    assert rows_of(b2a.read_arrow(os.path.join(MF, "glob", "*.arrow"))) == F1 + F2 + F3
    # :28-36 projection
    got = rows_of(b2a.read_arrow(os.path.join(MF, "glob", "*.arrow"), columns=["weight", "variety"]))
    assert got == [(w, v) for _, v, w in F1 + F2 + F3]


@pytest.mark.gpu
def test_file_list_count(b2a, cuda_device):
    s = b2a.read_arrow([os.path.join(GOLDEN, "test.arrows")] * 2, columns=["message"])
    assert len(rows_of(s)) == 30974
    assert sum(f.column(b, s.columns[0].index).rows for f in s.files for b in range(f.reader.num_batches)) == 30974


@pytest.mark.gpu
def test_union_by_name_rows(b2a, cuda_device):
    # multifile_reading.test:60-67
    got = rows_of(b2a.read_arrow(mf("fruit_extra.arrows", "glob/f1.arrow"), union_by_name=True))
    assert got == [("apple", "pink lady", 2.2, 10.0), ("orange", "jiha", None, None),
                   ("apple", "gala", 134.2, None), ("orange", "navel", 142.1, None)]


@pytest.mark.gpu
def test_different_column_order_rows(b2a, cuda_device):
    # multifile_reading.test:70-77 (ORDER BY ALL)
    got = rows_of(b2a.read_arrow(mf("different_order.arrows", "glob/f1.arrow")))
    assert got == [("apple", 2.2, "pink lady"), ("orange", None, "jiha"), ("apple", 134.2, "gala"),
                   ("orange", 142.1, "navel")]


@pytest.mark.gpu
def test_different_types_rows(b2a, cuda_device):
    # multifile_reading.test:80-96: VARCHAR global <- DOUBLE file, and DOUBLE global <- VARCHAR file
    got = rows_of(b2a.read_arrow(mf("different_type.arrows", "glob/f1.arrow")))
    assert got == [("apple", "pink lady", "2.2"), ("orange", "jiha", None), ("apple", "gala", "134.2"),
                   ("orange", "navel", "142.1")]
    got = rows_of(b2a.read_arrow(mf("glob/f1.arrow", "different_type.arrows")))
    assert got == F1 + [("apple", "pink lady", 2.2), ("orange", "jiha", None)]
    # :114-120 different type AND order
    got = rows_of(b2a.read_arrow(mf("glob/f1.arrow", "different_type_order.arrows")))
    assert got == F1 + [("apple", "pink lady", 2.2), ("orange", "jiha", None)]
    # :104-112: three files, BIGINT global from the first file: DOUBLE -> BIGINT rounds, VARCHAR -> BIGINT rounds
    three = mf("different_type_int.arrows", "glob/f1.arrow", "different_type.arrows")
    got = rows_of(b2a.read_arrow(three))
    assert got == [(2, "apple", "pink lady"), (8, "orange", "jiha"), (134, "apple", "gala"), (142, "orange", "navel"),
                   (2, "apple", "pink lady"), (None, "orange", "jiha")]
    got = rows_of(b2a.read_arrow(three, union_by_name=True))
    assert [r[0] for r in got] == ["2", "8", "134.2", "142.1", "2.2", None]


@pytest.mark.gpu
def test_filename_and_hive_rows(b2a, cuda_device):
    # multifile_reading.test:123-143
    got = rows_of(b2a.read_arrow(os.path.join(MF, "glob", "*.arrow"), filename=True))
    assert [(r[0], r[1], r[2], os.path.relpath(r[3], GOLDEN)) for r in got] == \
        [r + (f"multifile/glob/f{i}.arrow",) for i, fr in ((1, F1), (2, F2), (3, F3)) for r in fr]
    got = rows_of(b2a.read_arrow(os.path.join(MF, "hive", "*", "*.arrow"), hive_partitioning=True))
    assert got == [r + ("a",) for r in F1 + F2] + [r + ("b",) for r in F1 + F3]


def _write(path, table, batch=None):
    import pyarrow.ipc as ipc
    with ipc.new_stream(path, table.schema) as w:
        w.write_table(table, max_chunksize=batch)


@pytest.mark.gpu
def test_numeric_casts_against_numpy(b2a, cuda_device, tmp_path):
    """every numeric widening / float->int / int->varchar cast of the device kernel vs numpy / Python (DuckDB semantics
    restated: static_cast for widening, nearbyint for FLOAT/DOUBLE -> integer, decimal text for integer -> VARCHAR)"""
    rng = np.random.default_rng(5)
    n = 5000
    i8 = rng.integers(-128, 128, n, dtype=np.int8)
    u16 = rng.integers(0, 65536, n, dtype=np.uint16)
    i32 = rng.integers(-2**31, 2**31, n, dtype=np.int32)
    i64 = rng.integers(-2**62, 2**62, n, dtype=np.int64)
    f32 = rng.standard_normal(n).astype(np.float32) * 1000
    f64 = np.concatenate([rng.standard_normal(n - 6) * 1e6, [0.5, 1.5, 2.5, -0.5, -1.5, 1e15 + 0.5]])
    mask = rng.random(n) < 0.1
    first = pa.table({"a": pa.array(np.zeros(1, np.int64)), "b": pa.array(np.zeros(1, np.float64)),
                      "c": pa.array(np.zeros(1, np.int64)), "d": pa.array(["x"]), "e": pa.array(np.zeros(1, np.float64)),
                      "f": pa.array(np.zeros(1, np.int32)), "g": pa.array(np.zeros(1, np.float64)),
                      "h": pa.array(["x"])})
    second = pa.table({"a": pa.array(i8, mask=mask), "b": pa.array(i32, mask=mask), "c": pa.array(f64, mask=mask),
                       "d": pa.array(i64, mask=mask), "e": pa.array(f32, mask=mask), "f": pa.array(u16, mask=mask),
                       "g": pa.array(i64, mask=mask), "h": pa.array(u16, mask=mask)})
    _write(str(tmp_path / "a.arrows"), first)
    _write(str(tmp_path / "b.arrows"), second, batch=2100)
    s = b2a.read_arrow([str(tmp_path / "a.arrows"), str(tmp_path / "b.arrows")])
    assert s.types == ["BIGINT", "DOUBLE", "BIGINT", "VARCHAR", "DOUBLE", "INTEGER", "DOUBLE", "VARCHAR"]
    got = rows_of(s)[1:]
    exp = {"a": i8.astype(np.int64), "b": i32.astype(np.float64), "c": np.rint(f64).astype(np.int64),
           "d": [str(int(x)) for x in i64], "e": f32.astype(np.float64), "f": u16.astype(np.int32),
           "g": i64.astype(np.float64), "h": [str(int(x)) for x in u16]}
    for ci, name in enumerate("abcdefgh"):
        for r in range(n):
            want = None if mask[r] else exp[name][r]
            have = got[r][ci]
            if want is None:
                assert have is None
            elif isinstance(want, str):
                assert have == want, (name, r)
            else:
                assert have == want, (name, r, have, want)


@pytest.mark.gpu
def test_double_to_varchar_shortest_roundtrip(b2a, cuda_device, tmp_path):
    """DOUBLE / FLOAT -> VARCHAR is fmt's "{}" (shortest digits that round-trip; fixed notation for 1e-4 <= |v| < 1e16,
    else exponent form with a sign and >= 2 digits) — the same digits and layout as Python's repr(float); the
    reference pins '134.2' / '142.1' / '2.2' (multifile_reading.test:80-96)."""
    rng = np.random.default_rng(11)
    special = [0.0, -0.0, 1.0, -1.0, 0.1, 0.2, 0.3, 1e15, 1e16, 123456789012345680.0, 1e-4, 1e-5, 1.5e-5, 5e-324,
               2.2250738585072014e-308, 1.7976931348623157e308, 9007199254740993.0, 1e22, 1e23, 0.30000000000000004,
               float("inf"), float("-inf"), float("nan"), 134.2, 142.1, 2.2, 100.0, 1e100, 4.35, 0.000123456]
    bits = rng.integers(0, 2**63, 4000, dtype=np.uint64)
    rand = bits.view(np.float64)
    rand = rand[np.isfinite(rand)]
    vals = np.concatenate([np.array(special), rand, rng.standard_normal(2000), rng.integers(-10**6, 10**6, 1000) / 100.0])
    f32 = np.concatenate([rng.standard_normal(1000).astype(np.float32), np.array([0.1, 3.4028235e38, 1e-45, 16777216.0],
                                                                                  np.float32)])
    f32 = np.resize(f32, vals.size)
    _write(str(tmp_path / "a.arrows"), pa.table({"v": pa.array(["seed"]), "w": pa.array(["seed"])}))
    _write(str(tmp_path / "b.arrows"), pa.table({"v": pa.array(vals), "w": pa.array(f32)}), batch=3000)
    got = rows_of(b2a.read_arrow([str(tmp_path / "a.arrows"), str(tmp_path / "b.arrows")]))[1:]

    def fmt(x):
        if math.isnan(x):
            return "nan"
        return repr(float(x))

    def fmt32(x):
        if math.isnan(x):
            return "nan"
        # shortest digits that round-trip as FLOAT, laid out like repr
        s = np.format_float_scientific(np.float32(x), unique=True, trim="-")
        mant, e = s.split("e")
        digits = mant.replace("-", "").replace(".", "")
        e = int(e)
        sign = "-" if mant.startswith("-") else ""
        if -4 <= e < 16:
            if e < 0:
                return sign + "0." + "0" * (-e - 1) + digits
            ip = digits[: e + 1].ljust(e + 1, "0")
            fp = digits[e + 1:] or "0"
            return sign + ip + "." + fp
        return sign + digits[0] + ("." + digits[1:] if len(digits) > 1 else "") + "e" + ("-" if e < 0 else "+") + f"{abs(e):02d}"

    for r, (v, w) in enumerate(zip(vals, f32)):
        assert got[r][0] == fmt(v), (r, v, got[r][0], fmt(v))
        assert got[r][1] == fmt32(w), (r, w, got[r][1], fmt32(w))


@pytest.mark.gpu
def test_varchar_to_bigint_like_duckdb(b2a, cuda_device, tmp_path):
    """the device cast behind a BIGINT global column fed by a VARCHAR file = DuckDB's TrySimpleIntegerCast (the restatement
    in kernels/cast_parse.cuh is pinned against the compiled reference in tests/test_cast_parse.py; here: the kernel runs it)
    — exponents, '_' separators, hex / binary literals, fractions, the int64 limits; errors stay errors"""
    good = {"42": 42, " -7 ": -7, "1e3": 1000, "1.5e1": 15, "25e-1": 3, "-25e-1": -3, "1_000": 1000, "0x1F": 31,
            "0b101": 5, "2.5": 3, "-2.5": -3, "1.": 1, ".5": 1, "9223372036854775807": 2**63 - 1,
            "-9223372036854775808": -2**63, "12345e-2": 123, "+8": 8}
    _write(str(tmp_path / "a.arrows"), pa.table({"v": pa.array([1], pa.int64())}))
    _write(str(tmp_path / "b.arrows"), pa.table({"v": pa.array(list(good) + [None])}))
    got = [r[0] for r in rows_of(b2a.read_arrow([str(tmp_path / "a.arrows"), str(tmp_path / "b.arrows")]))[1:]]
    assert got == list(good.values()) + [None]
    for k, (text, what) in enumerate([("abc", "could not convert a string"), ("1e", "could not convert a string"),
                                      ("9223372036854775808", "out of range"), ("1e19", "out of range"),
                                      ("1__0", "could not convert a string")]):
        _write(str(tmp_path / f"bad{k}.arrows"), pa.table({"v": pa.array([text])}))
        with pytest.raises(b2a.B2AError, match=what):
            rows_of(b2a.read_arrow([str(tmp_path / "a.arrows"), str(tmp_path / f"bad{k}.arrows")]))


@pytest.mark.gpu
def test_varchar_to_double_equals_the_reference_or_loud(b2a, cuda_device, tmp_path):
    """the device cast (kernels/cast_parse_double.cuh: DuckDB's TryDoubleCast over fast_float, restated) against the bits the
    REFERENCE's CAST(s AS DOUBLE) gives for ~2 000 strings (tests/golden/cast_double.json, tools/gen_cast_double_golden.py):
    bit-exact, -0.0 and the NaN sign included; up to 800 digits (the big-decimal fallback) and '_' separators"""
    import json
    import struct
    with open(os.path.join(os.path.dirname(__file__), "golden", "cast_double.json")) as fh:
        doc = json.load(fh)
    texts = [s for s, _ in doc["values"]]
    _write(str(tmp_path / "a.arrows"), pa.table({"v": pa.array([1.0])}))
    _write(str(tmp_path / "b.arrows"), pa.table({"v": pa.array(texts + [None])}), batch=700)
    mfs = b2a.read_arrow([str(tmp_path / "a.arrows"), str(tmp_path / "b.arrows")])
    got = []  # the bits as the kernel wrote them (a Python float in between could lose a NaN's sign)
    for f in mfs.files:
        f.scan()
        for b in range(f.reader.num_batches):
            v = f.column(b, 0)
            bits = np.ascontiguousarray(v.data).view(np.uint64).ravel()
            got += [int(x) if ok else None for x, ok in zip(bits, v.valid)]
    assert got[0] == struct.unpack("<Q", struct.pack("<d", 1.0))[0]
    got = got[1:]
    assert len(got) == len(texts) + 1 and got[-1] is None
    bad = [(t, w, "%016x" % g) for (t, w), g in zip(doc["values"], got[:-1]) if g is None or "%016x" % g != w]
    assert not bad, f"{len(bad)} of {len(texts)} differ: {bad[:8]}"
    # where the reference itself reads past the end of the string: refuse loudly
    for k, (t, _) in enumerate(doc["undefined_in_reference"][:1]):
        _write(str(tmp_path / f"c{k}.arrows"), pa.table({"v": pa.array(["1.5", t])}))
        with pytest.raises(b2a.B2AError, match="not implemented on the device"):
            rows_of(b2a.read_arrow([str(tmp_path / "a.arrows"), str(tmp_path / f"c{k}.arrows")]))
    for k, t in enumerate(["1.12345678_9", "1e", "+", "nan(a-b)", "1_", "--1"]):
        assert t in doc["invalid"] or t == "+"
        _write(str(tmp_path / f"n{k}.arrows"), pa.table({"v": pa.array(["2", t])}))
        with pytest.raises(b2a.B2AError, match="could not convert a string"):
            rows_of(b2a.read_arrow([str(tmp_path / "a.arrows"), str(tmp_path / f"n{k}.arrows")]))
    _write(str(tmp_path / "d.arrows"), pa.table({"v": pa.array(["abc"])}))
    with pytest.raises(b2a.B2AError, match="could not convert a string"):
        rows_of(b2a.read_arrow([str(tmp_path / "a.arrows"), str(tmp_path / "d.arrows")]))
    _write(str(tmp_path / "e.arrows"), pa.table({"v": pa.array([1e300])}))
    _write(str(tmp_path / "i.arrows"), pa.table({"v": pa.array([1], pa.int32())}))
    with pytest.raises(b2a.B2AError, match="out of range"):
        rows_of(b2a.read_arrow([str(tmp_path / "i.arrows"), str(tmp_path / "e.arrows")]))


@pytest.mark.gpu
def test_temporal_decimal_boolean_to_varchar_equal_the_reference(b2a, cuda_device, tmp_path):
    """a VARCHAR column in the first file, BOOLEAN / DATE / TIME / TIMESTAMP / DECIMAL(w,s) in the next: the device cast
    (kernels/cast_format.cuh) against the text of the REFERENCE's CAST(c AS VARCHAR) (tests/golden/cast_to_varchar.json,
    tools/gen_cast_varchar_golden.py) — every physical decimal width, " (BC)" years, +-infinity, 24:00:00"""
    import json
    from test_cast_parse import arrow_array_of
    with open(os.path.join(os.path.dirname(__file__), "golden", "cast_to_varchar.json")) as fh:
        doc = json.load(fh)
    _write(str(tmp_path / "a.arrows"), pa.table({"v": pa.array(["first"])}))
    for k, (type_name, case) in enumerate(doc["types"].items()):
        raw = [int(x) for x in case["raw"]]
        arr = arrow_array_of(type_name, raw)
        arr = pa.concat_arrays([arr, pa.nulls(1, arr.type)])
        _write(str(tmp_path / f"t{k}.arrows"), pa.table({"v": arr}), batch=200)
        mfs = b2a.read_arrow([str(tmp_path / "a.arrows"), str(tmp_path / f"t{k}.arrows")])
        assert mfs.types == ["VARCHAR"]
        got = [r[0] for r in rows_of(mfs)]
        assert got[0] == "first" and got[-1] is None, type_name
        bad = [(v, w, g) for v, w, g in zip(raw, case["text"], got[1:-1]) if w != g]
        assert not bad and len(got) == len(raw) + 2, f"{type_name}: {len(bad)} differ, first {bad[:6]}"
    # the first partial day of the int64 range: the reference throws "Date out of range in timestamp conversion"
    low = [int(x) for x in doc["timestamp_out_of_range"]][:1]
    assert low
    _write(str(tmp_path / "low.arrows"), pa.table({"v": arrow_array_of("TIMESTAMP", [0] + low)}))
    with pytest.raises(b2a.B2AError, match="out of range"):
        rows_of(b2a.read_arrow([str(tmp_path / "a.arrows"), str(tmp_path / "low.arrows")]))


def test_union_by_name_numeric_and_decimal_types_match_the_reference(b2a, tmp_path):
    """LogicalType::ForceMaxLogicalType for every pair of 22 scalar types (types.cpp: CombineUnequalTypes, CombineEqualTypes,
    DecimalSizeCheck, CombineNumericTypes), bind only:
    the global type of a two-file union_by_name scan against typeof() of a UNION ALL of the same two types in the compiled
    reference (both go through MaxLogicalType's numeric rules)"""
    from oracle import ref_py as R
    if not R.available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    from test_cast_parse import INT_ARROW, arrow_array_of
    types = ["DECIMAL(15,2)", "DECIMAL(18,4)", "DECIMAL(4,0)", "DECIMAL(38,10)", "DECIMAL(38,0)", "DECIMAL(30,25)", "DECIMAL(9,9)",
             "TINYINT", "SMALLINT", "INTEGER", "BIGINT", "UTINYINT", "USMALLINT", "UINTEGER", "UBIGINT", "DOUBLE", "FLOAT",
             "BOOLEAN", "VARCHAR", "DATE", "TIMESTAMP", "TIME"]

    def array_of(t):
        if t in INT_ARROW:
            return pa.array([1], getattr(pa, INT_ARROW[t])())
        if t in ("DOUBLE", "FLOAT"):
            return pa.array([1.0], pa.float64() if t == "DOUBLE" else pa.float32())
        if t == "VARCHAR":
            return pa.array(["x"])
        return arrow_array_of(t, [1])

    for k, t in enumerate(types):
        _write(str(tmp_path / f"t{k}.arrows"), pa.table({"v": array_of(t)}))
    n = 0
    for i, a in enumerate(types):
        for j, b in enumerate(types):
            if i == j:
                continue
            want = R.query(f"SELECT typeof(x) FROM (SELECT CAST(NULL AS {a}) x UNION ALL SELECT CAST(NULL AS {b})) LIMIT 1")[0][0]
            try:
                got = b2a.read_arrow([str(tmp_path / f"t{i}.arrows"), str(tmp_path / f"t{j}.arrows")], union_by_name=True).types[0]
            except b2a.B2AError as e:  # the type is decided before the cast is looked up: the message names it
                assert ("not implemented on the device" in str(e) and f"to {want} " in str(e)) or \
                    "combining the column types" in str(e), (a, b, want, str(e))
                continue
            assert got == want, (a, b, got, want)
            n += 1
    assert n > 250


@pytest.mark.gpu
def test_casts_into_and_out_of_decimal_equal_the_reference(b2a, cuda_device, tmp_path):
    """the first file fixes the column's type, the second holds another one: integer / FLOAT / DOUBLE -> DECIMAL, DECIMAL ->
    DECIMAL (scale up / down with rounding), DECIMAL -> integer / DOUBLE / FLOAT on the device (kernels/cast_decimal.cuh) against the REFERENCE's results
    (tests/golden/cast_decimal.json, tools/gen_cast_decimal_golden.py): unscaled integers and floating-point bits, exact; a
    value the reference refuses to cast fails the scan with a conversion error"""
    import json
    from test_cast_parse import INT_ARROW, arrow_array_of
    with open(os.path.join(os.path.dirname(__file__), "golden", "cast_decimal.json")) as fh:
        doc = json.load(fh)

    def array_of(type_name, raw, bits=False):
        if type_name in INT_ARROW:
            return pa.array(raw, getattr(pa, INT_ARROW[type_name])())
        if type_name in ("DOUBLE", "FLOAT"):
            wide = type_name == "DOUBLE"
            if bits:  # the fixture holds the bit patterns of floating-point SOURCE values
                return pa.array(np.array(raw, dtype=np.uint64 if wide else np.uint32).view(np.float64 if wide else np.float32))
            return pa.array([float(v) for v in raw], pa.float64() if wide else pa.float32())
        return arrow_array_of(type_name, raw)

    def result_of(mfs, dst):
        out = []
        for f in mfs.files:
            f.scan()
            for b in range(f.reader.num_batches):
                v = f.column(b, 0)
                d = np.ascontiguousarray(v.data)
                if dst == "DOUBLE":
                    vals = d.view(np.uint64).ravel().tolist()
                elif dst == "FLOAT":
                    vals = d.view(np.uint32).ravel().tolist()
                elif dst in INT_ARROW:
                    vals = d.view(getattr(np, INT_ARROW[dst])).ravel().tolist()
                elif d.shape[1] == 16:
                    w = d.view(np.uint64).reshape(-1, 2)
                    vals = [(int(hi) << 64 | int(lo)) - ((1 << 128) if int(hi) >> 63 else 0) for lo, hi in w]
                else:
                    vals = d.view({2: np.int16, 4: np.int32, 8: np.int64}[d.shape[1]]).ravel().tolist()
                out += [int(x) if ok else None for x, ok in zip(vals, v.valid)]
        return out

    n = 0
    for k, case in enumerate(doc["cases"]):
        src, dst = case["from"], case["to"]
        first = str(tmp_path / f"first_{dst}.arrows")
        if not os.path.exists(first):
            _write(first, pa.table({"v": array_of(dst, [1])}))
        good = [(int(r), int(w)) for r, w in zip(case["raw"], case["result"]) if w is not None]
        bad = [int(r) for r, w in zip(case["raw"], case["result"]) if w is None]
        arr = array_of(src, [r for r, _ in good], bits=True)
        _write(str(tmp_path / f"c{k}.arrows"), pa.table({"v": pa.concat_arrays([arr, pa.nulls(1, arr.type)])}), batch=100)
        mfs = b2a.read_arrow([first, str(tmp_path / f"c{k}.arrows")])
        assert mfs.types == [dst]
        got = result_of(mfs, dst)[1:]
        assert got[-1] is None and len(got) == len(good) + 1, (src, dst)
        wrong = [(r, w, g) for (r, w), g in zip(good, got) if w != g]
        assert not wrong, f"{src} -> {dst}: {len(wrong)} of {len(good)} differ, first {wrong[:6]}"
        n += len(good)
        if bad:
            _write(str(tmp_path / f"b{k}.arrows"), pa.table({"v": array_of(src, [good[0][0] if good else 0, bad[0]], bits=True)}))
            with pytest.raises(b2a.B2AError, match="out of range"):
                rows = result_of(b2a.read_arrow([first, str(tmp_path / f"b{k}.arrows")]), dst)
                raise AssertionError(f"{src} -> {dst}: {bad[0]} was cast to {rows[-1]}")
    assert n > 13000


@pytest.mark.gpu
def test_varchar_to_date_and_timestamp_equal_the_reference(b2a, cuda_device, tmp_path):
    """a DATE / TIMESTAMP column in the first file, text in the second: the device parsers (kernels/cast_parse_temporal.cuh)
    against the REFERENCE's TRY_CAST (tests/golden/cast_temporal.json, tools/gen_cast_temporal_golden.py): BC years, the
    separators, infinity / epoch, UTC offsets, the limits of both ranges; a string the reference refuses fails the scan"""
    import json
    with open(os.path.join(os.path.dirname(__file__), "golden", "cast_temporal.json")) as fh:
        doc = json.load(fh)
    for sql_type, pa_type, np_type in (("DATE", pa.date32(), np.int32), ("TIMESTAMP", pa.timestamp("us"), np.int64)):
        case = doc[sql_type]
        first = str(tmp_path / f"first_{sql_type}.arrows")
        _write(first, pa.table({"v": pa.array([0], pa.int32() if sql_type == "DATE" else pa.int64()).cast(pa_type)}))
        good = [(s, w) for s, w in zip(case["strings"], case["result"]) if w is not None]
        bad = [s for s, w in zip(case["strings"], case["result"]) if w is None]
        assert len(good) > 400 and len(bad) > 300
        _write(str(tmp_path / f"good_{sql_type}.arrows"), pa.table({"v": pa.array([s for s, _ in good] + [None], pa.string())}), batch=300)
        mfs = b2a.read_arrow([first, str(tmp_path / f"good_{sql_type}.arrows")])
        assert mfs.types == [sql_type]
        got = []
        for f in mfs.files:
            f.scan()
            for b in range(f.reader.num_batches):
                v = f.column(b, 0)
                vals = np.ascontiguousarray(v.data).view(np_type).ravel().tolist()
                got += [int(x) if ok else None for x, ok in zip(vals, v.valid)]
        assert got[0] == 0 and got[-1] is None and len(got) == len(good) + 2
        wrong = [(s, w, g) for (s, w), g in zip(good, got[1:-1]) if w != g]
        assert not wrong, f"{sql_type}: {len(wrong)} of {len(good)} differ, first {wrong[:6]}"
        for k, s in enumerate(bad[:: max(1, len(bad) // 6)][:6]):
            _write(str(tmp_path / f"bad_{sql_type}_{k}.arrows"), pa.table({"v": pa.array([good[0][0], s], pa.string())}))
            with pytest.raises(b2a.B2AError, match="could not convert a string|out of range"):
                for f in b2a.read_arrow([first, str(tmp_path / f"bad_{sql_type}_{k}.arrows")]).files:
                    f.scan()
                    [f.column(b, 0) for b in range(f.reader.num_batches)]
                raise AssertionError(f"{sql_type}: {s!r} was accepted")


@pytest.mark.gpu
def test_varchar_to_decimal_equals_the_reference(b2a, cuda_device, tmp_path):
    """a DECIMAL(w,s) column in the first file, text in the second: the device parser (kernels/cast_parse_decimal.cuh) against
    the REFERENCE's TRY_CAST (tests/golden/cast_varchar_decimal.json): every physical width, rounding on the first digit past
    the scale, exponents, the width limit; a string the reference refuses fails the scan"""
    import json
    from test_cast_parse import arrow_array_of
    with open(os.path.join(os.path.dirname(__file__), "golden", "cast_varchar_decimal.json")) as fh:
        doc = json.load(fh)
    for k, (target, results) in enumerate(doc["targets"].items()):
        first = str(tmp_path / f"first{k}.arrows")
        _write(first, pa.table({"v": arrow_array_of(target, [1])}))
        good = [(s, int(w)) for s, w in zip(doc["strings"], results) if w is not None]
        bad = [s for s, w in zip(doc["strings"], results) if w is None]
        _write(str(tmp_path / f"good{k}.arrows"), pa.table({"v": pa.array([s for s, _ in good] + [None], pa.string())}), batch=250)
        mfs = b2a.read_arrow([first, str(tmp_path / f"good{k}.arrows")])
        assert mfs.types == [target]
        got = []
        for f in mfs.files:
            f.scan()
            for b in range(f.reader.num_batches):
                v = f.column(b, 0)
                d = np.ascontiguousarray(v.data)
                if d.shape[1] == 16:
                    w2 = d.view(np.uint64).reshape(-1, 2)
                    vals = [(int(hi) << 64 | int(lo)) - ((1 << 128) if int(hi) >> 63 else 0) for lo, hi in w2]
                else:
                    vals = d.view({2: np.int16, 4: np.int32, 8: np.int64}[d.shape[1]]).ravel().tolist()
                got += [int(x) if ok else None for x, ok in zip(vals, v.valid)]
        assert got[-1] is None and len(got) == len(good) + 2
        wrong = [(s, w, g) for (s, w), g in zip(good, got[1:-1]) if w != g]
        assert not wrong, f"{target}: {len(wrong)} of {len(good)} differ, first {wrong[:6]}"
        for j, s in enumerate(bad[:: max(1, len(bad) // 3)][:3]):
            _write(str(tmp_path / f"bad{k}_{j}.arrows"), pa.table({"v": pa.array([good[0][0], s], pa.string())}))
            with pytest.raises(b2a.B2AError, match="could not convert a string"):
                for f in b2a.read_arrow([first, str(tmp_path / f"bad{k}_{j}.arrows")]).files:
                    f.scan()
                    [f.column(b, 0) for b in range(f.reader.num_batches)]
                raise AssertionError(f"{target}: {s!r} was accepted")


@pytest.mark.gpu
def test_lineitem_glob_dictionary_union_by_name(b2a, cuda_device, tmp_path):
    """BASELINE config 4 in small: glob of lineitem files, dictionary-encoded strings, two schema variants under
    union_by_name (column order + INTEGER l_linenumber cast to BIGINT on the device); rows equal pyarrow's reading."""
    import pyarrow.ipc as ipc
    from tools.bench_multifile import write_files
    pattern = write_files(str(tmp_path), 30_000, 3, seed=9)
    s = b2a.read_arrow(pattern, union_by_name=True)
    assert s.types[3] == "BIGINT" and s.types[8] == "VARCHAR" and s.types[4] == "DECIMAL(15,2)"
    pa_t = dict(PA_TYPE)
    pa_t["DECIMAL"] = pa.decimal64(15, 2)
    got = []
    for f in s.files:
        f.scan()
        assert f.n_casts == (f.reader.num_batches if f.index % 2 else 0)
        for b in range(f.reader.num_batches):
            cols = []
            for gc in s.columns:
                v = f.column(b, gc.index)
                vals = column_to_py(v, pa.dictionary(pa.int32(), pa.string()) if v.is_dictionary else pa_t[gc.type_id])
                cols.append([x.decode() if isinstance(x, bytes) else x for x in vals])
            got += list(zip(*cols))
    exp = []
    for f in s.files:
        t = ipc.open_stream(f.path).read_all()
        cols = []
        for gc in s.columns:
            c = t.column(gc.name)
            if pa.types.is_dictionary(c.type):
                c = c.cast(pa.string())
            if pa.types.is_decimal(c.type):
                cols.append([int(x.scaleb(2)) for x in c.to_pylist()])
            elif pa.types.is_date(c.type):
                cols.append(c.cast(pa.int32()).to_pylist())
            else:
                cols.append(c.to_pylist())
        exp += list(zip(*cols))
    assert len(got) == len(exp) == 90_000
    assert got == exp
