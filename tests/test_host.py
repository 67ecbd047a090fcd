"""CPU-only tests: the C-ABI library loads and exports what include/b2arrow.h declares, the host slicer agrees
with pyarrow on every fixture, framing / projection errors match the reference's messages, and nothing in the
product can compute without a GPU (no CPU fallback, no oracle on the product path)."""
import ctypes as C
import glob
import os
import re
import struct
import subprocess

import numpy as np
import pyarrow as pa
import pyarrow.ipc as ipc
import pytest

from parity import GOLDEN, O, ROOT

LE = os.path.join(GOLDEN, "arrow-testing", "1.0.0-littleendian")


def test_abi_exports_every_declared_symbol(b2a):
    hdr = open(os.path.join(ROOT, "include", "b2arrow.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(b2a_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 25
    lib = C.CDLL(b2a.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in b2arrow.h but not exported: {missing}"


def test_library_is_sm100a_and_in_tree(b2a):
    assert os.path.dirname(b2a.LIB_PATH) == os.path.join(ROOT, "duckdb_arrow_b200")
    out = subprocess.run(["cuobjdump", "-lelf", b2a.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_type_enums_in_sync(b2a):
    hdr = open(os.path.join(ROOT, "include", "b2arrow.h")).read()
    body = hdr[hdr.index("typedef enum b2a_type_id"):hdr.index("} b2a_type_id")]
    ids = re.findall(r"B2A_T_([A-Z_0-9]+)", body)
    assert [O.T[i] for i in ids] == list(range(len(ids)))


def test_product_never_touches_the_oracle():
    for path in glob.glob(os.path.join(ROOT, "duckdb_arrow_b200", "**", "*"), recursive=True) + \
            glob.glob(os.path.join(ROOT, "include", "*")):
        if os.path.isfile(path) and path.endswith((".py", ".cu", ".cpp", ".hpp", ".h", ".cuh")):
            src = open(path).read()
            assert "oracle" not in src.replace("no oracle", ""), f"{path} mentions the oracle"


def test_no_cpu_fallback(b2a):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(b2a.B2AError) as e:
        b2a.read_arrow(os.path.join(GOLDEN, "data", "test.arrows"))
    assert e.value.kind == "CudaError"
    r = b2a.ArrowIPCReader.open_file(os.path.join(GOLDEN, "data", "test.arrows"))
    plan_only = b2a.ArrowScan(r, device=-1)
    with pytest.raises(b2a.B2AError, match="no CPU fallback"):
        plan_only.upload()


def _c_format(field: pa.Field) -> str:
    sch = O.ArrowSchema()
    field._export_to_c(C.addressof(sch))
    fmt = sch.format.decode()
    C.CFUNCTYPE(None, C.POINTER(O.ArrowSchema))(sch.release)(C.byref(sch))
    return fmt


def _walk(field: pa.Field):
    yield field
    t = field.type
    if isinstance(t, pa.BaseExtensionType):
        t = t.storage_type
    if pa.types.is_dictionary(t):
        return
    for i in range(t.num_fields):
        yield from _walk(t.field(i))


ALL_STREAMS = sorted(glob.glob(os.path.join(GOLDEN, "arrow-testing", "1.0.0-*", "*.stream")))


@pytest.mark.parametrize("path", ALL_STREAMS, ids=lambda p: "/".join(p.split("/")[-2:]))
def test_slicer_agrees_with_pyarrow(b2a, path):
    name = os.path.basename(path)
    if "decimal256" in name:
        with pytest.raises(b2a.B2AError, match="Unsupported Internal Arrow Type for Decimal"):
            b2a.ArrowIPCReader.open_file(path)
        return
    r = b2a.ArrowIPCReader.open_file(path)
    pr = ipc.open_stream(path)
    schema = pr.schema
    assert [c.name for c in r.columns] == schema.names
    try:
        batches = list(pr)
    except (KeyError, pa.ArrowNotImplementedError):
        batches = None
    if batches is not None:
        assert r.num_batches == len(batches)
        assert r.num_rows == sum(len(b) for b in batches)
        for i, b in enumerate(batches):
            assert r.batch_info(i).rows == len(b)
    if "nested_dictionary" in name or "interval" in name:
        return
    flat = [f for top in schema for f in _walk(top)]
    assert len(flat) == r.n_flat
    for got, exp in zip(r.fields, flat):
        t = exp.type
        if isinstance(t, pa.BaseExtensionType):
            t = t.storage_type
        fmt = _c_format(pa.field("x", t.value_type if pa.types.is_dictionary(t) else t))
        assert got.format == fmt, f"{got.name}: {got.format} != {fmt}"
        assert got.name == exp.name
        assert (got.dict_id >= 0) == pa.types.is_dictionary(t)


def test_reader_quirks(b2a, tmp_path):
    data = open(os.path.join(GOLDEN, "data", "test.arrows"), "rb").read()
    # EOS marker missing / truncated prefix = clean end of stream (ipc_file_stream_reader.cpp:126-129)
    r = b2a.ArrowIPCReader.open_buffers([data[:-8]])
    assert r.num_rows == 15487
    r = b2a.ArrowIPCReader.open_buffers([data[:-8] + b"\xff\xff"])
    assert r.num_rows == 15487
    # wrong continuation token
    bad = bytearray(data)
    bad[0:4] = struct.pack("<I", 1234)
    with pytest.raises(b2a.B2AError, match=r"Expected continuation token \(0xFFFFFFFF\) but got 1234"):
        b2a.ArrowIPCReader.open_buffers([bytes(bad)])
    bad = bytearray(data)
    bad[4:8] = struct.pack("<i", -5)
    with pytest.raises(b2a.B2AError, match="Expected metadata size >= 0 but got -5"):
        b2a.ArrowIPCReader.open_buffers([bytes(bad)])
    with pytest.raises(b2a.B2AError, match="Expected Schema Arrow IPC message but got end of stream"):
        b2a.ArrowIPCReader.open_buffers([b"\xff\xff\xff\xff\x00\x00\x00\x00"])
    with pytest.raises(b2a.B2AError, match="extends past the end"):
        b2a.ArrowIPCReader.open_buffers([data[: len(data) // 2]])
    # Arrow FILE format: magic skipped, footer ignored (ipc_file_stream_reader.cpp:107-119)
    rf = b2a.ArrowIPCReader.open_file(os.path.join(GOLDEN, "data", "test.arrow"))
    assert rf.num_rows == 15487 and rf.num_batches == 16
    # several buffers = one stream (scan_arrow_ipc list argument)
    cut = 4096 * 8
    rb = b2a.ArrowIPCReader.open_buffers([data[:cut], data[cut:]]) if False else None
    assert rb is None


def test_projection_errors(b2a):
    r = b2a.ArrowIPCReader.open_file(os.path.join(GOLDEN, "data", "test.arrows"))
    with pytest.raises(b2a.B2AError, match="Field 'nope' does not exist in IPC file schema"):
        r.set_projection(["nope"])
    r.set_projection(["files", "commit"])
    assert [f.name for f in r.output_columns()] == ["files", "commit"]
    s = b2a.ArrowScan(r, device=-1)
    full = b2a.ArrowScan(b2a.ArrowIPCReader.open_file(os.path.join(GOLDEN, "data", "test.arrows")), device=-1)
    assert s.n_tasks < full.n_tasks and s.out_bytes < full.out_bytes
    rd = b2a.ArrowIPCReader.open_file(os.path.join(LE, "generated_duplicate_fieldnames.stream"))
    names = [c.name for c in rd.columns]
    dup = [n for n in names if names.count(n) > 1]
    if dup:
        # DeduplicateColumns renames the second occurrence, so the raw duplicate name resolves to the first column
        rd.set_projection([dup[0]])


def test_plan_layout_lineitem(b2a):
    from tools import lineitem
    buf = lineitem.stream_bytes(300_000, seed=1)
    r = b2a.ArrowIPCReader.open_buffers([buf])
    s = b2a.ArrowScan(r, device=-1)
    assert r.num_rows == 300_000 and r.num_batches == 3
    assert [f.op for f in r.columns] == ["COPY"] * 8 + ["STRING32"] * 2 + ["COPY"] * 3 + ["STRING32"] * 3
    # algorithmic bytes = projected IPC buffers read once + DuckDB vectors written once: 156 B/row out
    per_row = s.algorithmic_bytes / r.num_rows
    assert 290 < per_row < 305
    offs = []
    for b in range(r.num_batches):
        for f in r.columns:
            co = s.column(b, f.index)
            assert co.rows == r.batch_info(b).rows
            assert co.validity_off == b2a.B2A_NONE  # null_count == 0 -> mask untouched (arrow_conversion.cpp:71)
            assert co.data_off % 256 == 0
            offs.append((co.data_off, co.rows * f.out_width))
    offs.sort()
    for (a, n), (b_, _) in zip(offs, offs[1:]):
        assert a + n <= b_, "output slabs overlap"
    assert s.n_chunks == sum((r.batch_info(b).rows + 2047) // 2048 for b in range(r.num_batches))


def test_writer_schema_message_types_read_by_pyarrow(b2a):
    """Host half of COPY ... (FORMAT ARROWS): the Schema message for every writable DuckDB type carries the Arrow type
    ArrowConverter::ToArrowSchema gives it (duckdb/src/common/arrow/arrow_converter.cpp:115-395) — read back by pyarrow.
    No device needed (b2a_writer_schema_message)."""
    import pyarrow.ipc as ipc
    from duckdb_arrow_b200 import DeviceVector as DV

    def v(t, name, kids=(), **kw):
        return DV(t, 0, None, None, list(kids), name, **kw)

    cols = [
        v("BOOLEAN", "b"), v("TINYINT", "i8"), v("USMALLINT", "u16"), v("BIGINT", "i64"), v("HUGEINT", "h"),
        v("FLOAT", "f"), v("DOUBLE", "g"), v("DECIMAL", "d4", decimal=(4, 1)), v("DECIMAL", "d15", decimal=(15, 2)),
        v("DECIMAL", "d38", decimal=(38, 4)), v("VARCHAR", "s"), v("BLOB", "z"), v("BIT", "bits"), v("UUID", "id"),
        v("DATE", "dt"), v("TIME", "t"), v("TIME_TZ", "ttz"), v("TIMESTAMP", "ts"), v("TIMESTAMP_S", "ts_s"),
        v("TIMESTAMP_MS", "ts_ms"), v("TIMESTAMP_NS", "ts_ns"),
        v("TIMESTAMP WITH TIME ZONE", "tstz", timezone="Europe/Amsterdam"), v("INTERVAL", "iv"), v("SQLNULL", "n"),
        v("LIST", "l", [v("INTEGER", "x")]), v("STRUCT", "st", [v("INTEGER", "a"), v("VARCHAR", "b")]),
        v("MAP", "m", [v("STRUCT", "e", [v("VARCHAR", "k"), v("BIGINT", "v")])]),
        v("ARRAY", "arr", [v("SMALLINT", "x")], param=3),
        v("UNION", "u", [v("BIGINT", "num"), v("VARCHAR", "str"), v("BOOLEAN", "flag")]),
    ]
    w = b2a.ArrowWriter(cols, 0, kv_metadata={"k": "v"}, device=-1)
    sch = ipc.read_schema(pa.py_buffer(w.schema_message()))
    w.close()
    exp = {
        "b": pa.bool_(), "i8": pa.int8(), "u16": pa.uint16(), "i64": pa.int64(), "h": pa.decimal128(38, 0),
        "f": pa.float32(), "g": pa.float64(), "d4": pa.decimal32(4, 1), "d15": pa.decimal64(15, 2),
        "d38": pa.decimal128(38, 4), "s": pa.string(), "z": pa.binary(), "bits": pa.binary(), "id": pa.string(),
        "dt": pa.date32(), "t": pa.time64("us"), "ttz": pa.time64("us"), "ts": pa.timestamp("us"),
        "ts_s": pa.timestamp("s"), "ts_ms": pa.timestamp("ms"), "ts_ns": pa.timestamp("ns"),
        "tstz": pa.timestamp("us", tz="Europe/Amsterdam"), "iv": pa.month_day_nano_interval(), "n": pa.null(),
        "l": pa.list_(pa.field("l", pa.int32())), "st": pa.struct([("a", pa.int32()), ("b", pa.string())]),
        "m": pa.map_(pa.string(), pa.int64()), "arr": pa.list_(pa.field("", pa.int16()), 3),
        "u": pa.sparse_union([pa.field("num", pa.int64()), pa.field("str", pa.string()), pa.field("flag", pa.bool_())]),
    }
    assert sch.names == list(exp)
    for name, t in exp.items():
        assert sch.field(name).type == t, (name, sch.field(name).type)
    assert sch.field("u").type.type_codes == [0, 1, 2] and sch.field("u").type.mode == "sparse"
    assert sch.metadata == {b"k": b"v"}


def test_writer_schema_message_large_offsets(b2a):
    """B2A_WRITE_LARGE_OFFSETS = SET arrow_large_buffer_size: "U" / "Z" / "+L" for VARCHAR + UUID / BLOB + BIT / LIST, at
    every nesting level; MAP stays a (32-bit) map (arrow_converter.cpp:179-196,262-293; arrow_appender.cpp:296-299)."""
    import pyarrow.ipc as ipc
    from duckdb_arrow_b200 import DeviceVector as DV

    def v(t, name, kids=(), **kw):
        return DV(t, 0, None, None, list(kids), name, **kw)

    cols = [v("VARCHAR", "s"), v("BLOB", "z"), v("BIT", "bits"), v("UUID", "id"), v("BIGINT", "i64"),
            v("LIST", "l", [v("VARCHAR", "x")]), v("LIST", "ll", [v("LIST", "x", [v("INTEGER", "y")])]),
            v("STRUCT", "st", [v("INTEGER", "a"), v("VARCHAR", "b")]),
            v("MAP", "m", [v("STRUCT", "e", [v("VARCHAR", "k"), v("LIST", "v", [v("BIGINT", "w")])])]),
            v("ARRAY", "arr", [v("VARCHAR", "x")], param=2)]
    w = b2a.ArrowWriter(cols, 0, device=-1, arrow_large_buffer_size=True)
    sch = ipc.read_schema(pa.py_buffer(w.schema_message()))
    w.close()
    exp = {"s": pa.large_string(), "z": pa.large_binary(), "bits": pa.large_binary(), "id": pa.large_string(),
           "i64": pa.int64(), "l": pa.large_list(pa.field("l", pa.large_string())),
           "ll": pa.large_list(pa.field("l", pa.large_list(pa.field("l", pa.int32())))),
           "st": pa.struct([("a", pa.int32()), ("b", pa.large_string())]),
           "m": pa.map_(pa.large_string(), pa.large_list(pa.field("l", pa.int64()))),
           "arr": pa.list_(pa.field("", pa.large_string()), 2)}
    for name, t in exp.items():
        assert sch.field(name).type == t, (name, sch.field(name).type)


def test_writer_schema_message_list_views(b2a):
    """B2A_WRITE_LIST_VIEW = SET arrow_output_list_view: "+vl" (with large offsets "+vL") for LIST at every level, MAP
    untouched (arrow_converter.cpp:282-287)."""
    import pyarrow.ipc as ipc
    from duckdb_arrow_b200 import DeviceVector as DV

    def v(t, name, kids=(), **kw):
        return DV(t, 0, None, None, list(kids), name, **kw)

    def cols():
        return [v("LIST", "l", [v("VARCHAR", "x")]), v("LIST", "ll", [v("LIST", "x", [v("INTEGER", "y")])]),
                v("MAP", "m", [v("STRUCT", "e", [v("VARCHAR", "k"), v("LIST", "v", [v("BIGINT", "w")])])])]
    for large in (False, True):
        w = b2a.ArrowWriter(cols(), 0, device=-1, arrow_large_buffer_size=large, arrow_output_list_view=True)
        sch = ipc.read_schema(pa.py_buffer(w.schema_message()))
        w.close()
        lv = pa.large_list_view if large else pa.list_view
        s = pa.large_string() if large else pa.string()
        assert sch.field("l").type == lv(pa.field("l", s))
        assert sch.field("ll").type == lv(pa.field("l", lv(pa.field("l", pa.int32()))))
        assert sch.field("m").type == pa.map_(s, lv(pa.field("l", pa.int64())))


def test_writer_schema_message_string_views(b2a):
    """B2A_WRITE_STRING_VIEW = SET produce_arrow_string_view: VARCHAR -> "vu" at every level, BLOB keeps "z"
    (arrow_converter.cpp:189-197,262-267); a UUID column is refused (the reference's schema and appender disagree there)."""
    import pyarrow.ipc as ipc
    from duckdb_arrow_b200 import B2AError, DeviceVector as DV

    def v(t, name, kids=(), **kw):
        return DV(t, 0, None, None, list(kids), name, **kw)

    cols = [v("VARCHAR", "s"), v("BLOB", "z"), v("LIST", "l", [v("VARCHAR", "x")]),
            v("MAP", "m", [v("STRUCT", "e", [v("VARCHAR", "k"), v("BIGINT", "v")])])]
    w = b2a.ArrowWriter(cols, 0, device=-1, produce_arrow_string_view=True, arrow_large_buffer_size=True)
    sch = ipc.read_schema(pa.py_buffer(w.schema_message()))
    w.close()
    assert sch.field("s").type == pa.string_view() and sch.field("z").type == pa.large_binary()
    assert sch.field("l").type == pa.large_list(pa.field("l", pa.string_view()))
    assert sch.field("m").type == pa.map_(pa.string_view(), pa.int64())
    with pytest.raises(B2AError, match="UUID"):
        b2a.ArrowWriter([v("UUID", "id")], 0, device=-1, produce_arrow_string_view=True)


def test_arrow_file_footer_index_equals_the_walk(b2a, tmp_path):
    """an Arrow FILE is indexed from its footer blocks; a stream, or a file whose footer is damaged, by
    walking the message prefixes like the reference does (ipc_file_stream_reader.cpp:96-132).  Same messages either way."""
    import pyarrow as pa
    import pyarrow.ipc as ipc
    for rel in ("data/fruit.arrow", "data/test.arrow", "arrow-testing/1.0.0-littleendian/generated_primitive.arrow_file"):
        path = os.path.join(GOLDEN, rel)
        r = b2a.ArrowIPCReader.open_file(path)
        assert r.used_footer
        f = ipc.open_file(path)
        assert r.num_batches == f.num_record_batches
        assert [r.batch_info(i).rows for i in range(r.num_batches)] == \
            [f.get_batch(i).num_rows for i in range(f.num_record_batches)]
        # damage the footer (its length word, then a block offset): the reader falls back to the walk, same index
        raw = bytearray(open(path, "rb").read())
        ref = [(r.batch_info(i).meta_offset, r.batch_info(i).body_offset, r.batch_info(i).body_length, r.batch_info(i).rows)
               for i in range(r.num_batches)]
        for damage in ("length", "block"):
            b = bytearray(raw)
            if damage == "length":
                b[-10:-6] = (2**31 - 1).to_bytes(4, "little")
            else:
                import struct
                off = struct.pack("<q", ref[0][0])
                k = bytes(b).rfind(off)  # the first record-batch Block of the footer
                assert k > ref[-1][1]
                b[k:k + 8] = struct.pack("<q", ref[0][0] + 8)
            p = tmp_path / f"damaged_{damage}.arrow"
            p.write_bytes(bytes(b))
            r2 = b2a.ArrowIPCReader.open_file(str(p))
            assert not r2.used_footer
            got = [(r2.batch_info(i).meta_offset, r2.batch_info(i).body_offset, r2.batch_info(i).body_length,
                    r2.batch_info(i).rows) for i in range(r2.num_batches)]
            assert got == ref
    assert not b2a.ArrowIPCReader.open_file(os.path.join(GOLDEN, "data", "test.arrows")).used_footer
    # a dictionary-encoded file: the footer's dictionary blocks come first
    t = pa.table({"d": pa.array(["a", "b", "a", None] * 100).dictionary_encode(), "i": pa.array(range(400))})
    p = tmp_path / "dict.arrow"
    with ipc.new_file(str(p), t.schema) as w:
        w.write_table(t, max_chunksize=150)
    r = b2a.ArrowIPCReader.open_file(str(p))
    assert r.used_footer and r.num_rows == 400 and r.num_batches == 3


def test_large_file_mappings_are_reused_and_never_stale(b2a, tmp_path):
    """a file of >= 64 MiB keeps its mapping for the next open (ipc_reader.cpp: idle mappings, keyed by device / inode / size
    / mtime); the same bytes come back through a reused mapping, two readers open at once get a mapping each, and a file
    rewritten under the same name (other inode or mtime) is read fresh"""
    import pyarrow as pa
    import pyarrow.ipc as ipc

    def write(path, seed):
        rng = np.random.default_rng(seed)
        t = pa.table({"v": pa.array(rng.integers(0, 2**62, 9_000_000), pa.int64())})
        with ipc.new_stream(path, t.schema) as w:
            w.write_table(t, max_chunksize=1_000_000)
        return t

    def first_values(r):
        img = r.image()
        bi = r.batch_info(0)
        return bytes(img[bi.body_offset: bi.body_offset + 64])

    p = str(tmp_path / "big.arrows")
    t1 = write(p, 1)
    assert os.path.getsize(p) >= 64 << 20
    want1 = t1.column("v").chunk(0).buffers()[1].to_pybytes()[:64]
    r = b2a.ArrowIPCReader.open_file(p)
    assert r.num_batches == 9 and first_values(r) == want1
    r.close()
    ra, rb = b2a.ArrowIPCReader.open_file(p), b2a.ArrowIPCReader.open_file(p)  # one reuses the idle mapping, one maps anew
    assert first_values(ra) == want1 and first_values(rb) == want1
    ra.close()
    rb.close()
    os.remove(p)
    t2 = write(p, 2)  # same name, same size, other content
    want2 = t2.column("v").chunk(0).buffers()[1].to_pybytes()[:64]
    assert want2 != want1
    r = b2a.ArrowIPCReader.open_file(p)
    assert first_values(r) == want2
    r.close()
    # modified IN PLACE (same inode and size; the mtime moves): no stale mapping either
    with open(p, "r+b") as f:
        bi_off = b2a.ArrowIPCReader.open_file(p).batch_info(0).body_offset
        f.seek(bi_off)
        f.write(b"\xAB" * 64)
    os.utime(p, ns=(1, 1))
    r = b2a.ArrowIPCReader.open_file(p)
    assert first_values(r) == b"\xAB" * 64
    r.close()
