"""GPU parity of the WRITE direction: encode kernels (through the C ABI) vs the numpy ArrowAppender oracle —
record-batch bodies bit-exact, FieldNodes equal — plus the reference's own write->read checks
(test/sql/write_arrow_stream.test:17-25, test_copy_to.test): the stream decodes with pyarrow AND with this
repo's reader to the source rows."""
import numpy as np
import pyarrow as pa
import pyarrow.ipc as ipc
import pytest

from oracle import oracle_write as OW
from tests_helpers import pylist, same_rows
from vectors import table_vectors, to_device
from writer_cases import large_type, nested_table, rare_columns, scalar_table, writer_type

pytestmark = pytest.mark.gpu


def _lineitem(n, batch_rows=4000):
    from tools import lineitem
    return ipc.open_stream(lineitem.stream_bytes(n, seed=9, batch_rows=batch_rows)).read_all()


def _write(b2a, table, rgs, **kw):
    host = table_vectors(table)
    dev = [to_device(h) for h in host]
    w = b2a.ArrowWriter(dev, table.num_rows, rgs, **kw)
    stream = w.write()
    return host, w, stream


def _check_against_oracle(b2a, table, rgs):
    host, w, stream = _write(b2a, table, rgs)
    exp = OW.serialize_batches(host, table.num_rows, rgs)
    assert w.n_batches == len(exp)
    # headers: parsed by this repo's own host slicer (CPU code) -> nodes / buffer slices
    r = b2a.ArrowIPCReader.open_buffers([stream])
    assert r.num_batches == len(exp) and r.num_rows == table.num_rows
    for i, (n, nodes, slices, body) in enumerate(exp):
        bi = w.batch(i)
        assert bi.rows == n and bi.body_length == len(body), f"batch {i}: body length {bi.body_length} != {len(body)}"
        got = bytes(stream[bi.body_offset: bi.body_offset + bi.body_length])
        if got != body:
            bad = next(k for k in range(len(body)) if got[k] != body[k])
            which = max(j for j, (o, l) in enumerate(slices) if o <= bad)
            raise AssertionError(f"batch {i}: body differs at byte {bad} (buffer #{which} {slices[which]})")
    # decode with pyarrow: FieldNodes (null counts) + content
    rd = ipc.open_stream(pa.py_buffer(stream))
    got = rd.read_all()
    assert same_rows(got, table)
    k = 0
    for batch, (n, nodes, _, _) in zip(ipc.open_stream(pa.py_buffer(stream)), exp):
        assert batch.num_rows == n
        for col, arr in enumerate(batch.columns):
            pass
    return w, stream, got


@pytest.mark.parametrize("case,rgs", [("scalar", 2048), ("scalar", 122_880), ("scalar", 1000), ("nested", 1000),
                                      ("nested", 4096), ("lineitem", 122_880), ("lineitem", 5000)])
def test_writer_bodies_bit_exact_vs_oracle(b2a, cuda_device, case, rgs):
    table = {"scalar": lambda: scalar_table(5000), "nested": lambda: nested_table(3000),
             "lineitem": lambda: _lineitem(300_000, 122_880)}[case]()
    _check_against_oracle(b2a, table, rgs)


def test_schema_formats_match_reference_converter(b2a, cuda_device):
    """lineitem written by COPY must carry the reference's Arrow types: l, d:15,2,64, u, tdD"""
    table = _lineitem(5000)
    _, w, stream = _write(b2a, table, 2048)
    sch = ipc.open_stream(pa.py_buffer(stream)).schema
    assert sch.field("l_orderkey").type == pa.int64()
    assert sch.field("l_quantity").type == pa.decimal64(15, 2)
    assert sch.field("l_shipdate").type == pa.date32()
    assert sch.field("l_comment").type == pa.string()
    r = b2a.ArrowIPCReader.open_buffers([stream])
    fmts = {f.name: f.format for f in r.columns}
    assert fmts["l_orderkey"] == "l" and fmts["l_quantity"] == "d:15,2,64" and fmts["l_shipdate"] == "tdD"
    assert fmts["l_comment"] == "u"


def test_round_trip_through_own_reader(b2a, cuda_device, tmp_path):
    """COPY ... TO 'x.arrows' then read_arrow('x.arrows'): write->read equality (write_arrow_stream.test:17-25)"""
    from parity import compare_scan_with_oracle
    table = scalar_table(7000, seed=5).drop_columns(["nul"])
    host = table_vectors(table)
    path = str(tmp_path / "t.arrows")
    b2a.copy_to_arrows([to_device(h) for h in host], path, row_group_size=3000)
    with open(path, "rb") as f:
        data = f.read()
    batches = list(ipc.open_stream(pa.BufferReader(data)))
    assert [len(b) for b in batches] == [3000, 3000, 1000]
    ds = b2a.read_arrow(path)
    assert ds.num_rows == 7000
    compare_scan_with_oracle(ds, batches)


def test_kv_metadata_and_chunk_size(b2a, cuda_device):
    """test/sql/test_copy_to.test: kv_metadata ends up in the schema's custom metadata; chunk_size = row_group_size"""
    table = _lineitem(6000)
    stream = b2a.copy_to_arrows([to_device(h) for h in table_vectors(table)], None, chunk_size=2500,
                                kv_metadata={"foo": "bar", "n": 42})
    rd = ipc.open_stream(pa.py_buffer(stream))
    assert rd.schema.metadata == {b"foo": b"bar", b"n": b"42"}
    assert [len(b) for b in rd] == [2500, 2500, 1000]


def test_to_arrow_ipc_blobs(b2a, cuda_device):
    """to_arrow_ipc: first row is the schema message (header=true), then one message per 245 760 rows"""
    table = _lineitem(300_000, 100_000)
    rows = b2a.to_arrow_ipc([to_device(h) for h in table_vectors(table)])
    assert rows[0][1] is True and all(not h for _, h in rows[1:])
    assert len(rows) == 1 + 2
    blob = b"".join(b for b, _ in rows) + b"\xff\xff\xff\xff\x00\x00\x00\x00"
    got = ipc.open_stream(pa.py_buffer(blob)).read_all()
    assert same_rows(got, table)


def test_zero_rows_and_single_row(b2a, cuda_device):
    t0 = scalar_table(5).slice(0, 0)
    stream = b2a.copy_to_arrows([to_device(h) for h in table_vectors(t0)], None)
    rd = ipc.open_stream(pa.py_buffer(stream))
    assert rd.read_all().num_rows == 0 and rd.schema.names == t0.schema.names
    t1 = nested_table(1, seed=3)
    _check_against_oracle(b2a, t1, 122_880)


def test_null_map_keys_are_rejected(b2a, cuda_device):
    """"Arrow doesn't accept NULL keys on Maps" (map_data.hpp:97-100)"""
    from vectors import HostVector
    keys = pa.array(["a", None, "c"], pa.string())
    vals = pa.array([1, 2, 3], pa.int64())
    from vectors import from_arrow
    ent = np.array([[0, 3]], np.uint64)
    child = HostVector("STRUCT", 3, None, None, [from_arrow(keys, "key"), from_arrow(vals, "value")], name="entries")
    m = HostVector("MAP", 1, ent.view(np.uint8).reshape(-1), None, [child], name="m")
    w = b2a.ArrowWriter([to_device(m)], 1)
    with pytest.raises(b2a.B2AError, match="NULL keys"):
        w.write()


def test_unsupported_type_is_not_implemented(b2a):
    """a type id the appender has no routine for (arrow_appender.cpp:300-302)"""
    from duckdb_arrow_b200 import DeviceVector
    with pytest.raises(b2a.B2AError, match="Unsupported type"):
        b2a.ArrowWriter([DeviceVector("INVALID", 0, None, None, name="u")], 0, device=-1)


@pytest.mark.parametrize("settings", [dict(arrow_lossless_conversion=True),
                                      dict(arrow_lossless_conversion=True, arrow_large_buffer_size=True),
                                      dict(arrow_lossless_conversion=True, produce_arrow_string_view=True),
                                      dict()])
@pytest.mark.parametrize("rgs", [2048, 1000])
def test_lossless_conversion_bit_exact_vs_oracle(b2a, cuda_device, rgs, settings):
    """SET arrow_lossless_conversion = true (arrow_converter.cpp:130-279, arrow_appender.cpp:144-191): BOOLEAN -> arrow.bool8
    bytes (root columns), UUID -> arrow.uuid big-endian w:16 (B2A_ENC_UUID_BLOB), HUGEINT / UHUGEINT / TIME_TZ ->
    arrow.opaque w:16 / w:8, BIT / VARINT -> arrow.opaque binary; bodies bit-exact vs the restatement that
    tests/test_ref_append.py pins to the reference appender, and pyarrow reads the extension types back.  Without the
    setting the same vectors take the lossy routes and UHUGEINT / VARINT stay extension types."""
    from writer_cases import lossless_table
    n = 4000
    table = lossless_table(n)
    host = table_vectors(table)
    lossless = settings.get("arrow_lossless_conversion", False)
    w = b2a.ArrowWriter([to_device(h) for h in host], n, rgs, **settings)
    stream = w.write()
    exp = OW.serialize_batches(host, n, rgs, large=settings.get("arrow_large_buffer_size", False),
                               sview=settings.get("produce_arrow_string_view", False), lossless=lossless)
    assert w.n_batches == len(exp)
    for i, (rows, nodes, slices, body) in enumerate(exp):
        bi = w.batch(i)
        assert bi.rows == rows and bi.body_length == len(body)
        got = bytes(stream[bi.body_offset: bi.body_offset + bi.body_length])
        if got != body:
            bad = next(k for k in range(len(body)) if got[k] != body[k])
            which = max(j for j, (o, l) in enumerate(slices) if o <= bad)
            raise AssertionError(f"batch {i}: body differs at byte {bad} (buffer #{which} {slices[which]})")
    got = ipc.open_stream(pa.py_buffer(stream)).read_all()
    types = {f.name: f.type for f in got.schema}
    binary = pa.large_binary() if settings.get("arrow_large_buffer_size") else pa.binary()
    assert types["uhuge"] == pa.opaque(pa.binary(16), "uhugeint", "DuckDB")
    assert types["vint"] == pa.opaque(binary, "varint", "DuckDB")
    if lossless:
        assert types["huge"] == pa.opaque(pa.binary(16), "hugeint", "DuckDB")
        assert types["id"] == pa.uuid() and types["flag"] == pa.bool8()
        assert types["ttz"] == pa.opaque(pa.binary(8), "time_tz", "DuckDB")
        assert types["bits"] == pa.opaque(binary, "bit", "DuckDB")
        assert types["st"].field("id").type == pa.uuid()
        assert types["st"].field("flag").type == pa.bool_()  # nested: ArrowBoolData's bits, labelled as what they are
        assert types["l_id"].value_type == pa.uuid()
        for name in ("huge", "uhuge", "id", "id_nn", "ttz", "bits", "vint"):
            assert got.column(name).combine_chunks().storage.to_pylist() == \
                table.column(name).combine_chunks().storage.to_pylist(), name
        want = [None if v is None else int(v != 0) for v in table.column("flag").combine_chunks().storage.to_pylist()]
        assert got.column("flag").combine_chunks().storage.to_pylist() == want
        assert got.column("l_id").to_pylist() == table.column("l_id").to_pylist()
    else:
        assert types["huge"] == pa.decimal128(38, 0) and types["id"] == pa.string() and types["flag"] == pa.bool_()
        assert types["ttz"] == pa.time64("us") and types["bits"] == binary
        ids = table.column("id").to_pylist()
        assert got.column("id").to_pylist() == [None if v is None else str(v) for v in ids]
    # this repo's reader takes the stream back to the same vectors
    r = b2a.ArrowIPCReader.open_buffers([stream])
    assert r.num_rows == n


@pytest.mark.parametrize("rgs", [122_880, 2048, 1000])
def test_uuid_time_tz_bit_union_bit_exact_vs_oracle(b2a, cuda_device, rgs):
    """the appenders of b2a_encode_kernel<true>: ArrowUUIDConverter (36-character strings), ArrowTimeTzConverter,
    BIT as binary, ArrowUnionData (type ids + members masked by tag / union validity, also below a LIST) — bodies
    bit-exact vs the oracle, and pyarrow reads the stream back to the expected rows"""
    n = 6000
    cols, expected, types = rare_columns(n)
    w = b2a.ArrowWriter([to_device(h) for h in cols], n, rgs)
    stream = w.write()
    exp = OW.serialize_batches(cols, n, rgs)
    assert w.n_batches == len(exp)
    for i, (rows, nodes, slices, body) in enumerate(exp):
        bi = w.batch(i)
        assert bi.rows == rows and bi.body_length == len(body)
        got = bytes(stream[bi.body_offset: bi.body_offset + bi.body_length])
        if got != body:
            bad = next(k for k in range(len(body)) if got[k] != body[k])
            which = max(j for j, (o, l) in enumerate(slices) if o <= bad)
            raise AssertionError(f"batch {i}: body differs at byte {bad} (buffer #{which} {slices[which]})")
    rd = ipc.open_stream(pa.py_buffer(stream))
    assert [f.type for f in rd.schema] == types and rd.schema.names == list(expected)
    got = rd.read_all()
    for name, e in expected.items():
        assert pylist(got.column(name)) == e, name
    # FieldNodes: a union node has null_count 0 and one buffer
    batch0 = next(iter(ipc.open_stream(pa.py_buffer(stream))))
    assert batch0.column("u").null_count == 0
    # and this repo's own reader decodes the union back (sparse-union read path, tests/test_union.py)
    r = b2a.ArrowIPCReader.open_buffers([stream])
    assert {f.name: f.format for f in r.columns}["u"].startswith("+us:") and r.num_rows == n


@pytest.mark.parametrize("settings", [dict(), dict(arrow_large_buffer_size=True), dict(produce_arrow_string_view=True)])
@pytest.mark.parametrize("rgs", [2048, 1000])
def test_enum_columns_dictionary_batches(b2a, cuda_device, rgs, settings):
    """ENUM -> dictionary-encoded field + one DictionaryBatch per ENUM field between the schema and the first record batch
    (arrow_converter.cpp:362-386, appender/enum_data.hpp:24-87): uint8 codes for a 4-value ENUM, uint16 for a 300-value
    one, also inside a struct and below a list; record-batch bodies and dictionary bodies bit-exact vs the restatement
    pinned in tests/test_ref_append.py::test_enum_columns_equal_reference; pyarrow and this repo's reader read it back."""
    rng = np.random.default_rng(21)
    n = 5000
    moods = ["sad", "ok", "happy", "a mood well past twelve bytes"]
    big = [f"value number {i:03d}" for i in range(300)]
    m = pa.DictionaryArray.from_arrays(pa.array(rng.integers(0, 4, n), pa.uint8(), mask=rng.random(n) < 0.15), pa.array(moods))
    b = pa.DictionaryArray.from_arrays(pa.array(rng.integers(0, 300, n), pa.uint16()), pa.array(big))
    k = pa.array(range(n), pa.int64())
    st = pa.StructArray.from_arrays([m, k], names=["m", "k"])
    lens = rng.integers(0, 4, n)
    lo = np.zeros(n + 1, np.int32)
    np.cumsum(lens, out=lo[1:])
    lb = pa.ListArray.from_arrays(pa.array(lo, pa.int32()), b.take(pa.array(rng.integers(0, n, int(lo[-1])))))
    s = pa.array(["s%d" % i if i % 3 else "a string well past twelve bytes %d" % i for i in range(n)])
    table = pa.table({"m": m, "b": b, "st": st, "lb": lb, "s": s})
    host = table_vectors(table)
    w = b2a.ArrowWriter([to_device(h) for h in host], n, rgs, **settings)
    stream = w.write()
    kw = dict(large=settings.get("arrow_large_buffer_size", False), sview=settings.get("produce_arrow_string_view", False))
    exp = OW.serialize_batches(host, n, rgs, **kw)
    assert w.n_batches == len(exp)
    for i, (rows, nodes, slices, body) in enumerate(exp):
        bi = w.batch(i)
        assert bi.rows == rows and bi.body_length == len(body)
        assert bytes(stream[bi.body_offset: bi.body_offset + bi.body_length]) == body, f"batch {i}"
    # the dictionary region: [schema, 4 x (DictionaryBatch header + body)) — every body is at the end of its message
    dicts = OW.serialize_dictionaries(host)
    assert len(dicts) == 4
    region = bytes(stream[len(w.schema_message()): w.preamble_bytes])
    pos = 0
    for d, (rows, nodes, slices, body) in enumerate(dicts):
        assert region[pos: pos + 4] == b"\xff\xff\xff\xff"
        meta = int.from_bytes(region[pos + 4: pos + 8], "little")
        assert region[pos + 8 + meta: pos + 8 + meta + len(body)] == body, f"dictionary {d}"
        pos += 8 + meta + len(body)
    assert pos == len(region)
    got = ipc.open_stream(pa.py_buffer(stream)).read_all()
    assert got.schema.field("m").type == pa.dictionary(pa.uint8(), pa.string())
    assert got.schema.field("b").type == pa.dictionary(pa.uint16(), pa.string())
    for name in table.column_names:
        assert got.column(name).to_pylist() == table.column(name).to_pylist(), name
    # this repo's reader: dictionary-encoded columns come back as VARCHAR, like the reference's scan
    r = b2a.ArrowIPCReader.open_buffers([stream])
    assert r.num_rows == n


def test_enum_needs_its_dictionary(b2a):
    from duckdb_arrow_b200 import DeviceVector
    with pytest.raises(b2a.B2AError, match="needs one VARCHAR child"):
        b2a.ArrowWriter([DeviceVector("ENUM", 0, None, None, name="e")], 0, device=-1)


def test_union_with_nested_member_is_rejected(b2a):
    from duckdb_arrow_b200 import DeviceVector
    m = DeviceVector("LIST", 0, None, None, [DeviceVector("INTEGER", 0, None, None, name="l")], name="m")
    with pytest.raises(b2a.B2AError, match="flat members only"):
        b2a.ArrowWriter([DeviceVector("UNION", 0, None, None, [m], name="u")], 0, device=-1)


@pytest.mark.parametrize("layout", ["contiguous", "reversed", "displaced", "gapped"])
def test_string_heap_layouts(b2a, cuda_device, layout):
    """enc_string's contiguous-heap fast path (one coalesced window copy + inlined strings patched from registers)
    against heaps that qualify, that do not, and that almost do — bodies bit-exact vs the oracle in every case."""
    rng = np.random.default_rng(11)
    n = 9000
    lens = rng.integers(13, 60, n)
    short = rng.random(n) < 0.35
    lens[short] = rng.integers(0, 13, int(short.sum()))
    lens[3000:3400] = rng.integers(0, 13, 400)  # whole warps of inlined strings: those tiles must take the slow path
    vals = [None if i % 37 == 5 else "".join(chr(97 + (i + k) % 26) for k in range(int(lens[i]))) for i in range(n)]
    table = pa.table({"s": pa.array(vals, pa.string()), "k": pa.array(range(n), pa.int32())})
    host = table_vectors(table)
    hv = host[0]
    start, length = hv.str_start.copy(), hv.str_len.copy()
    heap = hv.heap
    if layout == "reversed":
        order = np.arange(n)[::-1]
    elif layout in ("displaced", "gapped"):
        order = np.arange(n)
    if layout != "contiguous":
        gap = 3 if layout == "gapped" else 0
        new_start = np.zeros(n, start.dtype)
        pieces, pos = [], 0
        for i in order:
            new_start[i] = pos
            pieces.append(heap[start[i]: start[i] + length[i]])
            pos += int(length[i])
            if gap and i % 500 == 499:
                pieces.append(np.full(gap, 35, np.uint8))
                pos += gap
        new_heap = np.concatenate(pieces + [np.zeros(64, np.uint8)])
        if layout == "displaced":  # one long string of an otherwise contiguous tile lives somewhere else
            for i in (1500, 5000):
                while length[i] <= 12:
                    i += 1
                new_heap = np.concatenate([new_heap, heap[start[i]: start[i] + length[i]]])
                new_start[i] = len(new_heap) - int(length[i])
        hv.str_start, hv.heap = new_start, new_heap
    dev = [to_device(h) for h in host]
    for rgs in (122_880, 2048):
        w = b2a.ArrowWriter(dev, n, rgs)
        stream = w.write()
        exp = OW.serialize_batches(host, n, rgs)
        assert w.n_batches == len(exp)
        for i, (rows, nodes, slices, body) in enumerate(exp):
            bi = w.batch(i)
            assert bytes(stream[bi.body_offset: bi.body_offset + bi.body_length]) == body, (layout, rgs, i)
        assert same_rows(ipc.open_stream(pa.py_buffer(stream)).read_all(), table)


def _bodies_equal(w, stream, exp):
    assert w.n_batches == len(exp)
    for i, (rows, nodes, slices, body) in enumerate(exp):
        bi = w.batch(i)
        assert bi.rows == rows and bi.body_length == len(body), f"batch {i}: body length {bi.body_length} != {len(body)}"
        got = bytes(stream[bi.body_offset: bi.body_offset + bi.body_length])
        if got != body:
            bad = next(k for k in range(len(body)) if got[k] != body[k])
            which = max(j for j, (o, l) in enumerate(slices) if o <= bad)
            raise AssertionError(f"batch {i}: body differs at byte {bad} (buffer #{which} {slices[which]})")


@pytest.mark.parametrize("case,rgs", [("scalar", 2048), ("nested", 1000), ("nested", 4096), ("lineitem", 122_880),
                                      ("rare", 1000)])
def test_large_offsets_bit_exact_vs_oracle(b2a, cuda_device, case, rgs):
    """SET arrow_large_buffer_size = true: ArrowVarcharData<> / ArrowListData<> with int64 offsets (B2A_ENC_STRING64 /
    LIST64 / UUID64), MAP still int32 (arrow_appender.cpp:184-188,237-251,288-299) — bodies bit-exact vs the oracle, the
    schema carries large_utf8 / large_binary / large_list, pyarrow and this repo's reader decode the same rows"""
    if case == "rare":
        host, expected, types = rare_columns(6000)
        rows, names = 6000, list(expected)
    else:
        table = {"scalar": lambda: scalar_table(5000), "nested": lambda: nested_table(3000),
                 "lineitem": lambda: _lineitem(300_000, 122_880)}[case]()
        host, rows, names, types = table_vectors(table), table.num_rows, table.schema.names, [f.type for f in table.schema]
        expected = None
    w = b2a.ArrowWriter([to_device(h) for h in host], rows, rgs, arrow_large_buffer_size=True)
    stream = w.write()
    _bodies_equal(w, stream, OW.serialize_batches(host, rows, rgs, large=True))
    rd = ipc.open_stream(pa.py_buffer(stream))
    assert [f.type for f in rd.schema] == [large_type(t) for t in types] and rd.schema.names == names
    got = rd.read_all()
    got.validate(full=True)
    if expected is None:
        assert same_rows(got, table)
    else:
        for name, e in expected.items():
            assert pylist(got.column(name)) == e, name
    if case != "rare":  # read_arrow of the large stream: SetVectorString<uint64_t> / 64-bit list offsets vs the oracle
        from parity import compare_scan_with_oracle
        data = bytes(stream)
        compare_scan_with_oracle(b2a.scan_arrow_ipc([data]), list(ipc.open_stream(pa.BufferReader(data))))


@pytest.mark.parametrize("large", [False, True])
@pytest.mark.parametrize("case,rgs", [("nested", 1000), ("nested", 4096), ("nested", 999), ("rare", 1000)])
def test_list_views_bit_exact_vs_oracle(b2a, cuda_device, case, rgs, large):
    """SET arrow_output_list_view = true: ArrowListViewData (B2A_ENC_LISTVIEW32 / 64: offsets + sizes of `rows` entries,
    NULL row = {running offset, 0}, appender/list_view_data.hpp:59-103) — bodies bit-exact vs the oracle (odd row counts
    exercise the zero padding of the 4-byte sizes buffer), list_view types in the schema, same rows through pyarrow and
    through this repo's list-view read path"""
    if case == "rare":
        host, expected, types = rare_columns(6000)
        rows, names = 6000, list(expected)
    else:
        table = nested_table(3001)
        host, rows, names, types = table_vectors(table), table.num_rows, table.schema.names, [f.type for f in table.schema]
        expected = {n: pylist(table.column(n)) for n in names}
    w = b2a.ArrowWriter([to_device(h) for h in host], rows, rgs, arrow_large_buffer_size=large,
                        arrow_output_list_view=True)
    stream = w.write()
    _bodies_equal(w, stream, OW.serialize_batches(host, rows, rgs, large=large, view=True))
    rd = ipc.open_stream(pa.py_buffer(stream))
    assert [f.type for f in rd.schema] == [writer_type(t, large, True) for t in types] and rd.schema.names == names
    got = rd.read_all()
    got.validate(full=True)
    for name, e in expected.items():
        assert pylist(got.column(name)) == e, name
    if case != "rare":  # (the scan path refuses a list view directly below a list view, DESIGN 5e: "ll" stays out)
        from parity import compare_scan_with_oracle
        data = bytes(stream)
        keep = [n for n in names if n != "ll"]
        compare_scan_with_oracle(b2a.scan_arrow_ipc([data], columns=keep), list(ipc.open_stream(pa.BufferReader(data))),
                                 columns=keep)


@pytest.mark.parametrize("case,rgs", [("scalar", 2048), ("scalar", 999), ("nested", 1000), ("lineitem", 122_880)])
def test_string_views_bit_exact_vs_oracle(b2a, cuda_device, case, rgs):
    """SET produce_arrow_string_view = true: B2A_ENC_STRINGVIEW (ArrowVarcharToStringViewData, varchar_data.hpp:129-202) —
    views and the data buffer bit-exact vs the oracle, RecordBatch.variadicBufferCounts understood by pyarrow, which
    validates every view and reads the source rows back; this repo's Utf8View read path decodes the stream too"""
    table = {"scalar": lambda: scalar_table(5000), "nested": lambda: nested_table(3000),
             "lineitem": lambda: _lineitem(300_000, 122_880)}[case]()
    host, rows, types = table_vectors(table), table.num_rows, [f.type for f in table.schema]
    w = b2a.ArrowWriter([to_device(h) for h in host], rows, rgs, produce_arrow_string_view=True)
    stream = w.write()
    _bodies_equal(w, stream, OW.serialize_batches(host, rows, rgs, sview=True))
    rd = ipc.open_stream(pa.py_buffer(stream))
    assert [f.type for f in rd.schema] == [writer_type(t, sview=True) for t in types]
    got = rd.read_all()
    got.validate(full=True)
    for name in table.schema.names:
        assert pylist(got.column(name)) == pylist(table.column(name)), name
    from parity import compare_scan_with_oracle
    data = bytes(stream)
    compare_scan_with_oracle(b2a.scan_arrow_ipc([data]), list(ipc.open_stream(pa.BufferReader(data))))


@pytest.mark.parametrize("settings", [dict(arrow_large_buffer_size=True), dict(arrow_output_list_view=True),
                                      dict(produce_arrow_string_view=True),
                                      dict(arrow_large_buffer_size=True, arrow_output_list_view=True,
                                           produce_arrow_string_view=True)])
def test_writer_settings_edge_sizes(b2a, cuda_device, settings):
    """zero rows, one row, a batch boundary at an odd row count and every setting at once: bodies vs the oracle and the
    rows pyarrow reads back"""
    kw = dict(large=settings.get("arrow_large_buffer_size", False), view=settings.get("arrow_output_list_view", False),
              sview=settings.get("produce_arrow_string_view", False))
    for table, rgs in [(nested_table(5).slice(0, 0), 100), (nested_table(1, seed=3), 100), (nested_table(777, seed=6), 333),
                       (scalar_table(1001, seed=8), 1001)]:
        host = table_vectors(table)
        w = b2a.ArrowWriter([to_device(h) for h in host], table.num_rows, rgs, **settings)
        stream = w.write()
        _bodies_equal(w, stream, OW.serialize_batches(host, table.num_rows, rgs, **kw))
        rd = ipc.open_stream(pa.py_buffer(stream))
        assert [f.type for f in rd.schema] == [writer_type(f.type, **kw) for f in table.schema]
        got = rd.read_all()
        got.validate(full=True)
        assert got.num_rows == table.num_rows
        for name in table.schema.names:
            assert pylist(got.column(name)) == pylist(table.column(name)), name


def test_string_view_members_of_unions(b2a, cuda_device):
    """VARCHAR members of a sparse UNION (top level and below a LIST) as views: the member's mask (union valid AND tag ==
    member AND value valid) is what B2A_ENC_STRINGVIEW sees as validity; TIME_TZ / BIT next to them unchanged"""
    n = 6000
    cols, expected, types = rare_columns(n)
    keep = [i for i, c in enumerate(cols) if c.type != "UUID"]  # refused under this setting
    cols, names, types = [cols[i] for i in keep], [list(expected)[i] for i in keep], [types[i] for i in keep]
    w = b2a.ArrowWriter([to_device(h) for h in cols], n, 1000, produce_arrow_string_view=True)
    stream = w.write()
    _bodies_equal(w, stream, OW.serialize_batches(cols, n, 1000, sview=True))
    rd = ipc.open_stream(pa.py_buffer(stream))
    assert [f.type for f in rd.schema] == [writer_type(t, sview=True) for t in types]
    got = rd.read_all()
    got.validate(full=True)
    for name in names:
        assert pylist(got.column(name)) == expected[name], name
