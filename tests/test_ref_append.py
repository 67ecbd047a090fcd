"""Pins the WRITE-direction restatement (oracle/oracle_write.py) to the REFERENCE ITSELF: the same rows go through the
reference's own ArrowConverter::ToArrowArray / ArrowAppender (duckdb/src/common/arrow/arrow_converter.cpp:19-25,
arrow_appender.cpp:37-131, appender/*.hpp — compiled from /root/reference/duckdb into oracle/_ref), fed ONE DataChunk per
record batch the way ColumnDataCollectionSerializer::Serialize(ColumnDataCollection&) feeds it
(src/writer/column_data_collection_serializer.cpp:97-115), under each SET variant; every Arrow buffer it produces —
validity (0xFF fill, bits past the last row), offsets, characters, values incl. the NULL slots of fixed-width buffers
where the appender copies them — must equal the restatement's byte for byte, and so must (length, null_count) of every
node.  The GPU writer is compared against the same restatement in tests/test_gpu_writer.py."""
import ctypes

import numpy as np
import pyarrow as pa
import pytest

from oracle import oracle_write as OW
from oracle import ref_py as R
from vectors import table_vectors
from writer_cases import nested_table, scalar_table

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built (needs /root/reference)")


def _bytes(buf, size):
    if buf is None or size == 0:
        return b""
    return ctypes.string_at(buf.address, size)


def ref_nodes_buffers(arr: pa.Array, nodes: list, bufs: list, notes: list, path="", depth=0) -> None:
    """Pre-order (length, null_count) + buffers of what the reference appender produced; `notes[i]` says which kind of
    buffer bufs[i] is and the validity of its rows (for the buffers whose NULL slots the appender leaves unwritten)."""
    packed_bool8 = False
    if isinstance(arr.type, pa.BaseExtensionType):  # the buffers are those of the storage array
        # a NESTED arrow.bool8 of the reference carries ArrowBoolData's bit-packed buffer under its int8 label
        # (arrow_appender.cpp:40-48 converts root columns only): read it as the bits it holds
        packed_bool8 = arr.type.extension_name == "arrow.bool8" and depth > 0
        arr = arr.storage
    t, n = arr.type, len(arr)
    assert arr.offset == 0
    b = arr.buffers()
    if pa.types.is_null(t):
        nodes.append((n, n))
        return
    if pa.types.is_union(t):
        nodes.append((n, 0))
        bufs.append(_bytes(b[1], n))
        notes.append((path, "type_ids", None, 1))
        for i in range(t.num_fields):
            ref_nodes_buffers(arr.field(i), nodes, bufs, notes, f"{path}.{t.field(i).name}", depth + 1)
        return
    nodes.append((n, arr.null_count))
    valid = np.asarray(arr.is_valid().to_numpy(zero_copy_only=False), bool) if n else np.zeros(0, bool)
    bufs.append(_bytes(b[0], (n + 7) // 8))
    notes.append((path, "validity", None, 0))
    if pa.types.is_struct(t):
        for i in range(t.num_fields):
            ref_nodes_buffers(arr.field(i), nodes, bufs, notes, f"{path}.{t.field(i).name}", depth + 1)
    elif pa.types.is_list(t) or pa.types.is_large_list(t) or pa.types.is_map(t):
        w = 8 if pa.types.is_large_list(t) else 4
        bufs.append(_bytes(b[1], (n + 1) * w))
        notes.append((path, "offsets", None, w))
        ref_nodes_buffers(arr.values, nodes, bufs, notes, path + "[]", depth + 1)
    elif pa.types.is_list_view(t) or pa.types.is_large_list_view(t):
        w = 8 if pa.types.is_large_list_view(t) else 4
        bufs.append(_bytes(b[1], n * w))
        notes.append((path, "view_offsets", None, w))
        bufs.append(_bytes(b[2], n * w))
        notes.append((path, "view_sizes", None, w))
        ref_nodes_buffers(arr.values, nodes, bufs, notes, path + "[]", depth + 1)
    elif pa.types.is_fixed_size_list(t):
        ref_nodes_buffers(arr.values, nodes, bufs, notes, path + "[]", depth + 1)
    elif pa.types.is_string_view(t):
        bufs.append(_bytes(b[1], n * 16))
        notes.append((path, "views", valid, 16))
        views = np.frombuffer(bufs[-1], np.uint8).reshape(n, 16) if n else np.zeros((0, 16), np.uint8)
        lens = views[:, 0:4].copy().view(np.int32).ravel()
        total = int(lens[(lens > 12) & valid].sum())
        bufs.append(_bytes(b[2], total) if len(b) > 2 else b"")
        notes.append((path, "view_data", None, 1))
    elif (pa.types.is_string(t) or pa.types.is_binary(t) or pa.types.is_large_string(t) or pa.types.is_large_binary(t)):
        w = 8 if (pa.types.is_large_string(t) or pa.types.is_large_binary(t)) else 4
        bufs.append(_bytes(b[1], (n + 1) * w))
        notes.append((path, "offsets", None, w))
        last = int(np.frombuffer(bufs[-1], np.int64 if w == 8 else np.int32)[-1]) if n else 0
        bufs.append(_bytes(b[2], last))
        notes.append((path, "chars", None, 1))
    elif pa.types.is_boolean(t) or packed_bool8:
        bufs.append(_bytes(b[1], (n + 7) // 8))
        notes.append((path, "bool", None, 0))
    else:
        w = t.bit_width // 8
        bufs.append(_bytes(b[1], n * w))
        notes.append((path, "values", valid, w))


def _table(case):
    if case == "lossless":
        from writer_cases import lossless_table
        return lossless_table(3000)
    if case == "scalar":
        return scalar_table(5000)
    if case == "nested":
        return nested_table(3000)
    import pyarrow.ipc as ipc
    from tools import lineitem
    return ipc.open_stream(lineitem.stream_bytes(12_345, seed=9, batch_rows=4000)).read_all()


# Columns whose NULL slots are NOT a function of the input in the reference: the scan leaves them unwritten
# (ConvertDecimal narrowing, arrow_conversion.cpp:776-781) or the appender skips them (SkipNulls converters:
# ArrowIntervalConverter, scalar_data.hpp:34-52) — compared on valid rows only.
def _same_buffer(rb: bytes, ob: bytes, note) -> tuple[bool, str]:
    if rb == ob:
        return True, "exact"
    path, kind, valid, w = note
    if valid is not None and len(rb) == len(ob) and w:
        a = np.frombuffer(rb, np.uint8).reshape(-1, w)
        b = np.frombuffer(ob, np.uint8).reshape(-1, w)
        if np.array_equal(a[valid], b[valid]):
            return True, "valid rows"
    return False, ""


@pytest.mark.parametrize("flags", [0, R.FLAG_LARGE, R.FLAG_LIST_VIEW, R.FLAG_STRING_VIEW,
                                   R.FLAG_LARGE | R.FLAG_LIST_VIEW])
@pytest.mark.parametrize("case,rgs", [("scalar", 2048), ("scalar", 5000), ("nested", 1000), ("nested", 3000),
                                      ("lineitem", 5000)])
def test_appender_restatement_equals_reference(case, rgs, flags):
    _check_against_reference(case, rgs, flags)


@pytest.mark.parametrize("flags", [R.FLAG_LOSSLESS, R.FLAG_LOSSLESS | R.FLAG_LARGE, R.FLAG_LOSSLESS | R.FLAG_STRING_VIEW,
                                   0])
@pytest.mark.parametrize("rgs", [1000, 2048])  # ArrowBool8::DuckToArrow fills a 2 048-row Vector: one chunk at most
def test_lossless_conversion_equals_reference(rgs, flags):
    """SET arrow_lossless_conversion: arrow.bool8 / arrow.uuid / arrow.opaque{DuckDB, ...} storage buffers exactly as the
    reference appender lays them out; with flags = 0 the same vectors take the lossy routes (36-character UUIDs,
    time64[us], bit-packed booleans) and UHUGEINT / VARINT stay extension types."""
    _check_against_reference("lossless", rgs, flags)


def _check_against_reference(case, rgs, flags):
    table = _table(case)
    cols = table_vectors(table)
    kw = dict(large=bool(flags & R.FLAG_LARGE), view=bool(flags & R.FLAG_LIST_VIEW), sview=bool(flags & R.FLAG_STRING_VIEW),
              lossless=bool(flags & R.FLAG_LOSSLESS))
    row = 0
    loose = set()
    for n, nodes, slices, body in OW.serialize_batches(cols, table.num_rows, rgs, **kw):
        part = table.slice(row, n).combine_chunks()
        ref = R.append(part, flags)
        assert len(ref) == n
        rnodes, rbufs, notes = [], [], []
        for i in range(ref.type.num_fields):
            ref_nodes_buffers(ref.field(i), rnodes, rbufs, notes, ref.type.field(i).name)
        assert rnodes == nodes, f"{case} rows {row}+{n}: (length, null_count) differ"
        assert len(rbufs) == len(slices)
        for k, ((o, l), rb) in enumerate(zip(slices, rbufs)):
            ob = body[o:o + l]
            assert len(rb) == l, f"{case} rows {row}+{n}: buffer {k} {notes[k][:2]} has {len(rb)} bytes, restatement {l}"
            ok, how = _same_buffer(rb, ob, notes[k])
            assert ok, f"{case} rows {row}+{n}: buffer {k} {notes[k][:2]} differs"
            if how != "exact":
                loose.add(notes[k][0])
        row += n
    assert row == table.num_rows
    # only the columns whose NULL slots the reference leaves unwritten may fall back to the valid-row comparison
    allowed = {"dec4", "iv"}
    if case == "lossless" and flags & R.FLAG_LOSSLESS:  # ArrowUUIDBlobConverter::SkipNulls (scalar_data.hpp:80-82)
        allowed |= {"id", "st.id", "l_id[]"}
    elif case == "lossless":  # ArrowTimeTzConverter::SkipNulls (:60-62)
        allowed |= {"ttz"}
    assert loose <= allowed, f"NULL slots differ in {sorted(loose - allowed)}"


def test_appender_validity_tail_bits_and_null_slots():
    """append_data.hpp:117-149: 0xFF fill, bits past the last row stay 1; scalar_data.hpp:90-119: NULL slots of a plain
    scalar column carry the source vector's bytes (here: what arrow_scan aliased from the Arrow buffer)"""
    vals = pa.Array.from_buffers(pa.int32(), 5, [pa.py_buffer(bytes([0b01101])),
                                                 pa.py_buffer(np.array([7, 8, 9, 10, 11], np.int32).tobytes())])
    ref = R.append(pa.RecordBatch.from_arrays([vals], names=["v"]))
    col = ref.field(0)
    assert ctypes.string_at(col.buffers()[0].address, 1) == bytes([0b11101101])
    assert col.null_count == 2
    assert np.frombuffer(ctypes.string_at(col.buffers()[1].address, 20), np.int32).tolist() == [7, 8, 9, 10, 11]
    t = pa.table({"v": vals})
    (n, nodes, slices, body), = OW.serialize_batches(table_vectors(t), 5, 100)
    assert body[slices[0][0]] == 0b11101101 and nodes == [(5, 2)]
    assert np.frombuffer(body[slices[1][0]: slices[1][0] + 20], np.int32).tolist() == [7, 8, 9, 10, 11]


def test_dbgen_export_has_a_bitmap_per_column():
    """What a reference-written lineitem batch looks like (SURVEY App. A's open question, answered by the appender
    itself): every column carries a 0xFF validity buffer with null_count 0"""
    R.query("call dbgen(sf=0.01)")
    try:
        b = next(iter(R.export("select * from lineitem", 20_000)))
    finally:
        for t in ("lineitem", "orders", "partsupp", "part", "customer", "supplier", "nation", "region"):
            R.query(f"drop table {t}")
    assert b.num_rows == 20_000 and b.num_columns == 16
    assert [str(f.type) for f in b.schema][:5] == ["int64", "int64", "int64", "int64", "decimal64(15, 2)"]
    for col in b.columns:
        assert col.null_count == 0
        v = col.buffers()[0]
        assert v is not None and ctypes.string_at(v.address, 2500) == b"\xff" * 2500


@pytest.mark.parametrize("flags", [0, R.FLAG_LARGE])  # under produce_arrow_string_view the reference finalises the
# dictionary as a 4-buffer view array over int32 offsets + format "u" (FinalizeChild(VARCHAR)): not importable
def test_enum_columns_equal_reference(flags):
    """ENUM -> dictionary-encoded column (arrow_converter.cpp:362-386, appender/enum_data.hpp:24-87): code width by
    dictionary size, codes copied as they are, the dictionary a utf8 array with int32 offsets and a 0xFF validity buffer
    whatever the settings say — against what the reference appender exports for a small (uint8) and a 300-value (uint16)
    ENUM, top level and inside a struct."""
    big = ", ".join(f"'value number {i:03d}'" for i in range(300))
    R.query("CREATE TYPE mood_t AS ENUM ('sad', 'ok', 'happy', 'a mood well past twelve bytes')")
    R.query(f"CREATE TYPE big_t AS ENUM ({big})")
    try:
        sql = ("SELECT CASE WHEN i % 7 = 0 THEN NULL ELSE (['sad','ok','happy','a mood well past twelve bytes'])[1 + i % 4]::mood_t "
               "END AS m, ('value number ' || lpad((i * 7 % 300)::VARCHAR, 3, '0'))::big_t AS b, "
               "{'m': (['sad','ok'])[1 + i % 2]::mood_t, 'k': i} AS st FROM range(5000) t(i)")
        batches = list(R.export(sql, 2048, flags))
    finally:
        R.query("DROP TYPE mood_t")
        R.query("DROP TYPE big_t")
    assert [b.num_rows for b in batches] == [2048, 2048, 904]
    assert str(batches[0].schema.field("m").type) == "dictionary<values=string, indices=uint8, ordered=0>"
    assert str(batches[0].schema.field("b").type) == "dictionary<values=string, indices=uint16, ordered=0>"
    for b in batches:
        cols = table_vectors(pa.Table.from_batches([b]))
        (n, nodes, slices, body), = OW.serialize_batches(cols, b.num_rows, 1 << 20, large=bool(flags & R.FLAG_LARGE),
                                                         sview=bool(flags & R.FLAG_STRING_VIEW))
        rnodes, rbufs = [], []

        def walk(arr):
            bufs = arr.buffers()
            if pa.types.is_struct(arr.type):
                rnodes.append((len(arr), arr.null_count))
                rbufs.append(_bytes(bufs[0], (len(arr) + 7) // 8))
                for i in range(arr.type.num_fields):
                    walk(arr.field(i))
                return
            if pa.types.is_dictionary(arr.type):
                idx = arr.indices
                rnodes.append((len(arr), arr.null_count))
                rbufs.append(_bytes(idx.buffers()[0], (len(arr) + 7) // 8))
                rbufs.append(_bytes(idx.buffers()[1], len(arr) * idx.type.bit_width // 8))
                return
            rnodes.append((len(arr), arr.null_count))
            rbufs.append(_bytes(bufs[0], (len(arr) + 7) // 8))
            rbufs.append(_bytes(bufs[1], len(arr) * arr.type.bit_width // 8))

        for c in b.columns:
            walk(c)
        assert rnodes == nodes
        assert [body[o:o + l] for o, l in slices] == rbufs
        # the dictionaries, in field pre-order
        dicts = OW.serialize_dictionaries(cols)
        ref_dicts = [b.column("m").dictionary, b.column("b").dictionary, b.column("st").field(0).dictionary]
        assert len(dicts) == 3
        for (rows, dnodes, dslices, dbody), rd in zip(dicts, ref_dicts):
            assert rd.type == pa.string() and rows == len(rd) and dnodes == [(len(rd), 0)]
            rb = rd.buffers()
            last = int(np.frombuffer(_bytes(rb[1], (rows + 1) * 4), np.int32)[-1])
            want = [_bytes(rb[0], (rows + 7) // 8), _bytes(rb[1], (rows + 1) * 4), _bytes(rb[2], last)]
            assert [dbody[o:o + l] for o, l in dslices] == want


@pytest.mark.parametrize("seed0", [8000, 8030])
def test_appender_restatement_equals_reference_on_random_nested_tables(seed0):
    """a slice of the writer-side differential fuzz (2 000 random tables x 4 settings = 10 448 record batches were run by hand,
    no difference): random nested tables (tests/fuzz_arrays.py), cut into record batches; the restatement gets the vectors
    the scan leaves for exactly the arrays the reference appends (tests/scan_vectors.py — the scan's treatment of a sliced
    nested array is part of the input), and every (length, null_count) and buffer must match under each setting"""
    import fuzz_arrays
    from oracle import oracle_py as O
    from scan_vectors import slab_vector
    compared = 0
    for seed in range(seed0, seed0 + 30):
        rng = np.random.default_rng(seed)
        n = int(rng.choice([1, 100, 2047, 2048, 2049, 3000]))
        table = pa.table({f"c{i}": fuzz_arrays.gen(rng, n) for i in range(int(rng.integers(1, 4)))})
        rgs = int(rng.choice([1000, 2048, 5000]))
        for flags in (0, R.FLAG_LARGE, R.FLAG_LIST_VIEW, R.FLAG_STRING_VIEW):
            kw = dict(large=bool(flags & R.FLAG_LARGE), view=bool(flags & R.FLAG_LIST_VIEW),
                      sview=bool(flags & R.FLAG_STRING_VIEW), lossless=False)
            for row in range(0, n, rgs):
                part = pa.table({nm: table.column(nm).combine_chunks().slice(row, min(rgs, n - row)) for nm in table.column_names})
                try:
                    cols = [slab_vector(O.scan_column(part.column(i).chunk(0)), part.schema.field(i).type,
                                        part.schema.field(i).name) for i in range(part.num_columns)]
                except NotImplementedError:  # string views as INPUT
                    break
                (nb, nodes, slices, body), = OW.serialize_batches(cols, part.num_rows, part.num_rows, **kw)
                ref = R.append(part, flags)
                rnodes, rbufs, notes = [], [], []
                for i in range(ref.type.num_fields):
                    ref_nodes_buffers(ref.field(i), rnodes, rbufs, notes, ref.type.field(i).name)
                where = f"seed {seed} flags {flags} rows {row}+{part.num_rows} {[str(c.type) for c in table.columns]}"
                assert rnodes == nodes, where
                for k, ((o, ln), rb) in enumerate(zip(slices, rbufs)):
                    assert len(rb) == ln, (where, notes[k][:2])
                    assert _same_buffer(rb, body[o:o + ln], notes[k])[0], (where, notes[k][:2])
                compared += 1
    assert compared > 100
