"""CPU-only: pins the WRITE-direction oracle (oracle/oracle_write.py, a numpy restatement of DuckDB's ArrowAppender)
against pyarrow: the buffers it emits, reassembled as Arrow arrays, must equal the source data — for every writable
type, with NULLs, across several record batches (write_arrow_stream.test:17-25 checks the same write->read equality
in the reference)."""
import numpy as np
import pyarrow as pa
import pytest

from oracle import oracle_write as OW
from tests_helpers import pylist, same_rows
from vectors import table_vectors
from writer_cases import arrays_from_buffers, large_type, nested_table, rare_columns, scalar_table, writer_type


def _lineitem(n):
    import pyarrow.ipc as ipc
    from tools import lineitem
    return ipc.open_stream(lineitem.stream_bytes(n, seed=9, batch_rows=4000)).read_all()


@pytest.mark.parametrize("case,rgs", [("scalar", 2048), ("scalar", 122_880), ("nested", 1000), ("nested", 4096),
                                      ("lineitem", 5000)])
def test_write_oracle_round_trips_through_pyarrow(case, rgs):
    table = {"scalar": lambda: scalar_table(5000), "nested": lambda: nested_table(3000),
             "lineitem": lambda: _lineitem(12_345)}[case]()
    cols = table_vectors(table)
    batches = OW.serialize_batches(cols, table.num_rows, rgs)
    assert sum(b[0] for b in batches) == table.num_rows
    assert len(batches) == (table.num_rows + rgs - 1) // rgs
    row = 0
    for n, nodes, slices, body in batches:
        assert len(body) % 8 == 0
        bufs = [body[o:o + l] for o, l in slices]
        for (o, l) in slices:
            assert o % 8 == 0
        arrays = arrays_from_buffers([f.type for f in table.schema], nodes, bufs)
        got = pa.Table.from_arrays(arrays, names=table.schema.names)
        exp = table.slice(row, n)
        assert same_rows(got, exp), f"{case}: batch at row {row} differs"
        row += n


@pytest.mark.parametrize("rgs", [1000, 122_880])
def test_write_oracle_uuid_time_tz_bit_union(rgs):
    """UUID -> 36-character strings (checked against Python's uuid module), TIME_TZ -> time64[us], BIT -> binary, sparse
    UNION (type ids + masked members): the oracle's buffers, reassembled by pyarrow, give the expected rows"""
    cols, expected, types = rare_columns(3000)
    row = 0
    for n, nodes, slices, body in OW.serialize_batches(cols, 3000, rgs):
        bufs = [body[o:o + l] for o, l in slices]
        arrays = arrays_from_buffers(types, nodes, bufs)
        for arr, (name, exp) in zip(arrays, expected.items()):
            arr.validate(full=True)
            assert pylist(arr) == exp[row: row + n], name
        row += n
    assert row == 3000


def test_union_null_rows_get_tag_zero_and_all_members_null():
    """union_data.cpp:34-47: a NULL union row is written as type id 0 with every member NULL at that row"""
    from vectors import from_arrow, union_vector
    u = union_vector(np.array([1, 1, 0]), [from_arrow(pa.array([7, 8, 9], pa.int32()), "a"),
                                          from_arrow(pa.array(["x", None, "z"], pa.string()), "b")],
                     np.array([True, False, True]), "u")
    (n, nodes, slices, body), = OW.serialize_batches([u], 3, 100)
    assert nodes == [(3, 0), (3, 2), (3, 2)]
    assert body[slices[0][0]: slices[0][0] + 3] == bytes([1, 0, 0])
    assert body[slices[1][0]] == 0b11111100 and body[slices[3][0]] == 0b11111001


def test_validity_buffer_semantics():
    """0xFF fill, NULL bits cleared, bits past the last row stay 1 (append_data.hpp:116-149)"""
    buf, nulls = OW._validity_buffer(np.array([True, False, True]))
    assert buf == bytes([0b11111101]) and nulls == 1
    buf, nulls = OW._validity_buffer(np.ones(9, bool))
    assert buf == b"\xff\xff" and nulls == 0


def test_bool_null_rows_keep_bit_set():
    """ArrowBoolData clears the data bit only for a VALID false (bool_data.cpp:30-34)"""
    t = pa.table({"b": pa.array([True, None, False, None], pa.bool_())})
    (n, nodes, slices, body), = OW.serialize_batches(table_vectors(t), 4, 100)
    data = body[slices[1][0]: slices[1][0] + slices[1][1]]
    assert data == bytes([0b11111011])
    assert nodes == [(4, 2)]


@pytest.mark.parametrize("case,rgs", [("scalar", 2048), ("nested", 1000), ("lineitem", 5000), ("rare", 1000)])
def test_write_oracle_large_offsets(case, rgs):
    """SET arrow_large_buffer_size = true (arrow_appender.cpp:184-188,237-251,288-292): int64 offsets for VARCHAR / BLOB /
    BIT / UUID / LIST, int32 for MAP — the buffers reassemble, as large_utf8 / large_binary / large_list arrays, into the
    same rows, and apart from the offsets buffers the bodies are those of the regular run."""
    if case == "rare":
        cols, expected, types = rare_columns(3000)
        rows, names = 3000, list(expected)
    else:
        table = {"scalar": lambda: scalar_table(5000), "nested": lambda: nested_table(3000),
                 "lineitem": lambda: _lineitem(12_345)}[case]()
        cols, rows, names, types = table_vectors(table), table.num_rows, table.schema.names, [f.type for f in table.schema]
        expected = {n: pylist(table.column(n)) for n in names}
    ltypes = [large_type(t) for t in types]
    assert ltypes != types
    row = 0
    regular = OW.serialize_batches(cols, rows, rgs)
    for (n, nodes, slices, body), (n0, nodes0, slices0, _) in zip(OW.serialize_batches(cols, rows, rgs, large=True), regular):
        assert (n, nodes) == (n0, nodes0) and len(slices) == len(slices0)
        widened = sum(1 for (_, l), (_, l0) in zip(slices, slices0) if l != l0)
        assert widened and all(l in (l0, 2 * l0) for (_, l), (_, l0) in zip(slices, slices0))
        bufs = [body[o:o + l] for o, l in slices]
        for arr, name in zip(arrays_from_buffers(ltypes, nodes, bufs), names):
            arr.validate(full=True)
            assert pylist(arr) == expected[name][row: row + n], name
        row += n
    assert row == rows


@pytest.mark.parametrize("large", [False, True])
@pytest.mark.parametrize("case,rgs", [("nested", 1000), ("nested", 4096), ("rare", 1000)])
def test_write_oracle_list_views(case, rgs, large):
    """SET arrow_output_list_view = true (ArrowListViewData, appender/list_view_data.hpp:59-103): every LIST node carries an
    offsets and a sizes buffer of `rows` entries (NULL row = {running offset, 0}); MAPs stay maps. pyarrow rebuilds the
    same rows from the buffers as list_view / large_list_view arrays."""
    if case == "rare":
        cols, expected, types = rare_columns(3000)
        rows, names = 3000, list(expected)
    else:
        table = nested_table(3000)
        cols, rows, names, types = table_vectors(table), table.num_rows, table.schema.names, [f.type for f in table.schema]
        expected = {n: pylist(table.column(n)) for n in names}
    vtypes = [writer_type(t, large, True) for t in types]
    assert any("list_view" in str(t) for t in vtypes)
    row = 0
    for n, nodes, slices, body in OW.serialize_batches(cols, rows, rgs, large=large, view=True):
        bufs = [body[o:o + l] for o, l in slices]
        for arr, name in zip(arrays_from_buffers(vtypes, nodes, bufs), names):
            arr.validate(full=True)
            assert pylist(arr) == expected[name][row: row + n], name
        row += n
    assert row == rows


@pytest.mark.parametrize("case,rgs", [("scalar", 2048), ("nested", 1000), ("lineitem", 5000)])
def test_write_oracle_string_views(case, rgs):
    """SET produce_arrow_string_view = true (ArrowVarcharToStringViewData, varchar_data.hpp:129-202): VARCHAR columns (also
    below lists / structs / maps) become views + one data buffer holding only the strings longer than 12 bytes; pyarrow
    validates the views (prefix, buffer index, offset) and rebuilds the same rows."""
    table = {"scalar": lambda: scalar_table(5000), "nested": lambda: nested_table(3000),
             "lineitem": lambda: _lineitem(12_345)}[case]()
    cols, names, types = table_vectors(table), table.schema.names, [f.type for f in table.schema]
    vtypes = [writer_type(t, sview=True) for t in types]
    assert vtypes != types
    row = 0
    for n, nodes, slices, body in OW.serialize_batches(cols, table.num_rows, rgs, sview=True):
        bufs = [body[o:o + l] for o, l in slices]
        arrays = arrays_from_buffers(vtypes, nodes, bufs)
        for arr, name in zip(arrays, names):
            arr.validate(full=True)
            assert pylist(arr) == pylist(table.column(name).slice(row, n)), name
        row += n
    assert row == table.num_rows


def test_string_view_layout():
    """the three view forms: NULL = 16 zero bytes; <= 12 bytes inlined and zero padded; longer = prefix + index 0 + offset"""
    t = pa.table({"s": pa.array(["", "abc", None, "exactly12byt", "thirteen bytes", None, "a second long string"])})
    (n, nodes, slices, body), = OW.serialize_batches(table_vectors(t), 7, 100, sview=True)
    views = np.frombuffer(body[slices[1][0]: slices[1][0] + slices[1][1]], np.uint8).reshape(7, 16)
    data = body[slices[2][0]: slices[2][0] + slices[2][1]]
    assert nodes == [(7, 2)] and data == b"thirteen bytesa second long string"
    assert views[0].tobytes() == bytes(16) and views[2].tobytes() == bytes(16) and views[5].tobytes() == bytes(16)
    assert views[1].tobytes() == b"\x03\0\0\0abc" + bytes(9)
    assert views[3].tobytes() == b"\x0c\0\0\0exactly12byt"
    assert views[4].tobytes() == b"\x0e\0\0\0thir" + bytes(4) + bytes(4)
    assert views[6].tobytes() == b"\x14\0\0\0a se" + bytes(4) + b"\x0e\0\0\0"
