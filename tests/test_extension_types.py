"""Arrow extension types the reference registers by default (duckdb/src/common/arrow/
arrow_type_extension.cpp:366-390, looked up by ArrowType::GetTypeFromSchema, arrow_duck_schema.cpp:372-386, applied at
arrow_conversion.cpp:851-860): arrow.uuid, arrow.bool8, arrow.json, arrow.opaque {DuckDB: hugeint, uhugeint, time_tz, bit,
varint}.  The oracle is pinned against pyarrow's reading of the same columns; the GPU path against the oracle."""
import numpy as np
import pyarrow as pa
import pyarrow.ipc as ipc
import pytest

from logical import slab_to_py
from parity import compare_scan_with_oracle


def _fsb(values, width):
    return pa.array([None if v is None else v for v in values], pa.binary(width))


def _table(n, seed=0):
    rng = np.random.default_rng(seed)
    nul = lambda i, k: i % k == k - 1  # noqa: E731
    b8 = pa.ExtensionArray.from_storage(
        pa.bool8(), pa.array([None if nul(i, 9) else int(v) for i, v in enumerate(rng.integers(-3, 4, n))], pa.int8()))
    huge = [None if nul(i, 7) else int(rng.integers(-2**62, 2**62)) * int(rng.integers(1, 2**60)) for i in range(n)]
    hugeint = pa.ExtensionArray.from_storage(
        pa.opaque(pa.binary(16), "hugeint", "DuckDB"),
        _fsb([None if v is None else v.to_bytes(16, "little", signed=True) for v in huge], 16))
    uhuge = [int(rng.integers(0, 2**63)) * int(rng.integers(1, 2**63)) for _ in range(n)]
    uhugeint = pa.ExtensionArray.from_storage(pa.opaque(pa.binary(16), "uhugeint", "DuckDB"),
                                              _fsb([v.to_bytes(16, "little") for v in uhuge], 16))
    ttz = [None if nul(i, 5) else int(rng.integers(0, 2**62)) for i in range(n)]
    time_tz = pa.ExtensionArray.from_storage(
        pa.opaque(pa.binary(8), "time_tz", "DuckDB"), _fsb([None if v is None else v.to_bytes(8, "little") for v in ttz], 8))
    js = [None if nul(i, 6) else '{"k": %d, "pad": "%s"}' % (i, "x" * (i % 17)) for i in range(n)]
    json_s = pa.ExtensionArray.from_storage(pa.json_(pa.string()), pa.array(js, pa.string()))
    json_l = pa.ExtensionArray.from_storage(pa.json_(pa.large_string()), pa.array(js, pa.large_string()))
    bits = [None if nul(i, 8) else bytes([i % 8]) + rng.integers(0, 256, i % 20, dtype=np.uint8).tobytes() for i in range(n)]
    bit = pa.ExtensionArray.from_storage(pa.opaque(pa.binary(), "bit", "DuckDB"), pa.array(bits, pa.binary()))
    varint = pa.ExtensionArray.from_storage(pa.opaque(pa.large_binary(), "varint", "DuckDB"),
                                            pa.array(bits, pa.large_binary()))
    uuid = pa.ExtensionArray.from_storage(pa.uuid(), _fsb([None if nul(i, 4) else rng.bytes(16) for i in range(n)], 16))
    unknown = pa.ExtensionArray.from_storage(pa.opaque(pa.int32(), "mystery", "SomeoneElse"),
                                             pa.array(rng.integers(0, 100, n), pa.int32()))
    cols = {"b8": b8, "hugeint": hugeint, "uhugeint": uhugeint, "time_tz": time_tz, "json": json_s, "json_large": json_l,
            "bit": bit, "varint": varint, "uuid": uuid, "unknown": unknown}
    expect = {"b8": [None if v is None else v != 0 for v in b8.storage.to_pylist()], "hugeint": huge, "uhugeint": uhuge,
              "time_tz": ttz, "json": js, "json_large": js, "bit": bits, "varint": bits,
              "unknown": unknown.storage.to_pylist()}
    return pa.table(cols), expect


def _stream(table, chunk=None, **opt):
    sink = pa.BufferOutputStream()
    with ipc.new_stream(sink, table.schema, options=ipc.IpcWriteOptions(**opt)) as w:
        w.write_table(table, max_chunksize=chunk)
    return sink.getvalue()


def test_oracle_extension_types_match_expected_values():
    from oracle import oracle_py as O
    t, expect = _table(5000, seed=1)
    for name, exp in expect.items():
        arr = t.column(name).chunk(0)
        slab = O.scan_column(arr)
        if name in ("json", "json_large"):
            got = [None if v is None else v.decode() for v in slab_to_py(slab, arr.type.storage_type)]
        elif name in ("hugeint", "uhugeint", "time_tz"):
            raw = slab.data
            got = [int.from_bytes(raw[i].tobytes(), "little", signed=name == "hugeint") if slab.valid[i] else None
                   for i in range(slab.rows)]
        elif name == "b8":
            got = [bool(slab.data[i, 0]) if slab.valid[i] else None for i in range(slab.rows)]
            assert set(np.unique(slab.data[slab.valid])) <= {0, 1}
        else:
            got = slab_to_py(slab, arr.type.storage_type)
        assert got == exp, name


def test_extension_schema_on_host(b2a):
    t, _ = _table(50)
    r = b2a.ArrowIPCReader.open_buffers([_stream(t)])
    assert [c.type for c in r.columns] == ["BOOLEAN", "HUGEINT", "UHUGEINT", "TIME_TZ", "JSON", "JSON", "BIT", "VARINT",
                                           "UUID", "INTEGER"]
    assert [c.op for c in r.columns] == ["BOOL8", "COPY", "COPY", "COPY", "STRING32", "STRING64", "STRING32", "STRING64",
                                         "UUID", "COPY"]

    def field(fmt_type, name, meta=None):
        md = {"ARROW:extension:name": name}
        if meta:
            md["ARROW:extension:metadata"] = meta
        return pa.schema([pa.field("x", fmt_type, metadata=md)])

    def open_schema(schema):
        sink = pa.BufferOutputStream()
        with ipc.new_stream(sink, schema):
            pass
        return b2a.ArrowIPCReader.open_buffers([sink.getvalue()])

    # 1:n extensions are looked up without the format and then refuse formats they do not know
    with pytest.raises(b2a.B2AError, match='Arrow extension type "i" not supported for arrow.json'):
        open_schema(field(pa.int32(), "arrow.json"))
    with pytest.raises(b2a.B2AError, match='Arrow extension type "u" not supported for BIT type'):
        open_schema(field(pa.string(), "arrow.opaque", '{"type_name":"bit","vendor_name":"DuckDB"}'))
    with pytest.raises(b2a.B2AError, match='Arrow extension type "l" not supported for Varint'):
        open_schema(field(pa.int64(), "arrow.opaque", '{"type_name":"varint","vendor_name":"DuckDB"}'))
    # 1:1 extensions over another storage format are simply not found: the storage type stays
    assert open_schema(field(pa.int16(), "arrow.bool8")).columns[0].type == "SMALLINT"
    assert open_schema(field(pa.binary(8), "arrow.opaque",
                             '{"type_name":"hugeint","vendor_name":"DuckDB"}')).columns[0].type == "BLOB"
    assert open_schema(field(pa.binary(16), "arrow.opaque",
                             '{"type_name":"hugeint","vendor_name":"NotDuckDB"}')).columns[0].type == "BLOB"


@pytest.mark.gpu
@pytest.mark.parametrize("host_alias", [True, False])
def test_extension_scan_matches_oracle(b2a, cuda_device, host_alias):
    t, _ = _table(9000, seed=2)
    for chunk in (None, 2500):
        buf = _stream(t, chunk=chunk)
        ds = b2a.scan_arrow_ipc([buf], host_alias=host_alias)
        compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))


@pytest.mark.gpu
def test_extension_scan_compressed(b2a, cuda_device):
    t, _ = _table(6000, seed=3)
    buf = _stream(t, compression="zstd")
    ds = b2a.scan_arrow_ipc([buf])
    compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))
