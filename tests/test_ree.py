"""Run-end encoded arrays (ColumnArrowToDuckDBRunEndEncoded + FlattenRunEnds, arrow_conversion.cpp:551-767).
The reference's tests hold no REE fixture, so the oracle's restatement is pinned against pyarrow's own flattening of the
same arrays; the GPU path is then compared with the oracle bit for bit."""
import numpy as np
import pyarrow as pa
import pyarrow.ipc as ipc
import pytest

from logical import slab_to_py
from parity import compare_scan_with_oracle


def _ree_table(n=9000, seed=0):
    rng = np.random.default_rng(seed)

    def ends(run_type, max_len):
        e = np.cumsum(rng.integers(1, max_len, n))
        e = e[e < n]
        return pa.array(np.append(e, n).astype(run_type))

    cols, names = [], []
    e16, e32, e32b = ends(np.int16, 40), ends(np.int32, 9), ends(np.int32, 3000)
    cols.append(pa.RunEndEncodedArray.from_arrays(e32, pa.array(rng.integers(-10**12, 10**12, len(e32)), pa.int64())))
    names.append("i64")
    v = [None if i % 5 == 0 else int(i) for i in range(len(e16))]
    cols.append(pa.RunEndEncodedArray.from_arrays(e16, pa.array(v, pa.int32())))
    names.append("i32_nulls_run16")
    s = [None if i % 7 == 0 else ("x" * (i % 25)) + str(i) for i in range(len(e32b))]
    cols.append(pa.RunEndEncodedArray.from_arrays(e32b, pa.array(s, pa.string())))
    names.append("str_nulls")
    cols.append(pa.RunEndEncodedArray.from_arrays(e32, pa.array(rng.random(len(e32)) < 0.5, pa.bool_())))
    names.append("bool")
    cols.append(pa.RunEndEncodedArray.from_arrays(e16, pa.array(rng.standard_normal(len(e16)), pa.float64())))
    names.append("f64")
    cols.append(pa.array(rng.integers(0, 100, n), pa.int32()))
    names.append("plain")
    return pa.table(cols, names=names)


def _stream(table, chunk=None, **opt):
    sink = pa.BufferOutputStream()
    with ipc.new_stream(sink, table.schema, options=ipc.IpcWriteOptions(**opt)) as w:
        w.write_table(table, max_chunksize=chunk)
    return sink.getvalue()


def test_oracle_ree_matches_pyarrow_flattening():
    from oracle import oracle_py as O
    t = _ree_table()
    for name in t.column_names:
        arr = t.column(name).chunk(0)
        got = slab_to_py(O.scan_column(arr), arr.type)
        got = [g.decode() if isinstance(g, bytes) else g for g in got]
        assert got == arr.to_pylist(), name


def test_ree_schema_and_plan_on_host(b2a):
    buf = _stream(_ree_table())
    r = b2a.ArrowIPCReader.open_buffers([buf])
    assert [c.type for c in r.columns] == ["BIGINT", "INTEGER", "VARCHAR", "BOOLEAN", "DOUBLE", "INTEGER"]
    assert [c.op for c in r.columns][:5] == ["REE"] * 5
    s = b2a.ArrowScan(r, device=-1)
    assert s.n_tiles > 0
    # nested / dictionary values are refused like FlattenRunEndsSwitch's default branch
    bad = pa.table({"x": pa.RunEndEncodedArray.from_arrays(pa.array([2, 3], pa.int32()),
                                                           pa.array([[1], [2, 3]], pa.list_(pa.int32())))})
    with pytest.raises(b2a.B2AError, match="RunEndEncoded value type"):
        b2a.ArrowIPCReader.open_buffers([_stream(bad)])


@pytest.mark.gpu
@pytest.mark.parametrize("host_alias", [True, False])
def test_ree_scan_matches_oracle(b2a, cuda_device, host_alias):
    buf = _stream(_ree_table())
    ds = b2a.scan_arrow_ipc([buf], host_alias=host_alias)
    compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))


@pytest.mark.gpu
def test_ree_compressed_and_in_struct(b2a, cuda_device):
    t = _ree_table(5000, seed=4)
    st = pa.StructArray.from_arrays([t.column("i64").chunk(0), t.column("plain").chunk(0)], ["r", "p"],
                                    mask=pa.array([i % 9 == 0 for i in range(t.num_rows)]))
    t2 = pa.table({"s": st, "str": t.column("str_nulls")})
    for opt in ({}, {"compression": "zstd"}):
        buf = _stream(t2, **opt)
        ds = b2a.scan_arrow_ipc([buf])
        compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))
