"""List views (ConvertArrowListViewOffsetsTemplated, arrow_conversion.cpp:166-203).  The reference's tests
hold no list-view fixture: the oracle's restatement is pinned against pyarrow's own reading of well-formed (in-order,
gap-free) views, where the reference's window rule — child vector = sum(sizes) rows from the chunk's lowest offset — is
lossless; out-of-order / overlapping / gapped views are then compared GPU vs oracle bit for bit (entries, per-chunk child
windows, child rows), reproducing the reference's behaviour whatever it is."""
import numpy as np
import pyarrow as pa
import pyarrow.ipc as ipc
import pytest

from logical import slab_to_py
from parity import compare_scan_with_oracle


def _lengths(n, rng, max_len=6):
    ln = rng.integers(0, max_len, n)
    ln[rng.random(n) < 0.2] = 0
    return ln


def _wellformed(n, seed=0, large=False, child="int"):
    rng = np.random.default_rng(seed)
    ln = _lengths(n, rng)
    offs = np.concatenate([[0], np.cumsum(ln)])[:-1]
    total = int(ln.sum())
    if child == "int":
        values = pa.array([None if i % 9 == 4 else i * 3 for i in range(total)], pa.int64())
    elif child == "str":
        values = pa.array([None if i % 7 == 2 else "v" * (i % 20) + str(i) for i in range(total)], pa.string())
    else:
        values = pa.StructArray.from_arrays(
            [pa.array(range(total), pa.int32()), pa.array([str(i) if i % 3 else None for i in range(total)], pa.string())],
            ["a", "b"], mask=pa.array([i % 10 == 0 for i in range(total)], pa.bool_()))
    dt = np.int64 if large else np.int32
    cls = pa.LargeListViewArray if large else pa.ListViewArray
    mask = pa.array([i % 13 == 5 for i in range(n)], pa.bool_())
    return cls.from_arrays(pa.array(offs.astype(dt)), pa.array(ln.astype(dt)), values, mask=mask)


def _adversarial(n, seed=0):
    """views in random order, overlapping, with gaps, empty views with far-away offsets"""
    rng = np.random.default_rng(seed)
    total = 3 * n + 10
    values = pa.array(rng.integers(-1000, 1000, total), pa.int64())
    sizes = _lengths(n, rng, 5)
    offs = rng.integers(0, total - 6, n)
    return pa.ListViewArray.from_arrays(pa.array(offs.astype(np.int32)), pa.array(sizes.astype(np.int32)), values)


def _stream(table, chunk=None, **opt):
    sink = pa.BufferOutputStream()
    with ipc.new_stream(sink, table.schema, options=ipc.IpcWriteOptions(**opt)) as w:
        w.write_table(table, max_chunksize=chunk)
    return sink.getvalue()


def _norm(v):
    if isinstance(v, bytes):
        return v.decode()
    if isinstance(v, list):
        return [_norm(x) for x in v]
    if isinstance(v, tuple):
        return tuple(_norm(x) for x in v)
    return v


def _expect(arr):
    out = arr.to_pylist()

    def conv(x):
        if isinstance(x, dict):
            return tuple(conv(v) for v in x.values())
        if isinstance(x, list):
            return [conv(v) for v in x]
        return x
    return conv(out)


@pytest.mark.parametrize("child", ["int", "str", "struct"])
@pytest.mark.parametrize("large", [False, True])
def test_oracle_listview_matches_pyarrow(child, large):
    from oracle import oracle_py as O
    for n in (0, 1, 100, 2048, 5000):
        arr = _wellformed(n, seed=n + 1, large=large, child=child)
        got = _norm(slab_to_py(O.scan_column(arr), arr.type))
        assert got == _expect(arr), (n, child, large)


def test_listview_schema_on_host(b2a):
    t = pa.table([_wellformed(300), _wellformed(300, large=True, child="str")], names=["lv", "llv"])
    r = b2a.ArrowIPCReader.open_buffers([_stream(t)])
    assert [c.type for c in r.columns] == ["LIST", "LIST"]
    assert [c.op for c in r.columns] == ["LISTVIEW32", "LISTVIEW64"]
    assert b2a.ArrowScan(r, device=-1).n_tiles > 0
    inner = pa.ListViewArray.from_arrays(pa.array([0, 1], pa.int32()), pa.array([1, 1], pa.int32()),
                                         pa.array([[1], [2]], pa.list_(pa.int32())))
    with pytest.raises(b2a.B2AError, match="below a list view"):
        b2a.ArrowScan(b2a.ArrowIPCReader.open_buffers([_stream(pa.table([inner], names=["x"]))]), device=-1)


@pytest.mark.gpu
@pytest.mark.parametrize("host_alias", [True, False])
def test_listview_scan_matches_oracle(b2a, cuda_device, host_alias):
    n = 7000
    st = pa.StructArray.from_arrays([_wellformed(n, seed=4, child="str"), pa.array(range(n), pa.int32())], ["lv", "k"],
                                    mask=pa.array([i % 17 == 0 for i in range(n)], pa.bool_()))
    t = pa.table([_wellformed(n, seed=1), _wellformed(n, seed=2, large=True, child="str"),
                  _wellformed(n, seed=3, child="struct"), _adversarial(n, seed=5), st, pa.array(range(n), pa.int64())],
                 names=["lv_int", "llv_str", "lv_struct", "lv_adversarial", "st", "plain"])
    for chunk in (None, 3000):
        buf = _stream(t, chunk=chunk)
        ds = b2a.scan_arrow_ipc([buf], host_alias=host_alias)
        compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))


@pytest.mark.gpu
def test_listview_compressed_and_logical(b2a, cuda_device):
    from logical import column_to_py
    arr, sarr = _wellformed(5000, seed=9), _wellformed(5000, seed=10, child="str")
    t = pa.table([arr, sarr], names=["lv", "lv_str"])
    buf = _stream(t, compression="zstd")
    ds = b2a.scan_arrow_ipc([buf])
    compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))
    # end to end: the decoded vectors read back as the lists pyarrow reads
    assert column_to_py(ds.column(0, ds.reader.columns[0].index), arr.type) == _expect(arr)


@pytest.mark.gpu
def test_listview_out_of_bounds_raises(b2a, cuda_device):
    values = pa.array(range(10), pa.int64())
    bad = pa.ListViewArray.from_arrays(pa.array([0, 8], pa.int32()), pa.array([2, 5], pa.int32()), values)
    with pytest.raises(b2a.B2AError, match="offsets"):
        b2a.scan_arrow_ipc([_stream(pa.table([bad], names=["x"]))])
