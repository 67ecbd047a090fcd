"""GPU parity: CUDA decode (through the C ABI) vs the CPU oracle on the same inputs.  Bit-exact in the canonical
form of tests/parity.py.  Inputs: the reference's own fixtures + seeded synthetic lineitem."""
import glob
import os

import numpy as np
import pyarrow as pa
import pyarrow.ipc as ipc
import pytest

from parity import GOLDEN, compare_scan_with_oracle

pytestmark = pytest.mark.gpu

LE = os.path.join(GOLDEN, "arrow-testing", "1.0.0-littleendian")
BE = os.path.join(GOLDEN, "arrow-testing", "1.0.0-bigendian")

# the 14 files of the reference's test/python/test_integration.py:19 (+ dictionary files, beyond the reference)
INTEGRATION = ["generated_null_trivial", "generated_primitive_large_offsets", "generated_custom_metadata",
               "generated_datetime", "generated_decimal", "generated_map_non_canonical", "generated_map",
               "generated_nested_large_offsets", "generated_nested", "generated_null",
               "generated_primitive_no_batches", "generated_primitive_zerolength", "generated_primitive",
               "generated_recursive_nested"]
DICTIONARY = ["generated_dictionary", "generated_dictionary_unsigned"]


def _pa_batches(path):
    with open(path, "rb") as f:
        data = f.read()
    if data[:6] == b"ARROW1":
        r = ipc.open_file(pa.BufferReader(data))
        return [r.get_batch(i) for i in range(r.num_record_batches)]
    return list(ipc.open_stream(pa.BufferReader(data)))


@pytest.mark.parametrize("name", ["test.arrows", "test.arrow", "fruit.arrow"])
def test_reference_data_files(b2a, cuda_device, name):
    path = os.path.join(GOLDEN, "data", name)
    ds = b2a.read_arrow(path)
    batches = _pa_batches(path)
    assert ds.num_rows == sum(len(b) for b in batches)
    compare_scan_with_oracle(ds, batches)


def test_read_arrow_test_counts(b2a, cuda_device):
    """test/sql/read_arrow.test:36-38 — count(*) = 15487"""
    ds = b2a.read_arrow(os.path.join(GOLDEN, "data", "test.arrows"))
    assert ds.num_rows == 15487


@pytest.mark.parametrize("name", INTEGRATION + DICTIONARY)
@pytest.mark.parametrize("endian", ["little", "big"])
def test_arrow_testing_integration(b2a, cuda_device, name, endian):
    path = os.path.join(LE if endian == "little" else BE, name + ".stream")
    ds = b2a.read_arrow(path)
    compare_scan_with_oracle(ds, _pa_batches(path))


def test_arrow_file_format(b2a, cuda_device):
    path = os.path.join(LE, "generated_primitive.arrow_file")
    ds = b2a.read_arrow(path)
    compare_scan_with_oracle(ds, _pa_batches(path))


def test_projection(b2a, cuda_device):
    path = os.path.join(GOLDEN, "data", "test.arrows")
    ds = b2a.read_arrow(path, columns=["message", "files"])
    batches = _pa_batches(path)
    compare_scan_with_oracle(ds, batches)
    names = [f.name for f in ds.reader.output_columns()]
    assert names == ["message", "files"]


def test_lineitem_sf001_parquet_fixture(b2a, cuda_device, tmp_path):
    """data/parquet-testing/lineitem_sf0_01.parquet (decimal128 -> int64 narrowing, int32 keys) as IPC."""
    import pyarrow.parquet as pq
    t = pq.read_table(os.path.join(GOLDEN, "data", "lineitem_sf0_01.parquet"))
    p = str(tmp_path / "li.arrows")
    with ipc.new_stream(p, t.schema) as w:
        for b in t.to_batches(max_chunksize=20000):
            w.write_batch(b)
    ds = b2a.read_arrow(p)
    assert ds.num_rows == 60175
    compare_scan_with_oracle(ds, _pa_batches(p))


@pytest.mark.parametrize("host_alias", [True, False])
@pytest.mark.parametrize("rows,batch_rows", [(300_000, 122_880), (5000, 1000), (2048, 2048), (2049, 4096), (1, 1)])
def test_synthetic_lineitem(b2a, cuda_device, rows, batch_rows, host_alias):
    """host_alias=True: the drop-in host path (fixed-width vectors alias the host stream like DirectConversion);
    False: every column through upload -> decode kernel -> download."""
    from tools import lineitem
    buf = lineitem.stream_bytes(rows, seed=7, batch_rows=batch_rows)
    ds = b2a.scan_arrow_ipc([buf], host_alias=host_alias)
    compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))
    if host_alias:
        co = ds.scan.column(0, 0)
        assert co.data_off == b2a.B2A_NONE and co.image_off != b2a.B2A_NONE
        assert ds.scan.upload_bytes < ds.scan.image_bytes or rows < 100


@pytest.mark.parametrize("name", ["generated_primitive", "generated_nested", "generated_datetime"])
@pytest.mark.parametrize("endian", ["little", "big"])
def test_device_layout_path_on_integration_files(b2a, cuda_device, name, endian):
    """same files as above through the non-aliased (device-resident layout) path"""
    path = os.path.join(LE if endian == "little" else BE, name + ".stream")
    ds = b2a.read_arrow(path, host_alias=False)
    compare_scan_with_oracle(ds, _pa_batches(path))


@pytest.mark.parametrize("host_alias", [True, False])
def test_string_views_written_by_pyarrow(b2a, cuda_device, host_alias):
    """Utf8View columns (SetVectorStringView, arrow_conversion.cpp:425-452; a10): inlined, prefixed and NULL
    views, top level and below a list, several batches — vs the oracle over pyarrow's decoding of the same stream"""
    rng = np.random.default_rng(21)
    words = ["", "a", "twelve bytes", "thirteen bytes", "a considerably longer string value", "x" * 100]
    n = 9000
    pick = rng.integers(0, len(words), n)
    s = [None if rng.random() < 0.1 else words[int(k)] + str(i % 97) for i, k in enumerate(pick)]
    t = pa.table({
        "sv": pa.array(s, pa.string_view()),
        "l_sv": pa.array([None if i % 11 == 0 else [s[i], s[(i * 7) % n]][: i % 3] for i in range(n)],
                         pa.list_(pa.string_view())),
        "i": pa.array(np.arange(n, dtype=np.int64)),
    })
    sink = pa.BufferOutputStream()
    with ipc.new_stream(sink, t.schema) as w:
        for b in t.to_batches(max_chunksize=2500):
            w.write_batch(b)
    buf = sink.getvalue().to_pybytes()
    ds = b2a.scan_arrow_ipc([buf], host_alias=host_alias)
    compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))


def test_host_alias_scan_rejects_device_consumers(b2a, cuda_device):
    from tools import lineitem
    buf = lineitem.stream_bytes(5000, seed=2)
    r = b2a.ArrowIPCReader.open_buffers([buf])
    scan = b2a.ArrowScan(r, device=0, host_alias=True)
    with pytest.raises(b2a.B2AError):
        scan.decode()


def test_synthetic_lineitem_dictionary(b2a, cuda_device):
    from tools import lineitem
    buf = lineitem.stream_bytes(200_000, seed=11, dictionary_strings=True)
    ds = b2a.scan_arrow_ipc([buf])
    compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))


# ------------------------------------------------------------------------------------------------ K9 + consumer
def _device_scan(b2a, buf_or_path):
    r = (b2a.ArrowIPCReader.open_file(buf_or_path) if isinstance(buf_or_path, str)
         else b2a.ArrowIPCReader.open_buffers([buf_or_path]))
    scan = b2a.ArrowScan(r, device=0)
    scan.upload()
    scan.decode()
    scan.finish()
    return r, scan


def _check_filter(scan, table: pa.Table, preds, sel, count, total, batch_rows):
    """oracle for pushed-down filters = scan, then filter (numpy over pyarrow's decoding), per 2 048-row vector"""
    import operator
    ops = {"=": operator.eq, "!=": operator.ne, "<": operator.lt, "<=": operator.le, ">": operator.gt, ">=": operator.ge}
    mask = np.ones(table.num_rows, bool)
    for col, op, const in preds:
        c = table.column(col).combine_chunks()
        valid = np.asarray(c.is_valid())
        t = c.type
        if pa.types.is_decimal(t):
            vals = np.array([int(x.as_py().scaleb(t.scale)) if x.is_valid else 0 for x in c], dtype=np.int64) \
                if table.num_rows < 100000 else None
            if vals is None:
                vals = np.frombuffer(c.buffers()[1], dtype=np.int64 if t.bit_width == 64 else np.int64)[: len(c)]
        elif pa.types.is_date32(t):
            vals = np.frombuffer(c.buffers()[1], dtype=np.int32)[c.offset: c.offset + len(c)].astype(np.int64)
        else:
            vals = c.fill_null(0).to_numpy().astype(np.int64)
        mask &= valid & ops[op](vals, const)
    sel, count = sel.cpu().numpy().view(np.uint32), count.cpu().numpy().view(np.uint32)
    assert int(total.item()) == int(mask.sum())
    chunk, row = 0, 0
    for n in batch_rows:
        for k in range(0, n, 2048):
            m = mask[row + k: row + min(n, k + 2048)]
            exp = np.nonzero(m)[0].astype(np.uint32)
            assert count[chunk] == exp.size, f"chunk {chunk}: count {count[chunk]} != {exp.size}"
            assert np.array_equal(sel[chunk * 2048: chunk * 2048 + exp.size], exp), f"chunk {chunk}: sel differs"
            chunk += 1
        row += n
    return mask


def test_filter_and_q6_on_synthetic_lineitem(b2a, cuda_device):
    from tools import lineitem
    buf = lineitem.stream_bytes(400_000, seed=5)
    r, scan = _device_scan(b2a, buf)
    preds = [("l_shipdate", ">=", 8766), ("l_shipdate", "<", 9131), ("l_quantity", "<", 2400)]
    sel, count, total = scan.filter(preds)
    t = ipc.open_stream(buf).read_all()
    brows = [r.batch_info(i).rows for i in range(r.num_batches)]
    mask = _check_filter(scan, t, preds, sel, count, total, brows)
    price = np.frombuffer(t.column("l_extendedprice").combine_chunks().buffers()[1], dtype=np.int64)
    disc = np.frombuffer(t.column("l_discount").combine_chunks().buffers()[1], dtype=np.int64)
    exp = sum(int(a) * int(b) for a, b in zip(price[mask], disc[mask]))
    assert scan.sum_product("l_extendedprice", "l_discount", sel, count) == exp
    exp_all = int((price.astype(object) * disc.astype(object)).sum())
    assert scan.sum_product("l_extendedprice", "l_discount") == exp_all


def test_q6_known_answer_sf001(b2a, cuda_device, tmp_path):
    """TPC-H Q6 on data/parquet-testing/lineitem_sf0_01.parquet = 1193053.2253 (test/nodejs/arrow_test.js:424)."""
    import pyarrow.parquet as pq
    t = pq.read_table(os.path.join(GOLDEN, "data", "lineitem_sf0_01.parquet"))
    p = str(tmp_path / "li.arrows")
    with ipc.new_stream(p, t.schema) as w:
        for b in t.to_batches(max_chunksize=122880):
            w.write_batch(b)
    r, scan = _device_scan(b2a, p)
    preds = [("l_shipdate", ">=", 8766), ("l_shipdate", "<", 9131), ("l_discount", ">=", 5), ("l_discount", "<=", 7),
             ("l_quantity", "<", 24)]
    sel, count, total = scan.filter(preds)
    revenue = scan.sum_product("l_extendedprice", "l_discount", sel, count)
    assert revenue == 11930532253, revenue  # DECIMAL(30,4): 1193053.2253


def test_filter_with_nulls(b2a, cuda_device):
    rng = np.random.default_rng(1)
    n = 10_000
    vals = rng.integers(-50, 50, n)
    arr = pa.array(vals, type=pa.int32(), mask=rng.random(n) < 0.3)
    big = pa.array(rng.integers(-2**40, 2**40, n), type=pa.int64())
    t = pa.table({"a": arr, "b": big})
    sink = pa.BufferOutputStream()
    with ipc.new_stream(sink, t.schema) as w:
        for b in t.to_batches(max_chunksize=3000):
            w.write_batch(b)
    buf = sink.getvalue()
    r, scan = _device_scan(b2a, buf)
    preds = [("a", ">", -10), ("b", "<=", 12345)]
    sel, count, total = scan.filter(preds)
    _check_filter(scan, t, preds, sel, count, total, [r.batch_info(i).rows for i in range(r.num_batches)])
