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


@pytest.mark.parametrize("rows,batch_rows", [(300_000, 122_880), (5000, 1000), (2048, 2048), (2049, 4096), (1, 1)])
def test_synthetic_lineitem(b2a, cuda_device, rows, batch_rows):
    from tools import lineitem
    buf = lineitem.stream_bytes(rows, seed=7, batch_rows=batch_rows)
    ds = b2a.scan_arrow_ipc([buf])
    compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))


def test_synthetic_lineitem_dictionary(b2a, cuda_device):
    from tools import lineitem
    buf = lineitem.stream_bytes(200_000, seed=11, dictionary_strings=True)
    ds = b2a.scan_arrow_ipc([buf])
    compare_scan_with_oracle(ds, list(ipc.open_stream(buf)))
