"""b2a_scan_filter_no_alloc (include/b2arrow.h): the filter launch for callers that may not allocate device memory — its
per-vector tables live in a fixed `static __device__` array instead of cudaMalloc'ed memory.  No counterpart in the
reference.  Checked here through the C ABI against b2a_scan_filter on the
dbgen fixture: selection vectors, per-vector counts and the total must be IDENTICAL (integer work: bit-exact, the bar of
every other filter test), on two calls in a row with different predicate sets (the second call reuses the array with the
first call's contents still in it), and again after a third call with the first set.

Collected last on purpose (the file name)."""
import ctypes as C
import os

import numpy as np
import pytest

from parity import GOLDEN

DBGEN = os.path.join(GOLDEN, "dbgen", "lineitem_sf0.002_rg2500.arrows")

Q6 = [("l_shipdate", ">=", 8766), ("l_shipdate", "<", 9131), ("l_discount", ">=", 5), ("l_discount", "<=", 7),
      ("l_quantity", "<", 2400)]
OTHER = [("l_quantity", ">=", 1000), ("l_orderkey", "<", 6000)]


def _predicates(b2a, scan, preds):
    from duckdb_arrow_b200 import _lib
    names = [f.name for f in scan.reader.output_columns()]
    arr = (_lib.Predicate * len(preds))()
    for i, (col, op, const) in enumerate(preds):
        arr[i].column = names.index(col)
        arr[i].cmp = _lib.CMP[op]
        arr[i].constant = int(const)
    return arr


def _run(b2a, scan, entry, preds):
    """one filter call through the C ABI -> (sel int32[n_chunks, 2048], count, total) on the host"""
    import torch
    from duckdb_arrow_b200 import _lib
    dev = f"cuda:{scan.device}"
    n = max(1, scan.n_chunks)
    sel = torch.full((n * 2048,), -1, dtype=torch.int32, device=dev)
    count = torch.full((n,), -1, dtype=torch.int32, device=dev)
    total = torch.full((1,), -1, dtype=torch.int64, device=dev)
    arr = _predicates(b2a, scan, preds)
    fn = getattr(_lib.lib, entry)
    _lib.check(fn(scan._h, C.c_void_p(scan.d_out.data_ptr()), arr, len(preds), C.c_void_p(sel.data_ptr()),
                  C.c_void_p(count.data_ptr()), C.c_void_p(total.data_ptr()), scan._stream_ptr()))
    torch.cuda.synchronize()
    return sel.cpu().numpy().reshape(n, 2048), count.cpu().numpy(), int(total.item())


def _same(a, b):
    sel_a, cnt_a, tot_a = a
    sel_b, cnt_b, tot_b = b
    assert tot_a == tot_b
    assert np.array_equal(cnt_a, cnt_b)
    keep = np.arange(2048)[None, :] < cnt_a[:, None]  # entries past a vector's count are not written by either
    assert np.array_equal(sel_a[keep], sel_b[keep])


@pytest.mark.gpu
def test_filter_no_alloc_equals_filter_on_two_calls_in_a_row(b2a, cuda_device):
    r = b2a.ArrowIPCReader.open_file(DBGEN)
    scan = b2a.ArrowScan(r, device=0)
    scan.upload()
    scan.decode()
    scan.finish()
    # the no-alloc entry point FIRST (a fresh process has never touched the array), twice in a row, then the existing one
    first = _run(b2a, scan, "b2a_scan_filter_no_alloc", Q6)
    second = _run(b2a, scan, "b2a_scan_filter_no_alloc", OTHER)
    third = _run(b2a, scan, "b2a_scan_filter_no_alloc", Q6)
    want_q6 = _run(b2a, scan, "b2a_scan_filter", Q6)
    want_other = _run(b2a, scan, "b2a_scan_filter", OTHER)
    _same(first, want_q6)
    _same(second, want_other)
    _same(third, want_q6)
    assert first[2] == 232  # tests/golden/dbgen/answers.json: q6 count
    assert 0 < second[2] < r.num_rows


@pytest.mark.gpu
def test_filter_no_alloc_rejects_what_does_not_fit(b2a, cuda_device):
    """more than 4 MiB of tables: 8 predicates x 16 B + 4 B = 132 B per vector -> 31 775 vectors at most; a stream of
    32 000 one-row batches is 32 000 vectors.  B2A_EINVAL with a message that names the limit; b2a_scan_filter takes it."""
    import pyarrow as pa
    import pyarrow.ipc as ipc
    n = 32_000
    schema = pa.schema([("v", pa.int32())])
    sink = pa.BufferOutputStream()
    with ipc.new_stream(sink, schema) as w:
        one = pa.array(np.arange(n, dtype=np.int32))
        for i in range(n):
            w.write_batch(pa.RecordBatch.from_arrays([one.slice(i, 1)], schema=schema))
    stream = sink.getvalue().to_pybytes()
    r = b2a.ArrowIPCReader.open_buffers([stream])
    scan = b2a.ArrowScan(r, device=0)
    assert scan.n_chunks == n
    scan.upload()
    scan.decode()
    scan.finish()
    eight = [("v", ">=", 0)] * 7 + [("v", "<", 100)]
    with pytest.raises(b2a.B2AError, match="fixed device scratch holds 4194304"):
        _run(b2a, scan, "b2a_scan_filter_no_alloc", eight)
    # with fewer predicates the same scan fits (2 x 16 + 4 = 36 B per vector) and agrees with the existing entry point
    two = [("v", ">=", 10), ("v", "<", 100)]
    got = _run(b2a, scan, "b2a_scan_filter_no_alloc", two)
    _same(got, _run(b2a, scan, "b2a_scan_filter", two))
    assert got[2] == 90
    _same(_run(b2a, scan, "b2a_scan_filter", eight), _run(b2a, scan, "b2a_scan_filter", eight))
