"""GPU tests of this round's additions, through the C ABI:
  * a REFERENCE-written lineitem stream (tests/golden/dbgen: DuckDB dbgen -> the reference's ArrowAppender, validity
    bitmaps on every column) decodes bit-exactly, and its answers equal the reference's own (answers.json);
  * the same vectors compared DIRECTLY with the reference's arrow_scan (oracle/_ref) where that library is present;
  * b2a_read_arrow_device (host file -> device vectors for device consumers) + filter + aggregate;
  * pushed-down predicates on unsigned columns; malformed inputs (views / lists / dictionary indices out of range)."""
import json
import os

import numpy as np
import pyarrow as pa
import pyarrow.ipc as ipc
import pytest

from parity import GOLDEN, compare_column, compare_scan_with_oracle

pytestmark = pytest.mark.gpu

DBGEN = os.path.join(GOLDEN, "dbgen", "lineitem_sf0.002_rg2500.arrows")
ANSWERS = json.load(open(os.path.join(GOLDEN, "dbgen", "answers.json")))
Q6_PREDS = [("l_shipdate", ">=", 8766), ("l_shipdate", "<", 9131), ("l_discount", ">=", 5), ("l_discount", "<=", 7),
            ("l_quantity", "<", 2400)]


def _stream(table_or_batches, schema=None, chunk=None) -> pa.Buffer:
    sink = pa.BufferOutputStream()
    if isinstance(table_or_batches, pa.Table):
        schema = table_or_batches.schema
        batches = table_or_batches.to_batches(max_chunksize=chunk or 1 << 20)
    else:
        batches = table_or_batches
    with ipc.new_stream(sink, schema) as w:
        for b in batches:
            w.write_batch(b)
    return sink.getvalue()


@pytest.mark.parametrize("host_alias", [True, False])
def test_reference_written_lineitem_decodes_bit_exact(b2a, cuda_device, host_alias):
    batches = list(ipc.open_stream(DBGEN))
    assert sum(b.num_rows for b in batches) == ANSWERS["rows"] == 11957 and len(batches) == 5
    ds = b2a.read_arrow(DBGEN, host_alias=host_alias)
    assert ds.num_rows == ANSWERS["rows"]
    # every column of the file carries an (all-valid) bitmap the scan must skip like the reference (arrow_conversion.cpp:71)
    for c in batches[0].columns:
        assert c.null_count == 0
    compare_scan_with_oracle(ds, batches)


def test_reference_written_lineitem_against_the_reference_itself(b2a, cuda_device):
    """the GPU vectors against the vectors the reference's own ArrowToDuckDB produces (no restatement in between)"""
    from oracle import ref_py as R
    if not R.available():
        pytest.skip("oracle/_ref not in this tree")
    batches = list(ipc.open_stream(DBGEN))
    ds = b2a.read_arrow(DBGEN, host_alias=False)
    for bi, batch in enumerate(batches):
        for f in ds.reader.output_columns():
            arr = batch.column(f.name)
            compare_column(ds.column(bi, f.index), R.scan_column(arr), arr, f"batch{bi}.{f.name}")


def test_q6_on_the_reference_written_stream(b2a, cuda_device):
    """answers.json holds what the reference's DuckDB answers on the dbgen table the stream was exported from"""
    r = b2a.ArrowIPCReader.open_file(DBGEN)
    scan = b2a.ArrowScan(r, device=0)
    scan.upload()
    scan.decode()
    scan.finish()
    sel, count, total = scan.filter(Q6_PREDS)
    n, s = ANSWERS["q6_count_sum"]
    assert int(total.item()) == int(n)
    revenue = scan.sum_product("l_extendedprice", "l_discount", sel, count)
    assert f"{revenue // 10000}.{revenue % 10000:04d}" == s
    _, _, t2 = scan.filter([("l_quantity", "<", 2400)])
    assert int(t2.item()) == int(ANSWERS["count_quantity_lt_24"])
    _, _, t3 = scan.filter([("l_shipdate", ">=", 8766), ("l_shipdate", "<", 9131)])
    assert int(t3.item()) == int(ANSWERS["count_shipdate_1994"])


@pytest.mark.parametrize("source", ["file", "pinned"])
def test_read_arrow_device_feeds_device_consumers(b2a, cuda_device, source):
    """b2a_read_arrow_device: projected buffers only cross PCIe (staged from the mmap'ed file / copied from a pinned
    image), vectors stay in HBM, filter + aggregate give the reference's Q6 answer; repeated calls reuse the pipeline"""
    import torch
    if source == "file":
        r = b2a.ArrowIPCReader.open_file(DBGEN)
    else:
        raw = open(DBGEN, "rb").read()
        pinned = torch.empty(len(raw), dtype=torch.uint8, pin_memory=True)
        pinned.numpy()[:] = np.frombuffer(raw, np.uint8)
        r = b2a.ArrowIPCReader.open_buffers([(pinned.data_ptr(), len(raw))])
    r.set_projection(["l_quantity", "l_extendedprice", "l_discount", "l_shipdate"])
    scan = b2a.ArrowScan(r, device=0)
    full = b2a.ArrowScan(b2a.ArrowIPCReader.open_file(DBGEN), device=0)
    # 4 of 16 columns; on this 2 500-row-batch fixture neighbouring small ranges are merged into one copy (scan_plan.cpp:128)
    assert scan.upload_bytes * 3 < full.upload_bytes, "projection must narrow what crosses PCIe"
    d_out = torch.empty(scan.out_bytes, dtype=torch.uint8, device="cuda:0")
    n, s = ANSWERS["q6_count_sum"]
    for n_slices in (0, 2, 1):
        d_out.fill_(0xAB)
        scan.read_device(d_out, n_slices)
        sel, count, total = scan.filter(Q6_PREDS, d_out)
        res = scan.sum_product_async("l_extendedprice", "l_discount", sel, count, d_out)
        torch.cuda.synchronize()
        scan.finish()
        assert int(total.item()) == int(n)
        lo, hi = [int(x) for x in res.cpu().tolist()]
        revenue = (hi << 64) | (lo & 0xFFFFFFFFFFFFFFFF)
        assert f"{revenue // 10000}.{revenue % 10000:04d}" == s
    # and the vectors are the ones the plain device path decodes
    ref_scan = b2a.ArrowScan(r, device=0)
    ref_scan.upload()
    exp = ref_scan.decode()
    ref_scan.finish()
    for b in range(r.num_batches):
        for f in r.output_columns():
            co = scan.column(b, f.index)
            a = d_out[co.data_off: co.data_off + co.rows * f.out_width]
            e = exp[co.data_off: co.data_off + co.rows * f.out_width]
            assert torch.equal(a, e), f"batch {b} {f.name}"


def test_read_arrow_device_refuses_host_alias_scans(b2a, cuda_device):
    import torch
    r = b2a.ArrowIPCReader.open_file(DBGEN)
    scan = b2a.ArrowScan(r, device=0, host_alias=True)
    d_out = torch.empty(max(256, scan.out_bytes), dtype=torch.uint8, device="cuda:0")
    with pytest.raises(b2a.B2AError, match="HOST_ALIAS"):
        scan.read_device(d_out)


def test_filter_on_unsigned_columns(b2a, cuda_device):
    """UTINYINT / USMALLINT / UINTEGER predicate columns are zero-extended: values with the top bit set compare as the
    large positive numbers they are (DuckDB ConstantFilter on unsigned types)"""
    from test_gpu_parity import _check_filter, _device_scan
    rng = np.random.default_rng(3)
    n = 9000
    t = pa.table({
        "u8": pa.array(rng.integers(0, 256, n).astype(np.uint8), pa.uint8()),
        "u16": pa.array(rng.integers(0, 65536, n).astype(np.uint16), pa.uint16(), mask=rng.random(n) < 0.1),
        "u32": pa.array(rng.integers(0, 2**32, n).astype(np.uint32), pa.uint32()),
        "i32": pa.array(rng.integers(-2**31, 2**31, n).astype(np.int32), pa.int32()),
    })
    r, scan = _device_scan(b2a, _stream(t, chunk=4000))
    brows = [r.batch_info(i).rows for i in range(r.num_batches)]
    for preds in ([("u32", ">=", 2**31)], [("u32", "<", 2**31 + 12345), ("u8", ">", 200)],
                  [("u16", ">=", 40000), ("i32", "<", 0)], [("u8", "=", 255)], [("u32", ">", 2**32 - 2**20)]):
        sel, count, total = scan.filter(preds)
        mask = _check_filter(scan, t, preds, sel, count, total, brows)
        assert mask.any() and not mask.all()
    with pytest.raises(b2a.B2AError, match="signed integer-like|only"):
        t2 = pa.table({"u64": pa.array(rng.integers(0, 2**63, 10).astype(np.uint64), pa.uint64())})
        _, sc2 = _device_scan(b2a, _stream(t2))
        sc2.filter([("u64", ">", 1)])


# ------------------------------------------------------------------------------------------------ malformed inputs
def _patch_buffer(stream: bytes, needle: bytes, replacement: bytes) -> bytes:
    i = stream.find(needle)
    assert i >= 0 and stream.find(needle, i + 1) < 0, "needle must occur exactly once"
    return stream[:i] + replacement + stream[i + len(needle):]


def test_string_view_outside_its_data_buffer_is_refused(b2a, cuda_device):
    strs = ["short", "a string view longer than twelve bytes", "another long string view value here"]
    arr = pa.array(strs * 50, pa.string_view())
    good = _stream(pa.table({"v": arr})).to_pybytes()
    b2a.scan_arrow_ipc([good])
    # view of row 1: {len 38, prefix "a st", buffer index 0, offset 0}: move its offset beyond the data buffer
    import struct
    view = struct.pack("<I4sII", 38, b"a st", 0, 0)
    i = good.find(view)
    assert i >= 0
    bad = good[:i] + struct.pack("<I4sII", 38, b"a st", 0, 1 << 30) + good[i + 16:]
    with pytest.raises(b2a.B2AError, match="string view"):
        b2a.scan_arrow_ipc([bad])
    bad_idx = good[:i] + struct.pack("<I4sII", 38, b"a st", 7, 0) + good[i + 16:]
    with pytest.raises(b2a.B2AError, match="string view"):
        b2a.scan_arrow_ipc([bad_idx])


def test_list_offsets_beyond_the_child_are_refused(b2a, cuda_device):
    n = 300
    lists = pa.array([[i, i + 1] for i in range(n)], pa.list_(pa.int32()))
    good = _stream(pa.table({"l": lists})).to_pybytes()
    b2a.scan_arrow_ipc([good])
    import struct
    last = struct.pack("<ii", 2 * (n - 1), 2 * n)  # the two last offsets
    i = good.find(last)
    assert i >= 0
    bad = good[:i] + struct.pack("<ii", 2 * (n - 1), 2 * n + 1000) + good[i + 8:]
    with pytest.raises(b2a.B2AError, match="child array"):
        b2a.scan_arrow_ipc([bad])


@pytest.mark.parametrize("rows", [64, 5])  # 8-rows-per-thread fast path and the row-wise tail
def test_dictionary_index_out_of_range_is_refused(b2a, cuda_device, rows):
    idx = np.arange(rows, dtype=np.int8) % 3
    d = pa.DictionaryArray.from_arrays(pa.array(idx), pa.array(["a", "b", "c"]))
    good = _stream(pa.table({"d": d})).to_pybytes()
    b2a.scan_arrow_ipc([good])
    raw = idx.tobytes()
    i = good.rfind(raw)
    assert i >= 0
    for poison in (3, 0x80):  # just past the dictionary; negative
        bad = bytearray(good)
        bad[i + rows - 2] = poison
        with pytest.raises(b2a.B2AError, match="dictionary index"):
            b2a.scan_arrow_ipc([bytes(bad)])


def test_two_devices_in_one_process(b2a, cuda_device):
    """the host pipeline's streams / staging rings / arena are per device (a thread may scan on GPU 0, then GPU 1)"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    batches = list(ipc.open_stream(DBGEN))
    for dev in (0, 1, 0):
        ds = b2a.read_arrow(DBGEN, device=dev)
        compare_scan_with_oracle(ds, batches)


@pytest.mark.parametrize("mode", ["mmap", "pread", "direct"])
def test_file_io_modes_give_the_same_vectors(b2a, cuda_device, mode, tmp_path, monkeypatch):
    """the staged host path reads a file-backed image by memcpy from the mapping, by pread() into the
    pinned ring, or by O_DIRECT block reads (falls back to pread where the file system refuses O_DIRECT, e.g. tmpfs)"""
    import shutil
    p = str(tmp_path / "li.arrows")  # tmp_path is on a real file system on the box: O_DIRECT gets its chance
    shutil.copy(DBGEN, p)
    monkeypatch.setenv("B2A_FILE_IO", mode)
    batches = list(ipc.open_stream(p))
    for host_alias in (True, False):
        ds = b2a.read_arrow(p, host_alias=host_alias)
        compare_scan_with_oracle(ds, batches)


def test_arrow_file_through_the_footer_index(b2a, cuda_device, tmp_path):
    """.arrow inputs are indexed from the footer (no sequential header walk) and decode to the same vectors"""
    for rel in ("data/fruit.arrow", "arrow-testing/1.0.0-littleendian/generated_primitive.arrow_file"):
        path = os.path.join(GOLDEN, rel)
        ds = b2a.read_arrow(path)
        assert ds.reader.used_footer
        f = ipc.open_file(path)
        compare_scan_with_oracle(ds, [f.get_batch(i) for i in range(f.num_record_batches)])
    # the reference-written lineitem as a FILE (footer) vs as a stream (walk): identical vectors
    t = ipc.open_stream(DBGEN).read_all()
    p = str(tmp_path / "li.arrow")
    with ipc.new_file(p, t.schema) as w:
        for b in t.to_batches(max_chunksize=2500):
            w.write_batch(b)
    ds = b2a.read_arrow(p)
    assert ds.reader.used_footer and ds.num_rows == t.num_rows
    f = ipc.open_file(p)
    compare_scan_with_oracle(ds, [f.get_batch(i) for i in range(f.num_record_batches)])


def test_device_alias_scan_feeds_the_consumers_from_the_image(b2a, cuda_device):
    """B2A_SCAN_DEVICE_ALIAS: DirectConversion columns stay in the device image (the reference's zero-copy rule,
    arrow_conversion.cpp:454-461, for consumers that live on the device) — no decode work for a Q6-shaped projection, the
    filter and the aggregate read the Arrow buffers in place and give the same selection vectors and the reference's answer"""
    import torch
    cols = ["l_quantity", "l_extendedprice", "l_discount", "l_shipdate"]
    r = b2a.ArrowIPCReader.open_file(DBGEN)
    r.set_projection(cols)
    alias = b2a.ArrowScan(r, device=0, device_alias=True)
    r2 = b2a.ArrowIPCReader.open_file(DBGEN)
    r2.set_projection(cols)
    plain = b2a.ArrowScan(r2, device=0)
    assert alias.n_tiles == 0 and plain.n_tiles > 0, "nothing to convert: no decode launch"
    assert alias.upload_bytes == plain.upload_bytes, "the aliased buffers still cross PCIe"
    f = next(x for x in r.output_columns() if x.name == "l_quantity")
    co, cp = alias.column(0, f.index), plain.column(0, f.index)
    none = (1 << 64) - 1
    assert co.data_off == none and co.image_off != none and cp.data_off != none and cp.image_off == none
    for s in (alias, plain):
        s.upload()
        s.decode()
        s.finish()
    assert alias.gpu_launches == 0
    n, total_s = ANSWERS["q6_count_sum"]
    sel_a, cnt_a, tot_a = alias.filter(Q6_PREDS)
    sel_p, cnt_p, tot_p = plain.filter(Q6_PREDS)
    assert int(tot_a.item()) == int(tot_p.item()) == int(n)
    assert torch.equal(cnt_a, cnt_p)
    keep = torch.arange(2048, device="cuda:0")[None, :] < cnt_a[:, None]
    assert torch.equal(sel_a.view(-1, 2048)[keep], sel_p.view(-1, 2048)[keep])
    revenue = alias.sum_product("l_extendedprice", "l_discount", sel_a, cnt_a)
    assert f"{revenue // 10000}.{revenue % 10000:04d}" == total_s
    assert alias.sum_product("l_extendedprice", "l_discount") == plain.sum_product("l_extendedprice", "l_discount")
    # the same through the pipelined upload (read_device), on a fresh scan
    r3 = b2a.ArrowIPCReader.open_file(DBGEN)
    r3.set_projection(cols)
    piped = b2a.ArrowScan(r3, device=0, device_alias=True)
    piped.d_out = piped.read_device(n_slices=2)
    _, _, tot = piped.filter(Q6_PREDS)
    assert int(tot.item()) == int(n)


def test_device_alias_with_nulls_strings_and_host_consumers(b2a, cuda_device):
    """columns with NULLs keep their masks in the arena, strings are decoded as usual, only the fixed-width data is
    aliased; a host consumer (read_host) is refused, and so is the combination with the host alias"""
    import torch
    rng = np.random.default_rng(3)
    n = 10_000
    k = pa.array(rng.integers(-1000, 1000, n), pa.int64(), mask=rng.random(n) < 0.2)
    d = pa.array(rng.integers(0, 20000, n).astype(np.int32), pa.date32())
    s = pa.array([None if i % 9 == 0 else "s%d" % i for i in range(n)], pa.string())
    buf = _stream(pa.table({"k": k, "d": d, "s": s}), chunk=3000)
    preds = [("k", ">=", 0), ("d", "<", 15000)]
    res = []
    for device_alias in (True, False):
        r = b2a.ArrowIPCReader.open_buffers([buf])
        scan = b2a.ArrowScan(r, device=0, device_alias=device_alias)
        scan.upload()
        scan.decode()
        scan.finish()
        sel, cnt, tot = scan.filter(preds)
        res.append((sel.view(-1, 2048).cpu(), cnt.cpu(), int(tot.item())))
        if device_alias:
            assert scan.gpu_launches >= 1  # the strings and the masks were decoded
            with pytest.raises(b2a.B2AError, match="DEVICE_ALIAS"):
                scan.read_host()
    (sa, ca, ta), (sp, cp, tp) = res
    kk = np.asarray(k.fill_null(-1).to_numpy(zero_copy_only=False))
    want = int(((kk >= 0) & np.asarray(k.is_valid()) & (np.asarray(d.cast(pa.int32())) < 15000)).sum())
    assert ta == tp == want
    assert torch.equal(ca, cp)
    keep = torch.arange(2048)[None, :] < ca[:, None]
    assert torch.equal(sa[keep], sp[keep])
    with pytest.raises(b2a.B2AError, match="exclude each other"):
        b2a.ArrowScan(b2a.ArrowIPCReader.open_buffers([buf]), device=0, device_alias=True, host_alias=True)


def test_filter_sum_product_fused_equals_the_two_kernels(b2a, cuda_device):
    """b2a_scan_filter_sum_product: count + sum(a*b) over the passing rows in ONE launch — the same numbers as
    b2a_scan_filter followed by b2a_scan_sum_product, the reference's Q6 answer on the dbgen fixture (plain and
    device-alias scans), and NULL factors / NULL predicate values skipped like DuckDB's sum and filter skip them"""
    import torch
    n, total_s = ANSWERS["q6_count_sum"]
    for device_alias in (False, True):
        r = b2a.ArrowIPCReader.open_file(DBGEN)
        r.set_projection(["l_quantity", "l_extendedprice", "l_discount", "l_shipdate"])
        scan = b2a.ArrowScan(r, device=0, device_alias=device_alias)
        scan.upload()
        scan.decode()
        scan.finish()
        launches = scan.gpu_launches
        total, res = scan.filter_sum_product(Q6_PREDS, "l_extendedprice", "l_discount")
        torch.cuda.synchronize()
        assert scan.gpu_launches == launches + 1
        lo, hi = [int(x) for x in res.cpu().tolist()]
        revenue = (hi << 64) | (lo & 0xFFFFFFFFFFFFFFFF)
        assert int(total.item()) == int(n) and f"{revenue // 10000}.{revenue % 10000:04d}" == total_s
        # a second call into the same outputs (what a captured graph replays)
        scan.filter_sum_product(Q6_PREDS, "l_extendedprice", "l_discount", out=(total, res))
        torch.cuda.synchronize()
        assert int(total.item()) == int(n) and [int(x) for x in res.cpu().tolist()] == [lo, hi]
    # NULLs in a predicate column and in both factors, negative products, several batches
    rng = np.random.default_rng(11)
    rows = 50_000
    k = pa.array(rng.integers(-100, 100, rows), pa.int64(), mask=rng.random(rows) < 0.1)
    a = pa.array(rng.integers(-10**12, 10**12, rows), pa.int64(), mask=rng.random(rows) < 0.2)
    b = pa.array(rng.integers(-10**6, 10**6, rows), pa.int64(), mask=rng.random(rows) < 0.2)
    buf = _stream(pa.table({"k": k, "a": a, "b": b}), chunk=7000)
    r = b2a.ArrowIPCReader.open_buffers([buf])
    scan = b2a.ArrowScan(r, device=0)
    scan.upload()
    scan.decode()
    scan.finish()
    preds = [("k", ">=", -20), ("k", "<", 35)]
    total, res = scan.filter_sum_product(preds, "a", "b")
    torch.cuda.synchronize()
    kk, aa, bb = (np.asarray(x.fill_null(0).to_numpy(zero_copy_only=False), dtype=object) for x in (k, a, b))
    passing = np.asarray(k.is_valid()) & (kk >= -20) & (kk < 35)
    both = passing & np.asarray(a.is_valid()) & np.asarray(b.is_valid())
    want = int(sum(int(x) * int(y) for x, y in zip(aa[both], bb[both])))
    lo, hi = [int(x) for x in res.cpu().tolist()]
    got = ((hi << 64) | (lo & 0xFFFFFFFFFFFFFFFF))
    got = got - (1 << 128) if got >= 1 << 127 else got
    assert int(total.item()) == int(passing.sum()) and got == want


def test_filter_every_width_and_constants_outside_the_type(b2a, cuda_device):
    """predicate columns of every width the filter takes (1 / 2 / 4 / 8 bytes, signed and unsigned, date32, timestamp) are
    compared in their own type: constants at, just inside and far outside the type's range, all six operators, NULLs, a
    ragged last vector — per-vector selection vectors against numpy; the fused filter + aggregate sees the same rows"""
    import torch
    from test_gpu_parity import _check_filter, _device_scan
    rng = np.random.default_rng(17)
    n = 2048 * 3 + 777
    def col(lo, hi, dt, ty, null_p=0.1):
        return pa.array(rng.integers(lo, hi, n).astype(dt), ty, mask=rng.random(n) < null_p)
    t = pa.table({
        "i8": col(-128, 128, np.int8, pa.int8()), "i16": col(-2**15, 2**15, np.int16, pa.int16()),
        "i32": col(-2**31, 2**31, np.int32, pa.int32()), "i64": col(-2**62, 2**62, np.int64, pa.int64(), 0.0),
        "u8": col(0, 256, np.uint8, pa.uint8()), "u16": col(0, 2**16, np.uint16, pa.uint16(), 0.0),
        "u32": col(0, 2**32, np.uint32, pa.uint32()),
        "d": pa.array(rng.integers(-30000, 30000, n).astype(np.int32), pa.date32()),
        "w": pa.array(rng.integers(1, 1000, n), pa.int64()),
    })
    r, scan = _device_scan(b2a, _stream(t, chunk=2048 * 2 + 500))
    brows = [r.batch_info(i).rows for i in range(r.num_batches)]
    edges = {"i8": (-128, 127), "i16": (-2**15, 2**15 - 1), "i32": (-2**31, 2**31 - 1), "u8": (0, 255),
             "u16": (0, 2**16 - 1), "u32": (0, 2**32 - 1), "d": (-2**31, 2**31 - 1), "i64": (-2**63, 2**63 - 1)}
    checked = 0
    for name, (lo, hi) in edges.items():
        consts = {lo, hi, lo + 1, hi - 1, 0, (lo + hi) // 3}
        if name != "i64":
            consts |= {lo - 1, hi + 1, lo - 10**9, hi + 10**12, -2**62, 2**62}
        for c in sorted(consts):
            for op in ("=", "!=", "<", "<=", ">", ">="):
                preds = [(name, op, c)]
                sel, count, total = scan.filter(preds)
                mask = _check_filter(scan, t, preds, sel, count, total, brows)
                tot2, res = scan.filter_sum_product(preds, "w", "w")
                torch.cuda.synchronize()
                ww = np.asarray(t.column("w").to_numpy(), dtype=object)
                assert int(tot2.item()) == int(mask.sum())
                lo64, hi64 = [int(x) for x in res.cpu().tolist()]
                assert ((hi64 << 64) | (lo64 & 0xFFFFFFFFFFFFFFFF)) == int(sum(int(x) * int(x) for x in ww[mask]))
                checked += 1
    # runs: several predicates on one column next to predicates on another, given out of order
    preds = [("i32", ">=", -2**30), ("u8", "<", 200), ("i32", "<", 2**30), ("d", ">=", -20000), ("u8", ">", 3), ("d", "!=", 0)]
    sel, count, total = scan.filter(preds)
    _check_filter(scan, t, preds, sel, count, total, brows)
    assert checked > 400
