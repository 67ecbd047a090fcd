"""GPU parity at BASELINE.json's FULL size: lineitem SF10 (59 986 052 rows, 489 record batches of 122 880 rows + a
ragged last one, 8.45 GB of IPC bodies) — configs 2, 3 and 5.

The CPU oracle needs minutes for 60 M rows, so it checks three sampled batches (first, middle, ragged last) and every
other row is checked on the device, exactly, by statements that do not depend on the size:
  * DirectConversion columns (arrow_conversion.cpp:454-461): every decoded vector is byte-identical to the Arrow buffer
    pyarrow's reader slices out of the same stream (an independent slicer);
  * SetVectorString (arrow_conversion.cpp:409-423) + string_t(const char*, uint32_t) (string_type.hpp:43-64): for every
    row, length == offsets[i+1] - offsets[i]; length <= 12 -> the inlined bytes are the characters, zero padded; else the
    prefix is the first four characters and the pointer is characters + offsets[i] — restated with torch gathers over
    the device image;
  * pushed-down filters + the Q6 aggregate == numpy over pyarrow's decoding of all rows;
  * COPY -> read round trip (test/sql/write_arrow_stream.test:17-25): the vectors written by the encode kernels, read
    back by pyarrow, equal the source batches column by column.
Source = bench.build_stream: the dbgen SF10 stream written through the reference's appender (tools/dbgen_lineitem.py)
where oracle/_ref is in the tree — then the COPY test also demands that every message BODY the encode kernels write is
byte-identical to the body the reference's ArrowAppender produced for the same rows — else tools/lineitem."""
import numpy as np
import pyarrow as pa
import pyarrow.ipc as ipc
import pytest

from parity import O, compare_column

pytestmark = pytest.mark.gpu

ROWS = 59_986_052
STRINGS = ["l_returnflag", "l_linestatus", "l_shipinstruct", "l_shipmode", "l_comment"]


@pytest.fixture(scope="module")
def sf10(b2a, cuda_device):
    import torch
    from bench import build_stream
    host, size = build_stream(ROWS, 11, pin=False)
    reader = b2a.ArrowIPCReader.open_buffers([(host.data_ptr(), size)])
    scan = b2a.ArrowScan(reader, device=0)
    d_image = torch.empty(scan.image_bytes + 256, dtype=torch.uint8, device="cuda:0")
    d_image[:scan.image_bytes].copy_(host[scan.image_lo: scan.image_lo + scan.image_bytes])
    scan.set_device_image(d_image)
    d_out = scan.decode()
    scan.finish()
    batches = list(ipc.open_stream(pa.py_buffer(host.numpy()[:size])))  # zero-copy views of `host`
    assert reader.num_rows == ROWS and len(batches) == reader.num_batches == 489
    yield {"host": host, "size": size, "reader": reader, "scan": scan, "d_image": d_image, "d_out": d_out,
           "batches": batches}
    scan.close()


def _image_off(s, buf: pa.Buffer) -> int:
    """offset of an Arrow buffer (a view of the host stream) inside the device image"""
    return buf.address - s["host"].data_ptr() - s["scan"].image_lo


def test_sf10_sampled_batches_bit_exact_vs_oracle(b2a, sf10):
    scan, reader = sf10["scan"], sf10["reader"]
    arena = sf10["d_out"].cpu().numpy()
    image = sf10["host"].numpy()[: sf10["size"]]
    # string_t pointers are device addresses of the image: origin = where image byte 0 would be on the device
    ds = b2a.DecodedScan(scan, arena, image, sf10["d_image"].data_ptr() - scan.image_lo)
    assert ds.num_rows == ROWS
    for bi in (0, 244, 488):
        batch = sf10["batches"][bi]
        assert batch.num_rows == (122_880 if bi < 488 else ROWS - 488 * 122_880)
        for f in reader.output_columns():
            arr = batch.column([c.index for c in reader.columns].index(f.index))
            compare_column(ds.column(bi, f.index), O.scan_column(arr), arr, f"batch{bi}.{f.name}")


def test_sf10_every_fixed_width_vector_equals_its_arrow_buffer(b2a, sf10):
    import torch
    scan, reader, d_image, d_out = sf10["scan"], sf10["reader"], sf10["d_image"], sf10["d_out"]
    ok = []
    for bi, batch in enumerate(sf10["batches"]):
        for ci, f in enumerate(reader.columns):
            if f.name in STRINGS:
                continue
            arr = batch.column(ci)
            assert arr.null_count == 0
            co = scan.column(bi, f.index)
            n = batch.num_rows * f.out_width
            src = _image_off(sf10, arr.buffers()[1])
            ok.append(torch.equal(d_out[co.data_off: co.data_off + n], d_image[src: src + n]))
    assert len(ok) == 489 * 11 and all(ok)


def test_sf10_every_string_t_follows_set_vector_string(b2a, sf10):
    import torch
    scan, reader, d_image, d_out = sf10["scan"], sf10["reader"], sf10["d_image"], sf10["d_out"]
    base = d_image.data_ptr()
    ar4 = torch.arange(4, device="cuda:0")
    ar12 = torch.arange(12, device="cuda:0")
    bad = torch.zeros(1, dtype=torch.int64, device="cuda:0")
    n_long = n_rows = 0
    for bi, batch in enumerate(sf10["batches"]):
        n = batch.num_rows
        for ci, f in enumerate(reader.columns):
            if f.name not in STRINGS:
                continue
            _, b_off, b_chr = batch.column(ci).buffers()
            o = _image_off(sf10, b_off)
            c0 = _image_off(sf10, b_chr)
            offs = d_image[o: o + 4 * (n + 1)].view(torch.int32).long()
            lens = offs[1:] - offs[:-1]
            co = scan.column(bi, f.index)
            st = d_out[co.data_off: co.data_off + 16 * n]
            words = st.view(torch.int32).reshape(n, 4)
            rec = st.reshape(n, 16)
            bad += (words[:, 0].long() != lens).sum()
            long_ = lens > 12
            # long strings: 4-byte prefix + pointer to the characters inside the (device) image
            lo = offs[:-1][long_]
            ptr = st.view(torch.int64).reshape(n, 2)[:, 1][long_]
            bad += (ptr != base + c0 + lo).sum()
            bad += (d_image[(c0 + lo)[:, None] + ar4] != rec[long_][:, 4:8]).sum()
            # inlined strings: the characters, zero padded to 12
            so, sl = offs[:-1][~long_], lens[~long_]
            keep = ar12[None, :] < sl[:, None]
            chars = d_image[(c0 + so)[:, None] + torch.where(keep, ar12[None, :], 0)]
            bad += (torch.where(keep, chars, torch.zeros_like(chars)) != rec[~long_][:, 4:16]).sum()
            n_long += int(long_.sum().item()) if bi == 0 else 0
            n_rows += n
    assert n_rows == 5 * ROWS
    assert int(bad.item()) == 0
    assert n_long > 100_000  # l_comment / half of l_shipinstruct take the pointer form


def test_sf10_filters_and_q6_equal_numpy(b2a, sf10):
    scan = sf10["scan"]
    preds = [("l_shipdate", ">=", 8766), ("l_shipdate", "<", 9131), ("l_discount", ">=", 5), ("l_discount", "<=", 7),
             ("l_quantity", "<", 2400)]
    sel, count, total = scan.filter(preds)
    revenue = scan.sum_product("l_extendedprice", "l_discount", sel, count)
    names = sf10["batches"][0].schema.names
    exp_total, exp_rev, exp_counts = 0, 0, []
    for batch in sf10["batches"]:
        col = lambda nm, dt: np.frombuffer(batch.column(names.index(nm)).buffers()[1], dtype=dt)[: batch.num_rows]  # noqa: E731
        ship, disc = col("l_shipdate", np.int32), col("l_discount", np.int64)
        qty, price = col("l_quantity", np.int64), col("l_extendedprice", np.int64)
        m = (ship >= 8766) & (ship < 9131) & (disc >= 5) & (disc <= 7) & (qty < 2400)
        exp_total += int(m.sum())
        exp_rev += int((price[m] * disc[m]).sum())  # < 2^63 per batch: prices < 2^24, discounts <= 10
        pad = np.zeros(-batch.num_rows % 2048, bool)
        exp_counts.append(np.concatenate([m, pad]).reshape(-1, 2048).sum(1))
    assert int(total.item()) == exp_total and exp_total > 0
    assert np.array_equal(count.cpu().numpy().view(np.uint32), np.concatenate(exp_counts).astype(np.uint32))
    assert revenue == exp_rev
    # idempotence: the same launch again gives the same selection
    sel2, count2, total2 = scan.filter(preds)
    import torch
    assert torch.equal(count, count2) and torch.equal(total, total2)
    live = torch.arange(2048, device="cuda:0")[None, :] < count.long()[:, None]
    assert torch.equal(sel.reshape(-1, 2048)[live], sel2.reshape(-1, 2048)[live])


def test_sf10_copy_round_trip(b2a, sf10):
    """COPY lineitem TO (FORMAT ARROWS) of the decoded vectors, read back by pyarrow == the source batches"""
    import torch
    scan, reader, d_out = sf10["scan"], sf10["reader"], sf10["d_out"]
    cols = []
    for f in reader.columns:
        parts = []
        for b in range(reader.num_batches):
            co = scan.column(b, f.index)
            parts.append(d_out[co.data_off: co.data_off + co.rows * f.out_width])
        cols.append(b2a.DeviceVector(f.type, ROWS, torch.cat(parts), None, name=f.name,
                                     decimal=(f.decimal_width, f.decimal_scale)))
    w = b2a.ArrowWriter(cols, ROWS, 122_880)
    stream = w.write()
    assert w.n_batches == 489
    got = ipc.open_stream(pa.py_buffer(stream))
    assert got.schema.names == sf10["batches"][0].schema.names
    n = 0
    for a, b in zip(got, sf10["batches"]):
        assert a.num_rows == b.num_rows
        for i in range(a.num_columns):
            assert a.column(i).equals(b.column(i)), (n, a.schema.names[i])
        n += 1
    assert n == 489
    # where the source stream was written by the REFERENCE's appender (dbgen data): every record-batch body of the GPU
    # writer must be byte-identical to the reference's (validity 0xFF fill, values, offsets, characters, 8-byte padding)
    from tools import dbgen_lineitem as D
    if D.available():
        src = sf10["host"].numpy()[: sf10["size"]]
        pos, k = 0, 0
        import struct
        while pos + 8 <= len(src):
            cont, meta = struct.unpack_from("<Ii", src, pos)
            if cont != 0xFFFFFFFF or meta == 0:
                break
            msg = ipc.read_message(pa.py_buffer(src[pos:]))
            body = msg.body.size
            if msg.type == "record batch":
                bi = w.batch(k)
                mine = stream[bi.body_offset: bi.body_offset + bi.body_length]
                ref = src[pos + 8 + meta: pos + 8 + meta + body]
                assert bi.body_length == body and np.array_equal(mine, ref), f"body of batch {k} differs from the reference's"
                k += 1
            pos += 8 + meta + body
        assert k == 489
    w.close()
