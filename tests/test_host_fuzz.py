"""Host slicer / planner robustness: truncated and corrupted IPC streams must end in a B2AError (IOException /
InvalidInputException / NotImplementedException like the reference's reader) or in a valid plan whose every buffer lies
inside the stream image — never in a crash, and never in a plan the kernels would read out of bounds with."""
import os

import numpy as np

from parity import GOLDEN

FILES = ["data/test.arrows", "data/fruit.arrow", "arrow-testing/1.0.0-littleendian/generated_nested.stream",
         "arrow-testing/1.0.0-littleendian/generated_dictionary.stream",
         "arrow-testing/2.0.0-compression/generated_zstd.stream"]


def test_mutated_streams_never_crash(b2a):
    rng = np.random.default_rng(7)
    planned = rejected = 0
    for f in FILES:
        data = open(os.path.join(GOLDEN, f), "rb").read()
        for it in range(90):
            b = bytearray(data)
            if it % 3 == 0:
                b = b[:int(rng.integers(0, len(b)))]
            elif it % 3 == 1:
                for _ in range(int(rng.integers(1, 6))):
                    lim = min(len(b), 4096) if rng.random() < 0.7 else len(b)
                    b[int(rng.integers(0, lim))] = int(rng.integers(0, 256))
            else:
                p = int(rng.integers(0, max(1, len(b) - 8)))
                b[p:p + 8] = rng.integers(0, 256, 8, dtype=np.uint8).tobytes()
            try:
                r = b2a.ArrowIPCReader.open_buffers([bytes(b)])
                s = b2a.ArrowScan(r, device=-1)
            except b2a.B2AError:
                rejected += 1
                continue
            planned += 1
            assert s.image_lo + s.image_bytes <= len(b) + 16  # upload ranges are rounded to 16 bytes
            for i in range(r.num_batches):
                bi = r.batch_info(i)
                assert bi.body_offset + bi.body_length <= len(b)
    assert planned > 0 and rejected > 0


def test_mutated_streams_of_the_rarer_layouts_never_crash(b2a):
    """same contract for sparse unions, list views, run-end encoded and extension-typed columns (streams built here)"""
    import pyarrow as pa
    import pyarrow.ipc as ipc
    n = 300
    ids = pa.array(np.arange(n, dtype=np.int8) % 2)
    union = pa.UnionArray.from_sparse(ids, [pa.array(range(n), pa.int64()), pa.array([str(i) * 3 for i in range(n)])],
                                      ["i", "s"])
    lens = np.arange(n, dtype=np.int32) % 4
    offs = np.concatenate([[0], np.cumsum(lens)])[:-1].astype(np.int32)
    lv = pa.ListViewArray.from_arrays(pa.array(offs), pa.array(lens), pa.array(range(int(lens.sum())), pa.int32()))
    ree = pa.RunEndEncodedArray.from_arrays(pa.array([100, 200, n], pa.int32()), pa.array(["a", None, "ccc"]))
    b8 = pa.ExtensionArray.from_storage(pa.bool8(), pa.array((np.arange(n) % 3).astype(np.int8)))
    js = pa.ExtensionArray.from_storage(pa.json_(pa.string()), pa.array(['{"k":%d}' % i for i in range(n)]))
    t = pa.table([union, lv, ree, b8, js], names=["u", "lv", "ree", "b8", "js"])
    sink = pa.BufferOutputStream()
    with ipc.new_stream(sink, t.schema) as w:
        w.write_table(t, max_chunksize=128)
    data = sink.getvalue().to_pybytes()
    rng = np.random.default_rng(21)
    planned = rejected = 0
    for it in range(400):
        b = bytearray(data)
        if it % 4 == 0:
            b = b[:int(rng.integers(0, len(b)))]
        else:
            for _ in range(int(rng.integers(1, 5))):
                b[int(rng.integers(0, len(b)))] = int(rng.integers(0, 256))
        try:
            r = b2a.ArrowIPCReader.open_buffers([bytes(b)])
            s = b2a.ArrowScan(r, device=-1)
        except b2a.B2AError:
            rejected += 1
            continue
        planned += 1
        assert s.image_lo + s.image_bytes <= len(b) + 16
    assert planned > 0 and rejected > 0


def test_negative_variadic_buffer_counts_are_rejected(b2a):
    """RecordBatch.variadicBufferCounts entries are buffer counts of the message: a negative one (or a pair like {-1, +1}
    that sums to what the schema needs) must end in B2AError at parse time — it used to wrap the planner's bounds check."""
    import struct

    import pyarrow as pa
    import pyarrow.ipc as ipc
    long_ = ["a string view longer than twelve bytes %d" % i for i in range(50)]
    t = pa.table({"a": pa.array(long_, pa.string_view()), "b": pa.array(long_[::-1], pa.string_view())})
    sink = pa.BufferOutputStream()
    with ipc.new_stream(sink, t.schema) as w:
        w.write_table(t)
    data = sink.getvalue().to_pybytes()
    r = b2a.ArrowIPCReader.open_buffers([data])
    b2a.ArrowScan(r, device=-1)
    # the vector [1, 1] of int64 counts: u32 length 2, then two int64 ones (8-byte aligned after the length)
    pos = [i for i in range(0, len(data) - 24, 4) if data[i:i + 4] == struct.pack("<I", 2)
           and data[i + 4:i + 8] in (b"\0\0\0\0",) and data[i + 8:i + 24] == struct.pack("<qq", 1, 1)]
    pos += [i for i in range(0, len(data) - 20, 4) if data[i:i + 4] == struct.pack("<I", 2)
            and data[i + 4:i + 20] == struct.pack("<qq", 1, 1)]
    assert pos, "variadicBufferCounts vector not found in the RecordBatch message"
    hit = 0
    for i in pos:
        start = i + 8 if data[i + 8:i + 24] == struct.pack("<qq", 1, 1) else i + 4
        for counts in ((-1, 3), (3, -1), (-1, 1), (1 << 40, 1), (-(1 << 62), 1)):
            b = bytearray(data)
            b[start:start + 16] = struct.pack("<qq", *counts)
            try:
                rr = b2a.ArrowIPCReader.open_buffers([bytes(b)])
                b2a.ArrowScan(rr, device=-1)
            except b2a.B2AError as e:
                hit += 1
                assert "variadicBufferCounts" in str(e) or "buffers" in str(e)
                continue
            raise AssertionError(f"counts {counts} were accepted")
    assert hit >= 5
