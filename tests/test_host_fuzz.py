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
