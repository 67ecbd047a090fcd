"""IPC body compression (ZSTD, and LZ4_FRAME — SURVEY §8f N-4, which the reference cannot read): the device decoders
(src/kernels/zstd_decode.cuh, lz4_decode.cuh) pinned against the real libzstd / liblz4.
CPU: the same __host__ __device__ decoder compiled with g++ (tests/native/zstd_host.cpp) against frames produced by
pyarrow's libzstd.  GPU: compressed streams through the scan path vs the oracle fed by pyarrow's own decompression
(the reference's tests read arrow-testing 2.0.0-compression/generated_zstd and generated_uncompressible_zstd:
test/sql/arrow_testing.test:59-63)."""
import ctypes
import os
import subprocess

import numpy as np
import pyarrow as pa
import pyarrow.ipc as ipc
import pytest

from parity import GOLDEN, compare_scan_with_oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COMP = os.path.join(GOLDEN, "arrow-testing", "2.0.0-compression")


@pytest.fixture(scope="module")
def host_decoder():
    src = os.path.join(ROOT, "tests", "native", "zstd_host.cpp")
    so = os.path.join(ROOT, "tests", "native", "libzstd_host.so")
    hdr = os.path.join(ROOT, "duckdb_arrow_b200", "src", "kernels", "zstd_decode.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I",
                               os.path.join(ROOT, "duckdb_arrow_b200", "src"), src, "-o", so])
    lib = ctypes.CDLL(so)
    lib.b2a_test_zstd_decode.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t]

    def decode(comp: bytes, n: int):
        src_a = np.frombuffer(comp, np.uint8)
        dst = np.zeros(n + 8, np.uint8)
        rc = lib.b2a_test_zstd_decode(src_a.ctypes.data, len(comp), dst.ctypes.data, n)
        return rc, dst[:n].tobytes()
    return decode


@pytest.fixture(scope="module")
def host_lz4_decoder():
    src = os.path.join(ROOT, "tests", "native", "lz4_host.cpp")
    so = os.path.join(ROOT, "tests", "native", "liblz4_host.so")
    hdr = os.path.join(ROOT, "duckdb_arrow_b200", "src", "kernels", "lz4_decode.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I",
                               os.path.join(ROOT, "duckdb_arrow_b200", "src"), src, "-o", so])
    lib = ctypes.CDLL(so)
    lib.b2a_test_lz4_decode.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t]

    def decode(comp: bytes, n: int):
        src_a = np.frombuffer(comp, np.uint8) if comp else np.zeros(1, np.uint8)
        dst = np.full(n + 8, 0xA5, np.uint8)
        rc = lib.b2a_test_lz4_decode(src_a.ctypes.data, len(comp), dst.ctypes.data, n)
        assert (dst[n:] == 0xA5).all()  # never writes past the declared size
        return rc, dst[:n].tobytes()
    return decode


def _cases():
    from tools import lineitem
    rng = np.random.default_rng(1)
    yield "empty", b""
    yield "one", b"x"
    yield "zeros", bytes(1 << 20)
    yield "random", rng.integers(0, 256, 300_000, dtype=np.uint8).tobytes()
    yield "words", b" ".join(lineitem.WORDS[i] for i in rng.integers(0, len(lineitem.WORDS), 100_000))
    yield "seq", np.arange(300_000, dtype=np.int64).tobytes()
    yield "skewed", rng.choice(np.arange(256, dtype=np.uint8), 400_000, p=np.r_[0.5, np.full(255, 0.5 / 255)]).tobytes()
    for n in (2, 3, 15, 16, 17, 255, 256, 257, 65535, 65536, 131071, 131072, 131073, 262145):
        yield f"mixed{n}", (rng.integers(0, 3, n, dtype=np.uint8) * 37).tobytes()
    b = lineitem.make_batch(rng, 40_000, 0, lineitem._text_pool(rng))
    for i, c in enumerate(b.columns):
        for j, buf in enumerate(c.buffers()):
            if buf is not None and buf.size:
                yield f"lineitem{i}.{j}", buf.to_pybytes()


def test_host_build_of_device_decoder_vs_libzstd(host_decoder):
    n = 0
    for name, raw in _cases():
        for level in (1, 3, 19):
            comp = pa.Codec("zstd", compression_level=level).compress(raw, asbytes=True)
            rc, out = host_decoder(comp, len(raw))
            assert rc == 0 and out == raw, (name, level, rc)
            n += 1
    assert n > 100


def test_host_decoder_rejects_corruption(host_decoder):
    raw = b"hello hello hello hello world" * 100
    comp = bytearray(pa.Codec("zstd").compress(raw, asbytes=True))
    assert host_decoder(bytes(comp), len(raw) + 1)[0] != 0  # wrong content size
    assert host_decoder(bytes(comp[:-3]), len(raw))[0] != 0  # truncated
    bad = bytearray(comp)
    bad[0] ^= 0xFF
    assert host_decoder(bytes(bad), len(raw))[0] != 0  # magic


def test_host_build_of_device_lz4_decoder_vs_liblz4(host_lz4_decoder):
    n = 0
    for name, raw in _cases():
        comp = pa.Codec("lz4").compress(raw, asbytes=True)  # LZ4_FRAME, what Arrow IPC writes
        rc, out = host_lz4_decoder(comp, len(raw))
        assert rc == 0 and out == raw, (name, rc)
        n += 1
    # a frame larger than one 4 MiB block (linked blocks) and two concatenated frames
    rng = np.random.default_rng(5)
    big = (rng.integers(0, 4, 9_000_000, dtype=np.uint8) * 51).tobytes()
    rc, out = host_lz4_decoder(pa.Codec("lz4").compress(big, asbytes=True), len(big))
    assert rc == 0 and out == big
    a, b = b"left " * 1000, b"right" * 777
    two = pa.Codec("lz4").compress(a, asbytes=True) + pa.Codec("lz4").compress(b, asbytes=True)
    rc, out = host_lz4_decoder(two, len(a) + len(b))
    assert rc == 0 and out == a + b
    assert n > 30


def test_host_lz4_decoder_rejects_corruption(host_lz4_decoder):
    raw = b"hello hello hello hello world" * 100
    comp = bytearray(pa.Codec("lz4").compress(raw, asbytes=True))
    assert host_lz4_decoder(bytes(comp), len(raw) + 1)[0] != 0  # wrong size
    assert host_lz4_decoder(bytes(comp), len(raw) - 1)[0] != 0
    assert host_lz4_decoder(bytes(comp[:-5]), len(raw))[0] != 0  # truncated (no EndMark)
    bad = bytearray(comp)
    bad[0] ^= 0xFF
    assert host_lz4_decoder(bytes(bad), len(raw))[0] != 0  # magic
    bad = bytearray(comp)
    bad[4] |= 0x01  # dictionary id flag
    assert host_lz4_decoder(bytes(bad), len(raw))[0] != 0
    # every single-byte mutation ends in an error or an in-bounds result (the canary in the fixture checks the bounds)
    rng = np.random.default_rng(9)
    for _ in range(300):
        bad = bytearray(comp)
        bad[int(rng.integers(0, len(bad)))] = int(rng.integers(0, 256))
        host_lz4_decoder(bytes(bad), len(raw))


def test_golden_compressed_streams_plan_on_host(b2a):
    """host planning of the reference's compressed fixtures: jobs, scratch layout; no device needed"""
    for name in ("generated_zstd", "generated_uncompressible_zstd", "generated_lz4", "generated_uncompressible_lz4"):
        r = b2a.ArrowIPCReader.open_file(os.path.join(COMP, name + ".stream"))
        assert any(r.batch_info(b).compressed for b in range(r.num_batches))
        s = b2a.ArrowScan(r, device=-1)
        assert s.n_decompress_jobs > 0 and s.scratch_bytes > 0
        assert s.scratch_offset % 256 == 0


def _pa_batches(path_or_buf):
    if isinstance(path_or_buf, str):
        path_or_buf = pa.BufferReader(open(path_or_buf, "rb").read())
    return list(ipc.open_stream(path_or_buf))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["generated_zstd", "generated_uncompressible_zstd",
                                  "generated_lz4", "generated_uncompressible_lz4"])
def test_arrow_testing_compression_files(b2a, cuda_device, name):
    path = os.path.join(COMP, name + ".stream")
    ds = b2a.read_arrow(path)
    compare_scan_with_oracle(ds, _pa_batches(path))


def _compressed_lineitem(rows, codec="zstd", level=None, **kw):
    from tools import lineitem
    sink = pa.BufferOutputStream()
    opt = ipc.IpcWriteOptions(compression=pa.Codec(codec, compression_level=level) if level else codec)
    with ipc.new_stream(sink, lineitem.schema(kw.get("dictionary_strings", False)), options=opt) as w:
        for b in lineitem.iter_batches(rows, seed=21, batch_rows=kw.get("batch_rows", 30_000),
                                       dictionary_strings=kw.get("dictionary_strings", False)):
            w.write_batch(b)
    return sink.getvalue()


@pytest.mark.gpu
@pytest.mark.parametrize("level", [None, 1, 12])
def test_compressed_lineitem(b2a, cuda_device, level):
    buf = _compressed_lineitem(100_000, level=level)
    ds = b2a.scan_arrow_ipc([buf])
    assert ds.scan.n_decompress_jobs > 0
    compare_scan_with_oracle(ds, _pa_batches(pa.BufferReader(buf)))


@pytest.mark.gpu
def test_compressed_lineitem_dictionary_and_projection(b2a, cuda_device):
    buf = _compressed_lineitem(70_000, dictionary_strings=True)
    ds = b2a.scan_arrow_ipc([buf])
    compare_scan_with_oracle(ds, _pa_batches(pa.BufferReader(buf)))
    cols = ["l_comment", "l_shipdate", "l_returnflag"]
    dp = b2a.scan_arrow_ipc([buf], columns=cols)
    assert dp.scan.n_decompress_jobs < ds.scan.n_decompress_jobs  # only projected buffers are decompressed
    compare_scan_with_oracle(dp, _pa_batches(pa.BufferReader(buf)), columns=cols)


@pytest.mark.gpu
def test_compressed_nested_and_nulls(b2a, cuda_device):
    rng = np.random.default_rng(3)
    n = 20_000
    t = pa.table({
        "i": pa.array(rng.integers(0, 100, n), mask=rng.random(n) < 0.2),
        "s": pa.array([None if i % 7 == 0 else "v" * (i % 30) for i in range(n)]),
        "l": pa.array([[j for j in range(i % 5)] if i % 11 else None for i in range(n)], pa.list_(pa.int32())),
        "st": pa.array([{"a": i, "b": str(i)} if i % 13 else None for i in range(n)]),
        "b": pa.array(rng.random(n) < 0.5, mask=rng.random(n) < 0.1),
    })
    sink = pa.BufferOutputStream()
    with ipc.new_stream(sink, t.schema, options=ipc.IpcWriteOptions(compression="zstd")) as w:
        w.write_table(t, max_chunksize=6000)
    buf = sink.getvalue()
    ds = b2a.scan_arrow_ipc([buf])
    compare_scan_with_oracle(ds, _pa_batches(pa.BufferReader(buf)))


@pytest.mark.gpu
@pytest.mark.parametrize("kw", [{}, {"dictionary_strings": True}, {"batch_rows": 122_880}])
def test_lz4_compressed_lineitem(b2a, cuda_device, kw):
    buf = _compressed_lineitem(250_000, codec="lz4", **kw)
    ds = b2a.scan_arrow_ipc([buf])
    assert ds.scan.n_decompress_jobs > 0
    compare_scan_with_oracle(ds, _pa_batches(pa.BufferReader(buf)))
    cols = ["l_comment", "l_shipdate", "l_returnflag"]
    dp = b2a.scan_arrow_ipc([buf], columns=cols)
    assert dp.scan.n_decompress_jobs < ds.scan.n_decompress_jobs
    compare_scan_with_oracle(dp, _pa_batches(pa.BufferReader(buf)), columns=cols)


@pytest.mark.gpu
def test_lz4_long_runs_and_nested(b2a, cuda_device):
    """long matches / literal runs (the 16-byte copy path of lz4_decode.cuh), period-1..7 matches, nested + NULLs"""
    rng = np.random.default_rng(4)
    n = 300_000
    t = pa.table({
        "zeros": pa.array(np.zeros(n, np.int64)),
        "period3": pa.array(np.tile(np.array([1, 2, 3], np.int8), n // 3)),
        "noise": pa.array(rng.integers(-2**62, 2**62, n)),
        "steps": pa.array(np.repeat(rng.integers(0, 1000, n // 1000), 1000).astype(np.int32)),
        "s": pa.array(["x" * int(k) for k in rng.integers(0, 40, n)], mask=rng.random(n) < 0.1),
        "l": pa.array([[int(i), int(i) + 1] if i % 7 else None for i in range(n)], pa.list_(pa.int32())),
    })
    sink = pa.BufferOutputStream()
    with ipc.new_stream(sink, t.schema, options=ipc.IpcWriteOptions(compression="lz4")) as w:
        for b in t.to_batches(max_chunksize=100_000):
            w.write_batch(b)
    buf = sink.getvalue()
    ds = b2a.scan_arrow_ipc([buf])
    compare_scan_with_oracle(ds, _pa_batches(pa.BufferReader(buf)))


@pytest.mark.gpu
def test_corrupt_frames_reported(b2a, cuda_device):
    bad4 = bytearray(_compressed_lineitem(10_000, codec="lz4").to_pybytes())
    r4 = b2a.ArrowIPCReader.open_buffers([bytes(bad4)])
    b4 = r4.batch_info(0)
    for k in range(64):
        bad4[b4.body_offset + b4.body_length // 2 + k] ^= 0x5A
    try:
        b2a.scan_arrow_ipc([bytes(bad4)])  # LZ4 has no checksum here: a flipped literal may decode "successfully"
    except b2a.B2AError as e:
        assert "LZ4F_decompress" in str(e)
    good = bytearray(_compressed_lineitem(10_000).to_pybytes())
    r = b2a.ArrowIPCReader.open_buffers([bytes(good)])
    bi = r.batch_info(0)
    # flip bytes in the middle of the first compressed body
    for k in range(64):
        good[bi.body_offset + bi.body_length // 2 + k] ^= 0x5A
    with pytest.raises(b2a.B2AError, match="ZSTD_decompress"):
        b2a.scan_arrow_ipc([bytes(good)])
