"""Synthetic TPC-H lineitem as an Arrow IPC stream, shaped like `COPY lineitem TO 'x.arrows' (FORMAT ARROWS)`:
DuckDB dbgen's schema (duckdb/extension/tpch/dbgen/dbgen.cpp:409-426 — 4 x BIGINT, 4 x DECIMAL(15,2) stored as
int64 / Arrow `d:15,2,64`, 2 x VARCHAR(1), 3 x DATE, 3 x VARCHAR), TPC-H value distributions, batches of
122 880 rows (src/writer/write_arrow_stream.cpp:33).  DuckDB's dbgen itself is not available without building
the reference, so values are drawn with a seeded numpy generator ("data": "synthetic").  Test/bench tooling."""
from __future__ import annotations

import numpy as np
import pyarrow as pa

ROWS_PER_SF = {0.01: 60_175, 0.1: 600_572, 1: 6_001_215, 10: 59_986_052, 100: 600_037_902}
BATCH_ROWS = 122_880
EPOCH_1992 = 8035  # 1992-01-01 in days since 1970-01-01

SHIPINSTRUCT = [b"DELIVER IN PERSON", b"COLLECT COD", b"NONE", b"TAKE BACK RETURN"]
SHIPMODE = [b"REG AIR", b"AIR", b"RAIL", b"SHIP", b"TRUCK", b"MAIL", b"FOB"]
WORDS = (b"furiously quickly carefully blithely slyly regular express special pending final ironic even bold silent "
         b"unusual deposits requests accounts packages foxes ideas theodolites pinto beans instructions dependencies "
         b"excuses platelets asymptotes courts dolphins multipliers sauternes warthogs frets dinos attainments "
         b"sleep wake nag haggle cajole detect integrate boost use are above according to about after against along "
         b"alongside of among around at atop before behind beneath beside besides between beyond by despite during "
         b"except for from in place of inside instead of into near of on outside over past since through throughout "
         b"to toward under until up upon without with within ").split()


def schema(dictionary_strings: bool = False) -> pa.Schema:
    dec = pa.decimal64(15, 2)
    s = pa.string()
    d8 = pa.dictionary(pa.int8(), pa.string()) if dictionary_strings else s
    d32 = pa.dictionary(pa.int32(), pa.string()) if dictionary_strings else s
    return pa.schema([("l_orderkey", pa.int64()), ("l_partkey", pa.int64()), ("l_suppkey", pa.int64()),
                      ("l_linenumber", pa.int64()), ("l_quantity", dec), ("l_extendedprice", dec),
                      ("l_discount", dec), ("l_tax", dec), ("l_returnflag", d8), ("l_linestatus", d8),
                      ("l_shipdate", pa.date32()), ("l_commitdate", pa.date32()), ("l_receiptdate", pa.date32()),
                      ("l_shipinstruct", d32), ("l_shipmode", d32), ("l_comment", s)])


def _text_pool(rng, size=1 << 20) -> np.ndarray:
    idx = rng.integers(0, len(WORDS), size // 5)
    return np.frombuffer(b" ".join(WORDS[i] for i in idx)[:size], dtype=np.uint8)


def _fixed(t, a: np.ndarray) -> pa.Array:
    return pa.Array.from_buffers(t, len(a), [None, pa.py_buffer(a)])


def _strings_from_choices(choices, idx: np.ndarray) -> pa.Array:
    lens = np.array([len(c) for c in choices], dtype=np.int64)
    starts = np.concatenate([[0], np.cumsum(lens)[:-1]])
    pool = np.frombuffer(b"".join(choices), dtype=np.uint8)
    l = lens[idx]
    offsets = np.zeros(len(idx) + 1, dtype=np.int32)
    np.cumsum(l, out=offsets[1:])
    total = int(offsets[-1])
    pos = np.arange(total, dtype=np.int64) - np.repeat(offsets[:-1].astype(np.int64), l) + np.repeat(starts[idx], l)
    chars = pool[pos]
    return pa.Array.from_buffers(pa.string(), len(idx), [None, pa.py_buffer(offsets), pa.py_buffer(chars)])


def _dict_from_choices(choices, idx: np.ndarray, index_type) -> pa.Array:
    return pa.DictionaryArray.from_arrays(pa.array(idx.astype(index_type.to_pandas_dtype()), type=index_type),
                                          pa.array([c.decode() for c in choices], type=pa.string()))


def make_batch(rng, n: int, first_order: int, pool: np.ndarray, dictionary_strings=False) -> pa.RecordBatch:
    sch = schema(dictionary_strings)
    dec = pa.decimal64(15, 2)
    # ~4 lines per order, line numbers 1..7
    line = rng.integers(1, 8, n, dtype=np.int64)
    orderkey = first_order + np.cumsum(line == 1, dtype=np.int64)
    partkey = rng.integers(1, 2_000_001, n, dtype=np.int64)
    suppkey = rng.integers(1, 100_001, n, dtype=np.int64)
    quantity = rng.integers(1, 51, n, dtype=np.int64) * 100
    price = (quantity // 100) * rng.integers(90_000, 200_001, n, dtype=np.int64)
    discount = rng.integers(0, 11, n, dtype=np.int64)
    tax = rng.integers(0, 9, n, dtype=np.int64)
    shipdate = (EPOCH_1992 + rng.integers(1, 2526, n)).astype(np.int32)
    commitdate = (shipdate + rng.integers(-60, 61, n)).astype(np.int32)
    receiptdate = (shipdate + rng.integers(1, 31, n)).astype(np.int32)
    rf = rng.integers(0, 3, n)
    ls = rng.integers(0, 2, n)
    si = rng.integers(0, 4, n)
    sm = rng.integers(0, 7, n)
    # comments: dbgen copies a substring of a pre-generated text pool; here consecutive substrings of one window
    clen = rng.integers(10, 44, n)
    coff = np.zeros(n + 1, dtype=np.int32)
    np.cumsum(clen, out=coff[1:])
    total = int(coff[-1])
    start = int(rng.integers(0, len(pool) - total - 1)) if total < len(pool) - 1 else 0
    reps = (total + len(pool) - 1) // len(pool) + 1
    cchars = (np.tile(pool, reps) if reps > 1 else pool)[start:start + total]
    comment = pa.Array.from_buffers(pa.string(), n, [None, pa.py_buffer(coff), pa.py_buffer(np.ascontiguousarray(cchars))])
    if dictionary_strings:
        s_rf = _dict_from_choices([b"A", b"N", b"R"], rf, pa.int8())
        s_ls = _dict_from_choices([b"F", b"O"], ls, pa.int8())
        s_si = _dict_from_choices(SHIPINSTRUCT, si, pa.int32())
        s_sm = _dict_from_choices(SHIPMODE, sm, pa.int32())
    else:
        s_rf = _strings_from_choices([b"A", b"N", b"R"], rf)
        s_ls = _strings_from_choices([b"F", b"O"], ls)
        s_si = _strings_from_choices(SHIPINSTRUCT, si)
        s_sm = _strings_from_choices(SHIPMODE, sm)
    cols = [_fixed(pa.int64(), orderkey), _fixed(pa.int64(), partkey), _fixed(pa.int64(), suppkey),
            _fixed(pa.int64(), line), _fixed(dec, quantity), _fixed(dec, price), _fixed(dec, discount),
            _fixed(dec, tax), s_rf, s_ls, _fixed(pa.date32(), shipdate), _fixed(pa.date32(), commitdate),
            _fixed(pa.date32(), receiptdate), s_si, s_sm, comment]
    return pa.RecordBatch.from_arrays(cols, schema=sch)


def iter_batches(rows: int, seed: int = 42, batch_rows: int = BATCH_ROWS, dictionary_strings=False, unique: int = 0):
    """Yield record batches totalling `rows` rows.  unique > 0: only that many distinct batches are generated and
    then cycled (keeps SF10+ generation to seconds; every batch still occupies its own bytes in the stream)."""
    rng = np.random.default_rng(seed)
    pool = _text_pool(rng)
    done, order, cache = 0, 0, []
    while done < rows:
        n = min(batch_rows, rows - done)
        if unique and len(cache) >= unique and n == batch_rows:
            b = cache[(done // batch_rows) % unique]
        else:
            b = make_batch(rng, n, order, pool, dictionary_strings)
            if unique and n == batch_rows:
                cache.append(b)
        order += n // 4
        done += n
        yield b


def write_stream(sink, rows: int, seed: int = 42, batch_rows: int = BATCH_ROWS, dictionary_strings=False,
                 unique: int = 0) -> int:
    """Write an IPC stream (pyarrow encoder, 8-byte buffer alignment like the reference writer) to `sink`."""
    n = 0
    with pa.ipc.new_stream(sink, schema(dictionary_strings)) as w:
        for b in iter_batches(rows, seed, batch_rows, dictionary_strings, unique):
            w.write_batch(b)
            n += 1
    return n


def stream_bytes(rows: int, **kw) -> pa.Buffer:
    sink = pa.BufferOutputStream()
    write_stream(sink, rows, **kw)
    return sink.getvalue()
