"""CPU oracle of the WRITE direction (numpy).  TEST INFRASTRUCTURE ONLY — never imported by the product.

Restates DuckDB's ArrowAppender, one function per appender, over host vectors (tests/vectors.HostVector):
  AppendValidity            duckdb/src/include/duckdb/common/arrow/appender/append_data.hpp:127-149
  ArrowScalarBaseData       .../appender/scalar_data.hpp:90-119 (+ ArrowIntervalConverter :34-52, int16->int32 decimals)
  ArrowBoolData             duckdb/src/common/arrow/appender/bool_data.cpp:12-39
  ArrowVarcharData          .../appender/varchar_data.hpp:53-111
  ArrowListData             .../appender/list_data.hpp:26-98 (child_sel gather)
  ArrowStructData           duckdb/src/common/arrow/appender/struct_data.cpp:17-30
  ArrowFixedSizeListData    duckdb/src/common/arrow/appender/fixed_size_list_data.cpp:16-28
  ArrowMapData              .../appender/map_data.hpp:38-64
  ArrowTimeTzConverter      .../appender/scalar_data.hpp:54-68 (TIME_TZ -> time64[us], the offset is dropped)
  ArrowUUIDConverter        .../appender/varchar_data.hpp:33-43 + UUID::ToString duckdb/src/common/types/uuid.cpp:70-101
  ArrowUnionData            duckdb/src/common/arrow/appender/union_data.cpp:24-52 (flat members)
  ArrowEnumData             .../appender/enum_data.hpp:24-87: the codes copied as they are (NULL slots included); the
                            dictionary = serialize_dictionaries below
  arrow_lossless_conversion arrow_appender.cpp:144-191 + arrow_type_extension.cpp:345-390: BOOLEAN as one byte per row
                            (ArrowBool8::DuckToArrow), UUID through ArrowUUIDBlobConverter (scalar_data.hpp:70-87),
                            TIME_TZ / HUGEINT / UHUGEINT bits as they are, BIT / VARINT like BLOB, JSON like VARCHAR
and the body packing of the nanoarrow encoder the reference calls (ArrowIpcEncoderEncodeSimpleRecordBatch,
src/writer/column_data_collection_serializer.cpp:86-92): buffers in pre-order, each padded with zeros to 8 bytes.
nanoarrow is not in the reference tree (CMakeLists.txt:7-13), so the exact FILE bytes of the reference writer are
"parity unpinned"; what is pinned: (1) the Arrow buffers against pyarrow's own buffers for the same logical data
(tests/test_oracle_write.py), (2) the produced stream decodes with pyarrow to the source table.

The row_group_size split follows Serialize(ColumnDataCollection&) (:97-115): every batch is appended from one flat
DataChunk holding exactly the rows of that batch (so each batch starts with offsets 0 and fresh validity)."""
from __future__ import annotations

import numpy as np


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * ((-len(b)) % 8)


def _valid(hv, idx: np.ndarray) -> np.ndarray:
    if hv.validity is None:
        return np.ones(len(idx), bool)
    return hv.validity[idx]


def _validity_buffer(valid: np.ndarray) -> tuple[bytes, int]:
    """ResizeValidity(0xFF) + UnsetBit per NULL: bits past the last row of the last byte stay 1."""
    n = len(valid)
    bits = np.ones(((n + 7) // 8) * 8, bool)
    bits[:n] = valid
    return np.packbits(bits, bitorder="little").tobytes(), int(n - valid.sum())


def append(hv, idx: np.ndarray, nodes: list, buffers: list, large: bool = False, view: bool = False,
           sview: bool = False, lossless: bool = False):
    """Append rows `idx` (input row indices: the UnifiedVectorFormat selection) of vector hv; pre-order output.
    nodes: [(length, null_count)], buffers: [bytes] (unpadded).
    large: ClientProperties.arrow_offset_size == LARGE (SET arrow_large_buffer_size): int64 offsets for VARCHAR / BLOB /
    BIT / UUID / LIST (arrow_appender.cpp:184-188,237-251,288-292); MAP stays int32 (:296-299).
    view: ClientProperties.arrow_use_list_view (SET arrow_output_list_view): LIST -> ArrowListViewData, an offsets and a
    sizes buffer of n entries each (appender/list_view_data.hpp:59-103); MAP is not affected.
    sview: ClientProperties.produce_arrow_string_view: VARCHAR -> ArrowVarcharToStringViewData (varchar_data.hpp:129-202):
    16-byte views {length, 12 bytes zero padded} / {length, first 4 bytes, buffer index 0, int32 offset} (NULL: zeros) and
    one data buffer with the strings longer than 12 bytes; the fourth C-data buffer (data-buffer sizes) is not part of the
    IPC body (it becomes RecordBatch.variadicBufferCounts).
    lossless: ClientProperties.arrow_lossless_conversion (see the module header); True for a top-level column, "nested"
    below one: the bool8 route exists for ROOT columns only (ArrowAppender::Append converts root_data[i] through
    duckdb_to_arrow, arrow_appender.cpp:40-48; a nested BOOLEAN keeps ArrowBoolData, :147-149)."""
    below = "nested" if lossless else False
    off_t = np.int64 if large else np.int32
    n = len(idx)
    t = hv.type
    if t == "SQLNULL":  # ArrowNullData: no buffers, null_count = length
        nodes.append((n, n))
        return
    if t == "UNION":
        # per row: Value v = input.GetValue(row); tag = v.IsNull() ? 0 : tag(v); member[tag] gets the boxed value, every
        # other member NULL; one buffer (type ids), null_count 0, no validity buffer (Finalize: n_buffers = 1).
        # hv.data = the tag child of DuckDB's UNION vector (UTINYINT per row), hv.validity = the union's own mask.
        valid_full = np.ones(hv.rows, bool) if hv.validity is None else hv.validity
        tags_full = hv.data.view(np.uint8)
        nodes.append((n, 0))
        buffers.append(np.where(valid_full[idx], tags_full[idx], 0).astype(np.uint8).tobytes())
        for m, c in enumerate(hv.children):
            own = np.ones(c.rows, bool) if c.validity is None else c.validity
            import copy
            member = copy.copy(c)
            member.validity = valid_full & (tags_full == m) & own
            append(member, idx, nodes, buffers, large, view, sview, below)
        return
    valid = _valid(hv, idx)
    vbuf, nulls = _validity_buffer(valid)
    if getattr(hv, "_no_validity_buffer", False):
        vbuf, nulls = b"", 0
    nodes.append((n, nulls))
    buffers.append(vbuf)
    if t == "BOOLEAN" and lossless is True:  # arrow.bool8: the byte of every row, NULL slots included
        buffers.append(hv.data[idx].astype(np.uint8).tobytes())
    elif t == "UUID" and lossless:  # most significant byte first, the stored sign bit flipped back; NULLs skipped (0)
        raw = hv.data.view(np.uint64).reshape(-1, 2)[idx]
        out = np.zeros((n, 2), ">u8")
        out[:, 0] = raw[:, 1] ^ np.uint64(1 << 63)
        out[:, 1] = raw[:, 0]
        out[~valid] = 0
        buffers.append(out.tobytes())
    elif t == "TIME_TZ" and lossless:
        buffers.append(hv.data.reshape(-1, 8)[idx].tobytes() if n else b"")
    elif t == "BOOLEAN":
        data = hv.data[idx] != 0
        bits = np.ones(((n + 7) // 8) * 8, bool)
        bits[:n] = ~valid | data  # only a VALID false clears the bit
        buffers.append(np.packbits(bits, bitorder="little").tobytes())
    elif t == "UUID":  # 36 characters per valid row, NULL rows repeat the previous offset
        offs = np.zeros(n + 1, np.int64)
        np.cumsum(np.where(valid, 36, 0), out=offs[1:])
        raw = hv.data.view(np.uint64).reshape(-1, 2)[idx]
        chars = bytearray()
        for (lower, upper), ok in zip(raw.tolist(), valid.tolist()):
            if ok:
                h = "%016x%016x" % (upper ^ (1 << 63), lower)  # the sign bit of `upper` is flipped back
                chars += f"{h[:8]}-{h[8:12]}-{h[12:16]}-{h[16:20]}-{h[20:]}".encode()
        buffers.append(offs.astype(off_t).tobytes())
        buffers.append(bytes(chars))
    elif t == "TIME_TZ":  # dtime_tz_t: micros in the upper 40 bits (OFFSET_BITS = 24); NULL slots unwritten (0 here)
        bits = hv.data.view(np.uint64)[idx]
        buffers.append(np.where(valid, bits >> np.uint64(24), 0).astype(np.int64).tobytes())
    elif t in ("VARCHAR", "JSON") and sview:
        slen = np.where(valid, hv.str_len[idx], 0).astype(np.int64)
        is_long = slen > 12
        long_len = np.where(is_long, slen, 0)
        offs = np.zeros(n + 1, np.int64)
        np.cumsum(long_len, out=offs[1:])
        views = np.zeros((n, 16), np.uint8)
        views[:, 0:4] = slen.astype("<u4").view(np.uint8).reshape(n, 4)
        head = np.minimum(slen, np.where(is_long, 4, 12))  # bytes stored in the view itself
        tot = int(head.sum())
        if tot:
            r = np.repeat(np.arange(n), head)
            c = np.arange(tot, dtype=np.int64) - np.repeat(np.cumsum(head) - head, head)
            views[r, 4 + c] = hv.heap[np.repeat(hv.str_start[idx], head) + c]
        views[is_long, 12:16] = offs[:-1][is_long].astype("<i4").view(np.uint8).reshape(-1, 4)
        total = int(offs[-1])
        pos = (np.arange(total, dtype=np.int64) - np.repeat(offs[:-1], long_len)
               + np.repeat(hv.str_start[idx], long_len))
        buffers.append(views.tobytes())
        buffers.append(hv.heap[pos].tobytes() if total else b"")
    elif t in ("VARCHAR", "BLOB", "BIT", "VARINT", "JSON"):
        length = np.where(valid, hv.str_len[idx], 0).astype(np.int64)
        offs = np.zeros(n + 1, np.int64)
        np.cumsum(length, out=offs[1:])
        if n and not large and offs[-1] > 2147483647:
            raise OverflowError("Arrow Appender: The maximum total string size for regular string buffers is "
                                "2147483647 but the offset exceeds this")
        total = int(offs[-1])
        pos = np.arange(total, dtype=np.int64) - np.repeat(offs[:-1], length) + np.repeat(hv.str_start[idx], length)
        buffers.append(offs.astype(off_t).tobytes())
        buffers.append(hv.heap[pos].tobytes() if total else b"")
    elif t in ("LIST", "MAP"):
        ent = hv.data.view(np.uint64).reshape(-1, 2)[idx]
        length = np.where(valid, ent[:, 1], 0).astype(np.int64)
        offs = np.zeros(n + 1, np.int64)
        np.cumsum(length, out=offs[1:])
        total = int(offs[-1])
        child_sel = (np.arange(total, dtype=np.int64) - np.repeat(offs[:-1], length)
                     + np.repeat(ent[:, 0].astype(np.int64), length))
        if t == "LIST" and view:  # offsets[i] = running offset before row i (also for NULL rows), sizes[i] = length
            if n and not large and offs[-1] > 2147483647:
                raise OverflowError("Arrow Appender: The maximum combined list offset for regular list buffers is "
                                    "2147483647 but the offset exceeds this")
            buffers.append(offs[:-1].astype(off_t).tobytes())
            buffers.append(length.astype(off_t).tobytes())
        else:
            buffers.append(offs.astype(off_t if t == "LIST" else np.int32).tobytes())
        child = hv.children[0]
        if t == "MAP":
            child._no_validity_buffer = True  # the entries struct's validity buffer is never resized (map_data.hpp)
        append(child, child_sel, nodes, buffers, large, view, sview, below)
    elif t == "STRUCT":
        for c in hv.children:
            append(c, idx, nodes, buffers, large, view, sview, below)
    elif t == "ARRAY":
        k = hv.param
        child_idx = (idx[:, None] * k + np.arange(k)[None, :]).reshape(-1)
        append(hv.children[0], child_idx, nodes, buffers, large, view, sview, below)
    elif t == "DECIMAL" and hv.decimal[0] <= 4:
        buffers.append(hv.data.view(np.int16)[idx].astype(np.int32).tobytes())
    elif t == "INTERVAL":
        src = hv.data.view(np.dtype([("m", np.int32), ("d", np.int32), ("us", np.int64)]))[idx]
        out = np.zeros(n, np.dtype([("m", np.int32), ("d", np.int32), ("ns", np.int64)]))
        out["m"], out["d"] = np.where(valid, src["m"], 0), np.where(valid, src["d"], 0)
        out["ns"] = np.where(valid, src["us"] * 1000, 0)  # SkipNulls: NULL slots stay unwritten (0 here)
        buffers.append(out.tobytes())
    else:
        w = hv.data.size // hv.rows if hv.rows else 1
        buffers.append(hv.data.reshape(-1, w)[idx].tobytes() if n else b"")


def serialize_dictionaries(columns: list):
    """The DictionaryBatch bodies of a stream: one per ENUM field in field pre-order (= dictionary id order), each the
    utf8 array ArrowEnumData::Initialize builds from EnumType::GetValuesInsertOrder (appender/enum_data.hpp:34-75): 0xFF
    validity, int32 offsets whatever arrow_large_buffer_size says, the characters.
    -> [(rows, nodes, buffers[(offset, length)], body bytes)]"""
    out = []

    def walk(hv):
        if hv.type == "ENUM":
            d = hv.children[0]
            nodes, bufs = [], []
            append(d, np.arange(d.rows, dtype=np.int64), nodes, bufs)
            body, slices = b"", []
            for b in bufs:
                slices.append((len(body), len(b)))
                body += _pad8(b)
            out.append((d.rows, nodes, slices, body))
            return
        for c in hv.children:
            walk(c)

    for c in columns:
        walk(c)
    return out


def serialize_batches(columns: list, rows: int, row_group_size: int = 122_880, large: bool = False,
                      view: bool = False, sview: bool = False, lossless: bool = False):
    """-> [(rows, nodes, buffers[(offset, length)], body bytes)] per record batch"""
    out = []
    for r0 in range(0, rows, row_group_size):
        n = min(row_group_size, rows - r0)
        idx = np.arange(r0, r0 + n, dtype=np.int64)
        nodes, bufs = [], []
        for c in columns:
            append(c, idx, nodes, bufs, large, view, sview, lossless)
        body, slices = b"", []
        for b in bufs:
            slices.append((len(body), len(b)))
            body += _pad8(b)
        out.append((n, nodes, slices, body))
    return out
