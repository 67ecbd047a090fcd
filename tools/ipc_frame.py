"""Frames Arrow record batches into an IPC stream WITHOUT dropping buffers: every buffer the producer handed over is
written, validity bitmaps of all-valid columns included.

Why: the reference writes its files with nanoarrow's ArrowIpcEncoderEncodeSimpleRecordBatch over what DuckDB's
ArrowAppender produced (src/writer/column_data_collection_serializer.cpp:80-95); the appender always materialises a
0xFF-filled validity buffer (duckdb/src/common/arrow/arrow_appender.cpp:94, appender/append_data.hpp:117-134) and the
encoder writes every non-NULL buffer.  pyarrow's own IPC writer drops the bitmap of a column without NULLs, so a stream
written by pyarrow does not have the buffer layout of a reference-written file (+2 B/row, 32 more
buffers per lineitem batch).  nanoarrow is not on disk (CMakeLists.txt:7-13), hence this stand-in for its framing:
  encapsulated message = 0xFFFFFFFF, int32 metadata size (8 + size is a multiple of 8), flatbuffer Message, body;
  body = the buffers in pre-order, each starting on an 8-byte boundary, padded with zeros.

Bench / fixture infrastructure (tools/), not product code: the product's writer is duckdb_arrow_b200/src/ipc/ipc_writer.cpp.
"""
from __future__ import annotations

import struct

import pyarrow as pa


def _pad8(n: int) -> int:
    return (n + 7) & ~7


def _walk(arr: pa.Array, nodes: list, bufs: list) -> None:
    """Pre-order (length, null_count) nodes and (address-holding buffer, logical size) pairs."""
    t = arr.type
    n = len(arr)
    assert arr.offset == 0, "sliced arrays are not framed here"
    b = arr.buffers()

    def validity():
        bufs.append((b[0], (n + 7) // 8 if b[0] is not None else 0))

    if pa.types.is_null(t):
        nodes.append((n, n))
        return
    nodes.append((n, arr.null_count))
    if pa.types.is_struct(t):
        validity()
        for i in range(t.num_fields):
            _walk(arr.field(i), nodes, bufs)
    elif pa.types.is_list(t) or pa.types.is_large_list(t) or pa.types.is_map(t):
        w = 8 if pa.types.is_large_list(t) else 4
        validity()
        bufs.append((b[1], (n + 1) * w if b[1] is not None else 0))
        child = arr.values if not pa.types.is_map(t) else pa.StructArray.from_arrays(
            [arr.keys, arr.items], fields=[t.key_field, t.item_field])
        _walk(child, nodes, bufs)
    elif pa.types.is_fixed_size_list(t):
        validity()
        _walk(arr.values, nodes, bufs)
    elif (pa.types.is_string(t) or pa.types.is_binary(t) or pa.types.is_large_string(t)
          or pa.types.is_large_binary(t)):
        w = 8 if (pa.types.is_large_string(t) or pa.types.is_large_binary(t)) else 4
        validity()
        bufs.append((b[1], (n + 1) * w if b[1] is not None else 0))
        if b[1] is not None and n:
            last = struct.unpack_from("<q" if w == 8 else "<i", b[1], n * w)[0]
        else:
            last = 0
        bufs.append((b[2], last if b[2] is not None else 0))
    elif pa.types.is_boolean(t):
        validity()
        bufs.append((b[1], (n + 7) // 8))
    else:
        validity()
        bufs.append((b[1], n * (t.bit_width // 8)))


def record_batch_message(batch: pa.RecordBatch) -> bytes:
    """One encapsulated RecordBatch message (prefix + flatbuffer + body) holding every buffer of `batch`."""
    nodes, bufs = [], []
    for col in batch.columns:
        _walk(col, nodes, bufs)
    # body layout
    offsets, pos = [], 0
    for _, size in bufs:
        offsets.append(pos)
        pos += _pad8(size)
    body_len = pos
    # flatbuffer, laid out front to back (see the module docstring of duckdb_arrow_b200/src/ipc/flatbuf.hpp for the
    # format): [root][Message vtable][Message][RecordBatch vtable][RecordBatch][nodes][buffers]
    fb = bytearray()
    fb += struct.pack("<I", 16)                                   # root -> Message table at 16
    fb += struct.pack("<6H", 12, 20, 16, 18, 4, 8)                # vtable: version, header_type, header, bodyLength
    assert len(fb) == 16
    rb_table = 48
    fb += struct.pack("<iIqhBB", 12, rb_table - 20, body_len, 4, 3, 0)  # soffset, header, bodyLength, V5, RecordBatch
    assert len(fb) == 36
    fb += struct.pack("<5H", 10, 24, 16, 4, 8)                    # vtable: length, nodes, buffers
    fb += b"\0\0"
    assert len(fb) == rb_table
    nodes_vec = 76                                                # count at 76, 16-byte structs from 80
    buffers_vec = _pad8(nodes_vec + 4 + 16 * len(nodes)) + 4
    fb += struct.pack("<iIIIq", rb_table - 36, nodes_vec - (rb_table + 4), buffers_vec - (rb_table + 8), 0,
                      batch.num_rows)
    assert len(fb) == 72
    fb += b"\0" * 4
    fb += struct.pack("<I", len(nodes))
    for length, nulls in nodes:
        fb += struct.pack("<qq", length, nulls)
    fb += b"\0" * (buffers_vec - len(fb))
    fb += struct.pack("<I", len(bufs))
    for off, (_, size) in zip(offsets, bufs):
        fb += struct.pack("<qq", off, size)
    fb += b"\0" * (_pad8(len(fb)) - len(fb))
    out = bytearray(struct.pack("<Ii", 0xFFFFFFFF, len(fb)))
    out += fb
    for (buf, size), off in zip(bufs, offsets):
        if size:
            out += memoryview(buf)[:size]
            out += b"\0" * (_pad8(size) - size)
    assert len(out) == 8 + len(fb) + body_len
    return bytes(out)


def schema_message(schema: pa.Schema) -> bytes:
    return schema.serialize().to_pybytes()


EOS = struct.pack("<Ii", 0xFFFFFFFF, 0)


def write_stream(sink, schema: pa.Schema, batches) -> tuple[int, int]:
    """Writes schema + batches + end-of-stream to a binary file object -> (rows, bytes)."""
    total = sink.write(schema_message(schema))
    rows = 0
    for b in batches:
        total += sink.write(record_batch_message(b))
        rows += b.num_rows
    total += sink.write(EOS)
    return rows, total


def stream_bytes(schema: pa.Schema, batches) -> bytes:
    import io
    bio = io.BytesIO()
    write_stream(bio, schema, batches)
    return bio.getvalue()
